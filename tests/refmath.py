"""Independent numpy/OpenCV/SciPy restatement of the path's arithmetic, used ONLY to check the C++ oracle (tests/).

Nothing here shares code with oracle/ or the engine: rotation comes from cv2.Rodrigues, projection from
cv2.projectPoints, Jacobians from central finite differences, the LM step from scipy.sparse.linalg.spsolve.
Reference call sites restated: ReprojectCost::operator() src/adjuster/BundleAdjuster.h:38-68, HuberLoss(4) :109.
"""
import cv2
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def rotate(aa, X):
    """R(aa) X through OpenCV's Rodrigues (independent of the oracle's AngleAxisRotatePoint restatement)."""
    R, _ = cv2.Rodrigues(np.asarray(aa, np.float64).reshape(3, 1))
    return R @ np.asarray(X, np.float64)


def project(intr, pose, X):
    """Pinhole projection with cv2.projectPoints (no distortion, +z forward): returns (u, v)."""
    fx, fy, cx, cy = intr
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    uv, _ = cv2.projectPoints(np.asarray(X, np.float64).reshape(1, 1, 3), np.asarray(pose[:3], np.float64).reshape(3, 1),
                              np.asarray(pose[3:], np.float64).reshape(3, 1), K, None)
    return uv.reshape(2)


def project_np(intr, pose, X):
    """Same projection in plain numpy with the closed-form Rodrigues formula (used for finite differences: cv2 costs too much per call)."""
    aa = np.asarray(pose[:3], np.float64)
    th = np.linalg.norm(aa)
    X = np.asarray(X, np.float64)
    if th * th > np.finfo(np.float64).eps:
        k = aa / th
        P = X * np.cos(th) + np.cross(k, X) * np.sin(th) + k * np.dot(k, X) * (1 - np.cos(th))
    else:
        P = X + np.cross(aa, X)
    P = P + np.asarray(pose[3:], np.float64)
    return np.array([intr[0] * P[0] / P[2] + intr[2], intr[1] * P[1] / P[2] + intr[3]])


def rho(kind, a, s):
    """(rho, rho') of the robust kernel on s = ||r||^2: 0 none, 1 Huber, 2 Cauchy (Ceres definitions)."""
    b = a * a
    if kind == 1:
        return (s, 1.0) if s <= b else (2 * a * np.sqrt(s) - b, a / np.sqrt(s))
    if kind == 2:
        return b * np.log1p(s / b), 1.0 / (1.0 + s / b)
    return s, 1.0


def residuals(P, kind=0, a=4.0, pose=None, intr=None, pt=None):
    """Unrobustified residuals r[No,2] and cost = 1/2 sum rho(||r||^2) of a Problem, via project_np."""
    pose = P.pose if pose is None else pose
    intr = P.intr if intr is None else intr
    pt = P.pt if pt is None else pt
    r = np.empty((P.obs_cam.size, 2))
    for o, (c, p) in enumerate(zip(P.obs_cam, P.obs_pt)):
        r[o] = project_np(intr[P.cam_group[c]], pose[c], pt[p]) - P.obs_uv[o]
    s = (r * r).sum(axis=1)
    cost = 0.5 * sum(rho(kind, a, si)[0] for si in s)
    return r, cost


def fd_jacobian(intr, pose, X, h=1e-6):
    """Central-difference Jacobians of (u,v) w.r.t. intrinsics (2x4), pose (2x6), point (2x3)."""
    def f(v):
        return project_np(v[:4], v[4:10], v[10:13])
    v0 = np.concatenate([intr, pose, X]).astype(np.float64)
    J = np.empty((2, 13))
    for k in range(13):
        hk = h * max(1.0, abs(v0[k]))
        vp, vm = v0.copy(), v0.copy()
        vp[k] += hk; vm[k] -= hk
        J[:, k] = (f(vp) - f(vm)) / (2 * hk)
    return J[:, :4], J[:, 4:10], J[:, 10:13]


def schur_pairs(P):
    """Brute-force co-visibility: lexicographic list of (i<j) free, observed pose pairs sharing >= 1 point."""
    free = (P.cam_fixed == 0)
    by_pt = {}
    for c, p in zip(P.obs_cam, P.obs_pt):
        by_pt.setdefault(int(p), set()).add(int(c))
    pairs = set()
    for cams in by_pt.values():
        cs = sorted(c for c in cams if free[c])
        for i, a in enumerate(cs):
            for b in cs[i + 1:]:
                pairs.add((a, b))
    return np.array(sorted(pairs), np.int32).reshape(-1, 2)


def lm_first_step(P, Jc, Jp, Ji, r, radius=1e4, intrinsics_free=False, min_diag=1e-6, max_diag=1e32):
    """One Levenberg-Marquardt step exactly as SURVEY 8 (a-LM) states it, but solved on the FULL normal equations
    with a sparse direct solver (no Schur complement, no PCG): returns (dpose, dintr, dpt) to ADD to x.
    Inputs are the robustified, unscaled Jacobian blocks and residuals (what ba_evaluate returns)."""
    Nc, Ng, Np, No = P.sizes
    ni = 4
    nI = ni * Ng if intrinsics_free else 0
    n = 6 * Nc + nI + 3 * Np
    rows, cols, vals = [], [], []
    for o in range(No):
        c, p, g = int(P.obs_cam[o]), int(P.obs_pt[o]), int(P.cam_group[int(P.obs_cam[o])])
        for q in range(2):
            row = 2 * o + q
            for k in range(6):
                rows.append(row); cols.append(6 * c + k); vals.append(Jc[o, q, k])
            if intrinsics_free:
                for k in range(ni):
                    rows.append(row); cols.append(6 * Nc + ni * g + k); vals.append(Ji[o, q, k])
            for k in range(3):
                rows.append(row); cols.append(6 * Nc + nI + 3 * p + k); vals.append(Jp[o, q, k])
    J = sp.csr_matrix((vals, (rows, cols)), shape=(2 * No, n))
    colsq = np.asarray(J.multiply(J).sum(axis=0)).ravel()
    active = colsq > 0                              # constant / unreferenced blocks have zero columns: dropped like Ceres does
    s = 1.0 / (1.0 + np.sqrt(colsq))                # Jacobi scaling, frozen at iteration 0
    Js = J @ sp.diags(s)
    d2 = np.clip(np.asarray(Js.multiply(Js).sum(axis=0)).ravel(), min_diag, max_diag) / radius
    idx = np.flatnonzero(active)
    A = (Js.T @ Js + sp.diags(d2)).tocsc()[idx][:, idx]
    rhs = (Js.T @ r.reshape(-1))[idx]
    y = np.zeros(n)
    y[idx] = spla.spsolve(A, rhs)
    delta = -(s * y)
    return delta[:6 * Nc].reshape(Nc, 6), delta[6 * Nc:6 * Nc + nI].reshape(-1, ni), delta[6 * Nc + nI:].reshape(Np, 3)
