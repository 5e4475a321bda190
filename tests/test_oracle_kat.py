"""CPU known-answer tests of the oracle.  The reference pins nothing for this path (no tests,
no fixtures: SURVEY F6), so the oracle is anchored on hand-computed values and on independent implementations:
cv2.Rodrigues / cv2.projectPoints, central finite differences and scipy's sparse direct solver (tests/refmath.py).
"""
import math

import numpy as np
import pytest

import refmath as rm

FX, FY, CX, CY = 2905.88, 2905.88, 1416.0, 1064.0   # src/main.cpp:59


def one_obs(ba, pose, X, uv, intr=(FX, FY, CX, CY), fixed=0):
    return ba.Problem(pose=np.array([pose], np.float64), intr=np.array([intr], np.float64), cam_group=[0], cam_fixed=[fixed],
                      pt=np.array([X], np.float64), obs_cam=[0], obs_pt=[0], obs_uv=np.array([uv], np.float64))


def evaluate(oracle, P, **opts):
    with oracle.create(oracle.default_options(**opts)) as h:
        h.set_problem(P)
        return h.evaluate()


# ---- KAT-1: projection ----------------------------------------------------------------------------------
def test_kat1_projection_known_answers(ba, oracle):
    e = evaluate(oracle, one_obs(ba, [0] * 6, [0, 0, 1], [0, 0]), robust_kind=0)
    assert np.allclose(e["r"][0], [CX, CY], rtol=0, atol=0)                      # X on the axis -> principal point
    e = evaluate(oracle, one_obs(ba, [0] * 6, [1, 2, 4], [0, 0]), robust_kind=0)
    assert np.allclose(e["r"][0], [FX / 4 + CX, FY / 2 + CY], rtol=1e-15)
    assert e["cost"] == pytest.approx(0.5 * ((FX / 4 + CX) ** 2 + (FY / 2 + CY) ** 2), rel=1e-15)


def test_kat1_projection_matches_opencv(ba, oracle):
    rng = np.random.default_rng(1)
    for _ in range(50):
        pose = np.concatenate([rng.normal(0, 0.7, 3), rng.normal(0, 1, 3)])
        X = rng.normal(0, 1, 3) + [0, 0, 6]
        e = evaluate(oracle, one_obs(ba, pose, X, [0, 0]), robust_kind=0)
        assert np.allclose(e["r"][0], rm.project((FX, FY, CX, CY), pose, X), rtol=0, atol=2e-9)   # px; cv2 itself is ~1e-12 relative
        assert np.allclose(e["r"][0], rm.project_np((FX, FY, CX, CY), pose, X), rtol=1e-13)


# ---- KAT-2 / KAT-3 / KAT-4: Jacobians ---------------------------------------------------------------------
def test_kat2_jacobian_at_identity(ba, oracle):
    X = np.array([0.3, -0.2, 4.0])
    e = evaluate(oracle, one_obs(ba, [0] * 6, X, [0, 0]), robust_kind=0)
    x, y, z = X
    A = np.array([[FX / z, 0, -FX * x / z ** 2], [0, FY / z, -FY * y / z ** 2]])
    skew = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
    assert np.allclose(e["Jp"][0], A, rtol=1e-15)                # d(u,v)/dX = A R, R = I
    assert np.allclose(e["Jc"][0][:, 3:], A, rtol=1e-15)         # d/dt = A
    assert np.allclose(e["Jc"][0][:, :3], -A @ skew, rtol=1e-15)  # Taylor branch: d(RX)/d(aa) = -[X]x
    assert np.allclose(e["Ji"][0], [[x / z, 0, 1, 0], [0, y / z, 0, 1]], rtol=1e-15)


@pytest.mark.parametrize("rule", [0, 1])
def test_kat3_taylor_rodrigues_continuity(ba, oracle, rule):
    """Either side of theta^2 = DBL_EPSILON the residual and Jacobian agree to O(theta^2)."""
    eps = np.finfo(np.float64).eps
    X = [0.4, -0.3, 5.0]
    d = np.array([1.0, -2.0, 0.5]); d /= np.linalg.norm(d)
    lo = evaluate(oracle, one_obs(ba, np.r_[d * math.sqrt(eps * 0.99), 0.1, 0.2, 0.3], X, [0, 0]), robust_kind=0, taylor_rule=rule)
    hi = evaluate(oracle, one_obs(ba, np.r_[d * math.sqrt(eps * 1.01), 0.1, 0.2, 0.3], X, [0, 0]), robust_kind=0, taylor_rule=rule)
    assert np.allclose(lo["r"], hi["r"], rtol=0, atol=1e-6)
    for k in ("Jc", "Jp"):
        assert np.allclose(lo[k], hi[k], rtol=1e-6, atol=1e-4)


def test_kat4_jacobian_vs_finite_differences(ba, oracle):
    rng = np.random.default_rng(4)
    intr = np.array([FX, FY, CX, CY])
    for _ in range(25):
        pose = np.concatenate([rng.normal(0, 0.5, 3), rng.normal(0, 1, 3)])
        X = rng.normal(0, 1, 3) + [0, 0, 7]
        e = evaluate(oracle, one_obs(ba, pose, X, [10, 20]), robust_kind=0)
        Ji, Jc, Jp = rm.fd_jacobian(intr, pose, X)
        for got, ref in ((e["Ji"][0], Ji), (e["Jc"][0], Jc), (e["Jp"][0], Jp)):
            assert np.abs(got - ref).max() <= 1e-7 * max(np.abs(ref).max(), 1.0)


def test_fixed_pose_has_zero_columns(ba, oracle):
    e = evaluate(oracle, one_obs(ba, [0.1, 0.2, 0.3, 0, 0, 0], [0, 0, 5], [0, 0], fixed=1), robust_kind=0)
    assert not e["Jc"].any() and e["Jp"].any()


# ---- KAT-5: robust kernels ---------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,s,rho,rho1", [(1, 16.0, 16.0, 1.0), (1, 25.0, 24.0, 0.8), (2, 16.0, 16 * math.log(2.0), 0.5), (0, 25.0, 25.0, 1.0)])
def test_kat5_loss_values(ba, oracle, kind, s, rho, rho1):
    """||r||^2 = s is built from a point on the axis and an observation offset (sqrt(s), 0) from the principal point."""
    P = one_obs(ba, [0] * 6, [0, 0, 2], [CX - math.sqrt(s), CY])
    e = evaluate(oracle, P, robust_kind=kind, robust_a=4.0)
    assert e["cost"] == pytest.approx(0.5 * rho, rel=1e-14)
    assert np.allclose(e["r"][0], [math.sqrt(rho1) * math.sqrt(s), 0.0], rtol=1e-14)   # corrector: r <- sqrt(rho') r
    plain = evaluate(oracle, P, robust_kind=0)
    assert np.allclose(e["Jp"], math.sqrt(rho1) * plain["Jp"], rtol=1e-14)             # J <- sqrt(rho') J


def test_cost_matches_numpy_on_ladybug(ba, oracle):
    P = ba.synth("tiny")
    for kind in (0, 1, 2):
        e = evaluate(oracle, P, robust_kind=kind, intrinsics_free=1)
        r, cost = rm.residuals(P, kind, 4.0)
        assert e["cost"] == pytest.approx(cost, rel=1e-12)
        w = np.sqrt([rm.rho(kind, 4.0, s)[1] for s in (r * r).sum(axis=1)])
        assert np.allclose(e["r"], r * w[:, None], rtol=1e-9, atol=1e-9)


# ---- KAT-6: zero-noise scene ---------------------------------------------------------------------------------
def test_kat6_zero_noise_recovers_ground_truth(ba, oracle):
    P = ba.synth("tiny", noise_px=0.0, outlier_fraction=0.0, round_obs_to_float=0)
    with oracle.create(oracle.default_options(intrinsics_free=0, max_num_iterations=50, function_tolerance=0.0, parameter_tolerance=1e-14)) as h:
        h.set_problem(P)
        s = h.solve()
        pose, intr, pt = h.get_state()
    assert s.usable and 2 * s.final_cost < 1e-16 * P.obs_cam.size
    r, cost = rm.residuals(P, 0, 4.0, pose=pose, pt=pt)
    assert np.abs(r).max() < 1e-6                                     # reprojects exactly (the gauge may differ from pose_true)


# ---- KAT-7: one LM step against an independent sparse direct solve ------------------------------------------------
@pytest.mark.parametrize("fi", [0, 1])
@pytest.mark.parametrize("solver", [0, 1])
def test_kat7_first_step_vs_scipy(ba, oracle, fi, solver):
    P = ba.synth("tiny", seed=11)
    opts = dict(intrinsics_free=fi, robust_kind=1, robust_a=4.0, linear_solver=solver)
    e = evaluate(oracle, P, **opts)
    dpose, dintr, dpt = rm.lm_first_step(P, e["Jc"], e["Jp"], e["Ji"], e["r"], radius=1e4, intrinsics_free=bool(fi))
    with oracle.create(oracle.default_options(max_num_iterations=1, **opts)) as h:
        h.set_problem(P)
        s = h.solve()
        pose, intr, pt = h.get_state()
        t = h.trajectory()
    assert t[1].step_is_successful == 1
    scale = max(np.abs(dpt).max(), np.abs(dpose).max())
    assert np.abs((pose - P.pose) - dpose).max() <= 1e-8 * scale
    assert np.abs((pt - P.pt) - dpt).max() <= 1e-8 * scale
    if fi:
        assert np.abs((intr - P.intr) - dintr).max() <= 1e-8 * max(np.abs(dintr).max(), 1e-300)
    else:
        assert np.array_equal(intr, P.intr)
    # candidate cost of that step, recomputed independently
    _, cost = rm.residuals(P, 1, 4.0, pose=P.pose + dpose, intr=P.intr + (dintr if fi else 0), pt=P.pt + dpt)
    assert t[1].candidate_cost == pytest.approx(cost, rel=1e-9)


# ---- KAT-8: index structure ------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["tiny", "ladybug"])
def test_kat8_structure_vs_brute_force(ba, oracle, name):
    P = ba.synth(name)
    with oracle.create(oracle.default_options()) as h:
        h.set_problem(P)
        perm, cam_off, pt_off, pairs = h.structure()
    Nc, _, Np, No = P.sizes
    assert np.array_equal(perm, np.argsort(P.obs_cam, kind="stable").astype(np.int32))
    assert np.array_equal(cam_off, np.r_[0, np.cumsum(np.bincount(P.obs_cam, minlength=Nc))].astype(np.int32))
    assert np.array_equal(pt_off, np.r_[0, np.cumsum(np.bincount(P.obs_pt, minlength=Np))].astype(np.int32))
    assert np.array_equal(pairs, rm.schur_pairs(P))


# ---- KAT-9: trust-region radius schedule -------------------------------------------------------------------------
def check_radius_rule(traj, opts):
    dec = 2.0
    for a in traj[1:]:
        if a.step_is_valid and a.step_is_successful:
            want = min(opts.max_trust_region_radius, a.radius_used / max(1.0 / 3.0, 1.0 - (2.0 * a.relative_decrease - 1.0) ** 3))
            dec = 2.0
        else:
            want = a.radius_used / dec
            dec *= 2.0
        assert a.radius == pytest.approx(want, rel=1e-15), a.iteration
        if a.step_is_valid:
            assert (a.relative_decrease > opts.min_relative_decrease) == bool(a.step_is_successful)


def test_kat9_radius_schedule(ba, oracle):
    # a large first radius and a badly perturbed start give rejected steps, then recoveries
    P = ba.synth("tiny", seed=5, sigma_aa=1.0, sigma_pt_rel=1.0)
    opts = oracle.default_options(intrinsics_free=0, initial_trust_region_radius=1e12, max_num_iterations=30)
    with oracle.create(opts) as h:
        h.set_problem(P)
        h.solve()
        t = h.trajectory()
    assert any(not r.step_is_successful for r in t[1:]) and any(r.step_is_successful for r in t[1:])
    check_radius_rule(t, opts)
    assert t[1].radius_used == 1e12
    costs = [r.cost for r in t]
    assert all(b <= a for a, b in zip(costs, costs[1:]))               # monotonic steps only


def test_kat9_invalid_steps_fail_after_five(ba, oracle):
    """min_lm_diagonal = max = 0 and a rank-deficient problem (a point seen once) -> every linear solve fails -> FAILURE at the 5th."""
    P = one_obs(ba, [0] * 6, [0.1, 0.2, 3.0], [CX + 1, CY - 2])
    with oracle.create(oracle.default_options(intrinsics_free=0, min_lm_diagonal=0.0, max_lm_diagonal=0.0, jacobi_scaling=0)) as h:
        h.set_problem(P)
        s = h.solve()
        t = h.trajectory()
    assert s.termination == 2 and not s.usable
    assert len(t) == 6 and all(not r.step_is_valid for r in t[1:])
    assert [r.radius for r in t[1:5]] == [1e4 / 2, 1e4 / 8, 1e4 / 64, 1e4 / 1024]


def test_termination_usable_flags(ba, oracle):
    P = ba.synth("tiny")
    with oracle.create(oracle.default_options(intrinsics_free=0, max_num_iterations=2)) as h:
        h.set_problem(P)
        s = h.solve()
    assert s.termination == 1 and s.usable and s.num_iterations == 3      # NO_CONVERGENCE is usable (BundleAdjuster.h:128)
    with oracle.create(oracle.default_options(intrinsics_free=0)) as h:
        h.set_problem(P)
        s = h.solve()
    assert s.termination == 0 and s.usable and s.num_residuals == 2 * P.obs_cam.size


# ---- g2o policy: Levenberg rules and the SE3 exponential-map pose update, against SciPy ---------------------
def twist_update(pose, w, u):
    """exp((w, u)^) * T with scipy.linalg.expm on the 4x4 twist (independent of the oracle's closed forms); returns (R', t')."""
    import cv2
    from scipy.linalg import expm
    X = np.zeros((4, 4))
    X[:3, :3] = [[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]]
    X[:3, 3] = u
    E = expm(X)
    R = cv2.Rodrigues(np.asarray(pose[:3], np.float64).reshape(3, 1))[0]
    return E[:3, :3] @ R, E[:3, :3] @ np.asarray(pose[3:]) + E[:3, 3]


def test_g2o_left_twist_jacobian_vs_finite_differences(ba, oracle):
    import cv2
    rng = np.random.default_rng(8)
    intr = (FX, FY, CX, CY)
    for _ in range(10):
        pose = np.concatenate([rng.normal(0, 0.4, 3), rng.normal(0, 1, 3)])
        X = rng.normal(0, 1, 3) + [0, 0, 7]
        e = evaluate(oracle, one_obs(ba, pose, X, [10, 20]), robust_kind=0, intrinsics_free=0, lm_policy=1)

        def proj(d):
            R, t = twist_update(pose, d[:3], d[3:])
            return rm.project_np(intr, np.r_[cv2.Rodrigues(R)[0].ravel(), t], X)
        h = 1e-6
        J = np.stack([(proj(h * np.eye(6)[k]) - proj(-h * np.eye(6)[k])) / (2 * h) for k in range(6)], axis=1)
        assert np.abs(e["Jc"][0] - J).max() <= 1e-6 * np.abs(J).max()


def test_g2o_first_iteration_vs_scipy(ba, oracle):
    """lambda0 = 1e-5 max diag(H), uniform damping, dense solve of the FULL normal equations, exp-map pose update."""
    import cv2
    P = ba.synth("tiny", seed=4)
    opts = dict(intrinsics_free=0, lm_policy=1, robust_kind=1, robust_a=4.0)
    e = evaluate(oracle, P, **opts)
    Nc, _, Np, No = P.sizes
    J = np.zeros((2 * No, 6 * Nc + 3 * Np))
    for o in range(No):
        c, p = int(P.obs_cam[o]), int(P.obs_pt[o])
        J[2 * o:2 * o + 2, 6 * c:6 * c + 6] = e["Jc"][o]
        J[2 * o:2 * o + 2, 6 * Nc + 3 * p:6 * Nc + 3 * p + 3] = e["Jp"][o]
    H = J.T @ J
    lam = 1e-5 * np.diag(H).max()
    act = np.diag(H) > 0
    y = np.zeros(H.shape[0])
    y[act] = np.linalg.solve((H + lam * np.eye(H.shape[0]))[np.ix_(act, act)], (J.T @ e["r"].ravel())[act])
    with oracle.create(oracle.default_options(max_num_iterations=1, **opts)) as h:
        h.set_problem(P)
        h.solve()
        t = h.trajectory()
        pose, intr, pt = h.get_state()
    assert t[1].step_is_successful and t[1].radius_used == pytest.approx(1.0 / lam, rel=1e-12)
    assert np.abs(pt - (P.pt - y[6 * Nc:].reshape(Np, 3))).max() <= 1e-9 * np.abs(y).max()
    for c in range(Nc):
        R, tt = twist_update(P.pose[c], -y[6 * c:6 * c + 3], -y[6 * c + 3:6 * c + 6])
        assert np.abs(cv2.Rodrigues(pose[c, :3].reshape(3, 1))[0] - R).max() <= 1e-10
        assert np.abs(pose[c, 3:] - tt).max() <= 1e-9 * max(1.0, np.abs(tt).max())
    # gain ratio and lambda update of that trial
    _, cost1 = rm.residuals(P, 1, 4.0, pose=pose, pt=pt)
    scale = float(y @ (lam * y + J.T @ e["r"].ravel()))
    rho = (2 * t[0].cost - 2 * cost1) / (scale + 1e-3)
    assert t[1].relative_decrease == pytest.approx(rho, rel=1e-7)
    assert t[1].radius == pytest.approx(1.0 / (lam * max(1 / 3, min(1 - (2 * rho - 1) ** 3, 2 / 3))), rel=1e-7)


def test_g2o_lambda_schedule_and_minimum(ba, oracle):
    P = ba.synth("tiny", seed=5, sigma_aa=1.0, sigma_pt_rel=1.0)
    with oracle.create(oracle.default_options(intrinsics_free=0, lm_policy=1, max_num_iterations=40)) as h:
        h.set_problem(P)
        s = h.solve()
        t = h.trajectory()
    assert any(not r.step_is_successful for r in t[1:])
    nu = 2.0
    for r in t[1:]:
        lam = 1.0 / r.radius_used
        if r.step_is_successful:
            assert r.relative_decrease > 0
            want = lam * max(1 / 3, min(1 - (2 * r.relative_decrease - 1) ** 3, 2 / 3)); nu = 2.0
        else:
            want = lam * nu; nu *= 2.0
        assert 1.0 / r.radius == pytest.approx(want, rel=1e-12)
    with oracle.create(oracle.default_options(intrinsics_free=0, lm_policy=0, max_num_iterations=50)) as h:
        h.set_problem(ba.synth("tiny", seed=5))
        sc = h.solve()
    with oracle.create(oracle.default_options(intrinsics_free=0, lm_policy=1, max_num_iterations=50)) as h:
        h.set_problem(ba.synth("tiny", seed=5))
        sg = h.solve()
    assert sg.final_cost == pytest.approx(sc.final_cost, rel=1e-5)        # both policies reach the same minimum
    with pytest.raises(ba.BAError):
        oracle.create(oracle.default_options(intrinsics_free=1, lm_policy=1))


def test_schur_jacobi_preconditioner_in_the_oracle(ba, oracle):
    """The oracle's two PCG preconditioners (block diagonal of B = Ceres JACOBI, of S = SCHUR_JACOBI) reach the exact-solve trajectory, and
    the block diagonal of the system actually being solved needs fewer iterations (what the engine's opt-in Schur-Jacobi is measured against)."""
    P = ba.synth("ladybug")
    common = dict(intrinsics_free=0, max_num_iterations=5, num_threads=4)
    runs = {}
    for key, extra in (("exact", dict(linear_solver=0)), ("jacobi", dict(linear_solver=1, preconditioner=0)), ("schur", dict(linear_solver=1, preconditioner=1))):
        with oracle.create(oracle.default_options(**common, **extra)) as h:
            h.set_problem(P)
            s = h.solve()
            runs[key] = (s, h.trajectory())
    for key in ("jacobi", "schur"):
        for a, b in zip(runs[key][1], runs["exact"][1]):
            assert a.step_is_successful == b.step_is_successful
            assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
    assert 0 < runs["schur"][0].total_pcg_iterations < runs["jacobi"][0].total_pcg_iterations


def test_forcing_sequence_q_criterion(ba, oracle):
    """KAT for the throughput mode's stop rule (Ceres' inexact-Newton Q-criterion, eta = 0.1): the oracle's PCG iteration count of the
    first LM step equals that of an independent dense numpy PCG on the explicitly formed reduced system, with Q(x) = 1/2 x'Sx - b'x
    evaluated from its definition (the oracle uses the identity Q = -1/2 x'(b + r))."""
    P = ba.synth("tiny", seed=3)
    Nc, Ng, Np, No = P.sizes
    with oracle.create(oracle.default_options(intrinsics_free=0)) as h:
        h.set_problem(P)
        e = h.evaluate()
    Jc, Jp, r = e["Jc"], e["Jp"], e["r"].reshape(-1)
    F = np.zeros((2 * No, 6 * Nc)); E = np.zeros((2 * No, 3 * Np))
    for o in range(No):
        F[2 * o:2 * o + 2, 6 * P.obs_cam[o]:6 * P.obs_cam[o] + 6] = Jc[o]
        E[2 * o:2 * o + 2, 3 * P.obs_pt[o]:3 * P.obs_pt[o] + 3] = Jp[o]
    J = np.hstack([F, E])
    s = 1.0 / (1.0 + np.sqrt((J * J).sum(axis=0)))                  # Jacobi scaling
    Js = J * s
    d2 = np.clip((Js * Js).sum(axis=0), 1e-6, 1e32) / 1e4           # LM diagonal at the initial radius
    Fs, Es = Js[:, :6 * Nc], Js[:, 6 * Nc:]
    B = Fs.T @ Fs + np.diag(d2[:6 * Nc])
    C = Es.T @ Es + np.diag(d2[6 * Nc:])
    W = Fs.T @ Es
    Ci = np.linalg.inv(C)
    S = B - W @ Ci @ W.T
    b = Fs.T @ r - W @ Ci @ (Es.T @ r)
    Minv = np.zeros_like(B)
    for c in range(Nc):
        sl = slice(6 * c, 6 * c + 6)
        Minv[sl, sl] = np.linalg.inv(B[sl, sl])

    def pcg_iterations(eta):
        x = np.zeros_like(b); res = b.copy(); z = Minv @ res; p = z.copy(); rho = res @ z; Q0 = 0.0
        for k in range(1, 500):
            q = S @ p
            alpha = rho / (p @ q)
            x += alpha * p; res -= alpha * q
            z = Minv @ res
            rho_new = res @ z
            Q1 = 0.5 * x @ S @ x - b @ x
            if k * (Q1 - Q0) / Q1 < eta:
                return k
            Q0 = Q1
            p = z + (rho_new / rho) * p; rho = rho_new
        return 500

    for eta in (0.1, 0.01, 1e-4):
        with oracle.create(oracle.default_options(intrinsics_free=0, linear_solver=1, pcg_forcing=1, eta=eta, pcg_rel_tolerance=0.0, max_num_iterations=1)) as h:
            h.set_problem(P)
            h.solve()
            t = h.trajectory()
        assert t[1].pcg_iterations == pcg_iterations(eta), eta
    # and the stop rule buys what it is for: far fewer iterations than the fixed 1e-12 tolerance
    with oracle.create(oracle.default_options(intrinsics_free=0, linear_solver=1, max_num_iterations=1)) as h:
        h.set_problem(P)
        h.solve()
        assert h.trajectory()[1].pcg_iterations > 3 * pcg_iterations(0.1)
