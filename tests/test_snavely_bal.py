"""The BAL / Snavely model (src/adjuster/SnavelyReprojectionError.h:16-48): residual and Jacobian against an
independent numpy restatement + finite differences, the BAL text reader/writer round trip, PLY export; GPU: engine against oracle."""
import os

import cv2
import numpy as np
import pytest


def snavely_project(cam9, X):
    """camera = (angle-axis, t, f, l1, l2); -z forward (numpy + cv2.Rodrigues, independent of oracle/engine)."""
    R = cv2.Rodrigues(np.asarray(cam9[:3], np.float64).reshape(3, 1))[0]
    P = R @ np.asarray(X) + cam9[3:6]
    xp, yp = -P[0] / P[2], -P[1] / P[2]
    r2 = xp * xp + yp * yp
    d = 1.0 + r2 * (cam9[7] + cam9[8] * r2)
    return np.array([cam9[6] * d * xp, cam9[6] * d * yp])


def snavely_problem(ba, name="tiny", seed=1, noise=0.5):
    """A BAL-style problem from a synthetic pinhole scene: flip the camera frame to -z forward (P' = diag(1,-1,-1) P), give every
    camera its own (f, l1, l2), synthesise observations with the Snavely model."""
    S = ba.synth(name, seed=seed)
    rng = np.random.default_rng(seed)
    Nc, _, Np, No = S.sizes
    D = np.diag([1.0, -1.0, -1.0])
    pose_true = np.empty((Nc, 6))
    for c in range(Nc):
        R = cv2.Rodrigues(S.pose_true[c, :3].reshape(3, 1))[0]
        pose_true[c, :3] = cv2.Rodrigues(D @ R)[0].ravel()
        pose_true[c, 3:] = D @ S.pose_true[c, 3:]
    intr_true = np.stack([2900 + rng.normal(0, 30, Nc), -0.05 + rng.normal(0, 0.005, Nc), 0.01 + rng.normal(0, 0.002, Nc), np.zeros(Nc)], axis=1)
    uv = np.array([snavely_project(np.r_[pose_true[c], intr_true[c, :3]], S.pt_true[p]) for c, p in zip(S.obs_cam, S.obs_pt)])
    uv += rng.normal(0, noise, uv.shape)
    pose0 = pose_true + np.c_[rng.normal(0, 0.005, (Nc, 3)), rng.normal(0, 0.005 * S.extent, (Nc, 3))]
    intr0 = intr_true * np.array([1.01, 1.0, 1.0, 1.0]) + np.array([0, 0.002, -0.001, 0])
    pt0 = S.pt_true + rng.normal(0, 0.01 * S.extent, (Np, 3))
    fixed = np.zeros(Nc, np.uint8); fixed[0] = 1
    return ba.Problem(pose=pose0, intr=intr0, cam_group=np.arange(Nc, dtype=np.int32), cam_fixed=fixed, pt=pt0, obs_cam=S.obs_cam,
                      obs_pt=S.obs_pt, obs_uv=uv, pose_true=pose_true, pt_true=S.pt_true, extent=S.extent)


def check_residual_and_jacobian(make_handle, ba):
    P = snavely_problem(ba)
    with make_handle(model=1, intrinsics_free=1, robust_kind=0) as h:
        h.set_problem(P)
        e = h.evaluate()
    for o in range(0, P.obs_cam.size, 7):
        c, p = int(P.obs_cam[o]), int(P.obs_pt[o])
        v0 = np.r_[P.pose[c], P.intr[c, :3], P.pt[p]]

        def f(v):
            return snavely_project(v[:9], v[9:]) - P.obs_uv[o]
        assert np.allclose(e["r"][o], f(v0), rtol=0, atol=1e-9)
        J = np.empty((2, 12))
        for k in range(12):
            h_ = 1e-6 * max(1.0, abs(v0[k])); vp, vm = v0.copy(), v0.copy(); vp[k] += h_; vm[k] -= h_
            J[:, k] = (f(vp) - f(vm)) / (2 * h_)
        free = 0.0 if P.cam_fixed[c] else 1.0
        for got, ref in ((e["Jc"][o], J[:, :6] * free), (e["Ji"][o][:, :3], J[:, 6:9]), (e["Jp"][o], J[:, 9:])):
            assert np.abs(got - ref).max() <= 2e-6 * max(np.abs(ref).max(), 1.0)
        assert not e["Ji"][o][:, 3].any()


def test_oracle_snavely_residual_and_jacobian(ba, oracle):
    check_residual_and_jacobian(lambda **kw: oracle.create(oracle.default_options(**kw)), ba)


def test_oracle_snavely_solve_reduces_cost(ba, oracle):
    P = snavely_problem(ba)
    with oracle.create(oracle.default_options(model=1, intrinsics_free=1, num_threads=4, max_num_iterations=25)) as h:
        h.set_problem(P)
        s = h.solve()
    assert s.usable and s.final_cost < 1e-2 * s.initial_cost
    assert 2 * s.final_cost / (2 * P.obs_cam.size) < 1.0          # mean squared residual below 1 px^2 (noise 0.5 px)


def test_bal_round_trip_and_ply(ba, tmp_path):
    P = snavely_problem(ba)
    path = str(tmp_path / "problem-6-40-pre.txt")
    # BAL files may list observations in any order: write them shuffled, the reader sorts stably by point
    rng = np.random.default_rng(0)
    perm = rng.permutation(P.obs_cam.size)
    Q = ba.Problem(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt, obs_cam=P.obs_cam[perm], obs_pt=P.obs_pt[perm], obs_uv=P.obs_uv[perm])
    ba.write_bal(path, Q)
    head = open(path).readline().split()
    assert [int(x) for x in head] == [P.sizes[0], P.sizes[2], P.sizes[3]]
    B = ba.read_bal(path)
    assert B.sizes == (P.sizes[0], P.sizes[0], P.sizes[2], P.sizes[3])
    assert np.array_equal(B.pose, P.pose) and np.array_equal(B.pt, P.pt) and np.array_equal(B.intr[:, :3], P.intr[:, :3]) and not B.intr[:, 3].any()
    assert np.all(np.diff(B.obs_pt) >= 0) and not B.cam_fixed.any() and np.array_equal(B.cam_group, np.arange(P.sizes[0]))
    key = lambda X: sorted(zip(X.obs_pt.tolist(), X.obs_cam.tolist(), map(tuple, X.obs_uv.tolist())))
    assert key(B) == key(P)
    with pytest.raises(ba.BAError):
        ba.read_bal(str(tmp_path / "missing.txt"))
    ply = str(tmp_path / "map.ply")
    ba.write_ply(ply, P.pt, P.pose)
    lines = open(ply).read().split("\n")
    assert lines[0] == "ply" and f"element vertex {P.sizes[2] + P.sizes[0]}" in lines
    body = lines[lines.index("end_header") + 1:]
    assert np.allclose([float(x) for x in body[0].split()[:3]], P.pt[0])
    c = P.sizes[0] - 1
    R = cv2.Rodrigues(P.pose[c, :3].reshape(3, 1))[0]
    assert np.allclose([float(x) for x in body[P.sizes[2] + c].split()[:3]], -R.T @ P.pose[c, 3:], atol=1e-12)
    assert body[P.sizes[2] + c].split()[3:] == ["0", "255", "0"]


@pytest.mark.gpu
def test_engine_snavely_residual_and_jacobian(ba):
    check_residual_and_jacobian(lambda **kw: ba.Engine(**kw), ba)


@pytest.mark.gpu
@pytest.mark.parametrize("fi,solver,scatter", [(0, 2, 0), (1, 2, 0), (0, 1, 6), (1, 1, 6), (0, 1, 4)])
def test_engine_snavely_trajectory_parity(ba, oracle, fi, solver, scatter):
    """solver 2 = AUTO (direct solve at this size); solver 1 = PCG with the streaming (6) / CTA-per-tile (4) matvec: nine stored Jacobian columns."""
    if scatter == 6:
        from test_engine_gpu import need_stream
        need_stream(ba, 6)
    P = snavely_problem(ba, "ladybug", seed=2)
    with ba.Engine(model=1, intrinsics_free=fi, max_num_iterations=10, linear_solver=solver, scatter_mode=scatter) as eng, \
            oracle.create(oracle.default_options(model=1, intrinsics_free=fi, num_threads=4, max_num_iterations=10)) as h:
        eng.set_problem(P); h.set_problem(P)
        se, so = eng.solve(), h.solve()
        te, to = eng.trajectory(), h.trajectory()
        xe, xo = eng.get_state(), h.get_state()
    assert len(te) == len(to) and se.termination == so.termination
    for a, b in zip(te, to):
        assert a.step_is_successful == b.step_is_successful and abs(a.cost - b.cost) <= 1e-6 * abs(b.cost)
    for k in range(3):
        assert np.linalg.norm(xe[k] - xo[k]) <= 1e-5 * np.linalg.norm(xo[k])
