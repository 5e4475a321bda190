"""GPU parity tests: the CUDA engine against the CPU oracle on the same seeded inputs, through the C-ABI.

Bars (BASELINE.json north_star / SURVEY §8d): per-iteration chi2 within 1e-6 relative with identical accept/reject
flags; final poses and points within 1e-5 relative; index structure bit-exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CHI2_RTOL = 1e-6
STATE_RTOL = 1e-5


def solve_both(ba, oracle, P, **opts):
    oopts = dict(opts)
    threads = oopts.pop("oracle_threads", 4)
    oopts.pop("scatter_mode", None)
    oopts.pop("engine_solver", None)
    opts.pop("oracle_threads", None)
    solver = opts.pop("engine_solver", None)      # engine only: 0 dense Schur + Cholesky, 1 PCG, None = AUTO; the oracle stays EXACT
    if solver is not None:
        opts["linear_solver"] = solver
    with ba.Engine(**opts) as eng:
        eng.set_problem(P)
        se = eng.solve()
        te = eng.trajectory()
        xe = eng.get_state()
    with oracle.create(oracle.default_options(num_threads=threads, **oopts)) as h:
        h.set_problem(P)
        so = h.solve()
        to = h.trajectory()
        xo = h.get_state()
    return (se, te, xe), (so, to, xo)


def assert_trajectory_parity(te, to, upto=None):
    n = min(len(te), len(to)) if upto is None else upto
    assert n > 1
    for a, b in zip(te[:n], to[:n]):
        assert a.iteration == b.iteration
        assert a.step_is_valid == b.step_is_valid, f"iteration {a.iteration}: valid flag differs"
        assert a.step_is_successful == b.step_is_successful, f"iteration {a.iteration}: accept/reject differs"
        assert abs(a.cost - b.cost) <= CHI2_RTOL * abs(b.cost), f"iteration {a.iteration}: chi2 {2*a.cost!r} vs oracle {2*b.cost!r}"
        assert abs(a.radius - b.radius) <= 1e-6 * abs(b.radius)


def need_stream(ba, mode):
    """scatter_mode 6 exists only in libraries built with -DBA_WITH_STREAM (tools/build_variants.sh stream:"-DBA_WITH_STREAM")."""
    if mode != 6:
        return
    try:
        ba.Engine(scatter_mode=6).close()
    except ba.BAError as e:
        if "BA_WITH_STREAM" in str(e):
            pytest.skip("library built without the streaming kernel")
        raise


def rel_err(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def test_backend_is_cuda(engine_lib):
    assert engine_lib.backend() == "cuda-sm100a"


@pytest.mark.parametrize("name", ["tiny", "ladybug"])
def test_evaluate_parity(ba, oracle, name):
    """K0+K1+K2: residuals, analytic Jacobian blocks, cost and gradient against the Jet-differentiated oracle."""
    P = ba.synth(name)
    for fi in (0, 1):
        with ba.Engine(intrinsics_free=fi) as eng, oracle.create(oracle.default_options(intrinsics_free=fi)) as h:
            eng.set_problem(P)
            h.set_problem(P)
            a, b = eng.evaluate(), h.evaluate()
        assert abs(a["cost"] - b["cost"]) <= 1e-12 * abs(b["cost"])
        for k in ("r", "Jc", "Jp", "Ji"):
            scale = max(np.abs(b[k]).max(), 1e-300)
            assert np.abs(a[k] - b[k]).max() <= 1e-10 * scale, k
        assert np.abs(a["g"] - b["g"]).max() <= 1e-10 * np.abs(b["g"]).max()


@pytest.mark.parametrize("solver", [0, 1])
@pytest.mark.parametrize("name,fi", [("tiny", 0), ("tiny", 1), ("ladybug", 0), ("ladybug", 1), ("trafalgar", 0), ("trafalgar", 1)])
def test_trajectory_parity(ba, oracle, name, fi, solver):
    """Both engine solvers of the reduced system (0: dense Schur complement + Cholesky, 1: PCG) against the oracle's exact solve."""
    P = ba.synth(name)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, intrinsics_free=fi, max_num_iterations=12, engine_solver=solver)
    assert se.usable and so.usable
    assert len(te) == len(to)
    assert_trajectory_parity(te, to)
    assert se.termination == so.termination
    assert rel_err(xe[0], xo[0]) <= STATE_RTOL
    assert rel_err(xe[2], xo[2]) <= STATE_RTOL
    assert rel_err(xe[1], xo[1]) <= STATE_RTOL


@pytest.mark.parametrize("mode", [1, 4, 6])
@pytest.mark.parametrize("shape,fi", [("ladybug", 0), ("ladybug", 1), ("trafalgar", 0)])
def test_scatter_modes(ba, oracle, mode, shape, fi):
    """The camera-side scatter implementations (1: L2 atomics, 4: partial rows with one CTA per tile, 6: the same rows from the persistent
    cp.async.bulk streaming kernel — an experiment that is compiled only with -DBA_WITH_STREAM) agree with the oracle."""
    need_stream(ba, mode)
    P = ba.synth(shape, seed=123)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, intrinsics_free=fi, max_num_iterations=6, scatter_mode=mode, engine_solver=1)
    assert_trajectory_parity(te, to)


@pytest.mark.parametrize("shape,fi,model", [("ladybug", 0, 0), ("ladybug", 1, 0), ("trafalgar", 0, 0), ("trafalgar", 1, 0)])
def test_streaming_kernel_is_bit_identical_to_cta_per_tile(ba, shape, fi, model):
    """k_schur_stream (scatter_mode 6) runs the same tile math on the same static schedule as k_schur_rows (4): same rows, same sums,
    so the whole solve is bit-identical (pinhole and Snavely models, fixed and free intrinsics; Trafalgar: several tiles per CTA, ring reuse)."""
    need_stream(ba, 6)
    P = ba.synth(shape, seed=5)
    out = []
    for mode in (4, 6):
        with ba.Engine(intrinsics_free=fi, model=model, max_num_iterations=5, linear_solver=1, scatter_mode=mode) as eng:
            eng.set_problem(P)
            s = eng.solve()
            out.append((s, [(r.cost, r.pcg_iterations, r.model_cost_change) for r in eng.trajectory()], eng.get_state()))
    assert out[0][1] == out[1][1]
    for a, b in zip(out[0][2], out[1][2]):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("shape", ["ladybug", "trafalgar"])
def test_forcing_sequence_and_target_chi2(ba, oracle, shape):
    """Throughput mode (hard part 1): Ceres' inexact-Newton stop rule (eta = 0.1, Q-criterion) in engine and oracle — same PCG
    iteration counts (to within the rounding of the stop test), same chi2 per iteration to 1e-6, far fewer PCG iterations than the fixed
    tolerance; and the target_chi2 stop rule ends the solve at the first iteration whose chi2 is at or below the target."""
    P = ba.synth(shape)
    common = dict(intrinsics_free=0, max_num_iterations=12, linear_solver=1, pcg_forcing=1, eta=0.1, pcg_max_iterations=500)
    with ba.Engine(**common) as eng:
        eng.set_problem(P)
        se = eng.solve()
        te = eng.trajectory()
    with oracle.create(oracle.default_options(num_threads=8, **common)) as h:
        h.set_problem(P)
        so = h.solve()
        to = h.trajectory()
    assert len(te) == len(to)
    for a, b in zip(te, to):
        assert a.step_is_successful == b.step_is_successful
        assert abs(a.pcg_iterations - b.pcg_iterations) <= 2, (a.iteration, a.pcg_iterations, b.pcg_iterations)
        assert abs(a.cost - b.cost) <= 1e-6 * abs(b.cost), (a.iteration, a.cost, b.cost)
    with ba.Engine(**dict(common, pcg_forcing=0)) as eng:
        eng.set_problem(P)
        st = eng.solve()
    assert se.total_pcg_iterations < st.total_pcg_iterations / 3
    assert abs(se.final_cost - st.final_cost) <= 1e-4 * st.final_cost            # judged on final chi2 only
    target = 2.0 * te[4].cost * (1 + 1e-9)
    with ba.Engine(**dict(common, target_chi2=target)) as eng:
        eng.set_problem(P)
        s2 = eng.solve()
        t2 = eng.trajectory()
    assert s2.termination == 3 and len(t2) == 5 and t2[-1].cost == te[4].cost    # BA_TERM_TARGET_REACHED at iteration 4


@pytest.mark.parametrize("solver,fi", [(0, 0), (0, 1), (1, 0)])
def test_solve_is_bit_reproducible(ba, solver, fi):
    """Every sum has a fixed order (partial rows / by-block pair lists): two solves give bit-identical states."""
    P = ba.synth("ladybug", seed=9)
    states = []
    for _ in range(2):
        with ba.Engine(intrinsics_free=fi, max_num_iterations=5, linear_solver=solver) as eng:
            eng.set_problem(P)
            eng.solve()
            states.append(eng.get_state())
    for a, b in zip(*states):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_robust_kernels(ba, oracle, kind):
    P = ba.synth("ladybug", seed=77)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, intrinsics_free=0, robust_kind=kind, robust_a=4.0, max_num_iterations=8)
    assert_trajectory_parity(te, to)


def test_structure_bit_exact(ba, oracle):
    P = ba.synth("ladybug")
    with ba.Engine(intrinsics_free=0) as eng, oracle.create(oracle.default_options(intrinsics_free=0)) as h:
        eng.set_problem(P)
        h.set_problem(P)
        for x, y in zip(eng.structure(), h.structure()):
            assert x.dtype == y.dtype and x.shape == y.shape and np.array_equal(x, y)


def test_reset_state_reproduces(ba):
    P = ba.synth("ladybug")
    with ba.Engine(intrinsics_free=0, max_num_iterations=5) as eng:
        eng.set_problem(P)
        s1 = eng.solve()
        eng.reset_state()
        s2 = eng.solve()
    assert abs(s1.final_cost - s2.final_cost) <= 1e-9 * s1.final_cost
    assert s1.num_iterations == s2.num_iterations


@pytest.mark.parametrize("solver,scatter", [(0, 0), (1, 4), (1, 6)])
def test_long_track_and_ragged(ba, oracle, solver, scatter):
    """A point seen by more cameras than one tile holds (long-point path of every K4/K5 kernel), empty points and an unobserved camera."""
    need_stream(ba, scatter)
    P = ba.synth(num_cameras=300, num_points=60, num_observations=60 * 2 + 298 + 100, seed=5, max_track=300)
    # make point 0 visible from every camera: rebuild the observation list by hand
    rng = np.random.default_rng(3)
    Nc, Np = 300, 60
    cams, pts = [], []
    for p in range(Np):
        k = Nc if p == 0 else (40 if p == 20 else int(rng.integers(2, 6)))
        cs = np.sort(rng.choice(Nc - 1, size=min(k, Nc - 1), replace=False))   # camera Nc-1 is never observed
        if p in (7, 8):
            continue                                                            # points 7 and 8 have no observation
        cams.append(cs); pts.append(np.full(cs.size, p))
    obs_cam = np.concatenate(cams).astype(np.int32); obs_pt = np.concatenate(pts).astype(np.int32)
    # exact projections of the true scene + noise
    from math import sqrt
    def rot(aa):
        th = np.linalg.norm(aa)
        if th < 1e-300: return np.eye(3)
        k = aa / th; K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    uv = np.empty((obs_cam.size, 2))
    for i, (c, p) in enumerate(zip(obs_cam, obs_pt)):
        Pc = rot(P.pose_true[c, :3]) @ P.pt_true[p] + P.pose_true[c, 3:]
        uv[i] = [P.intr[0, 0] * Pc[0] / Pc[2] + P.intr[0, 2], P.intr[0, 1] * Pc[1] / Pc[2] + P.intr[0, 3]]
    uv += rng.normal(0, 0.5, uv.shape)
    Q = ba.Problem(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt, obs_cam=obs_cam, obs_pt=obs_pt, obs_uv=uv)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, Q, intrinsics_free=1, max_num_iterations=6, engine_solver=solver, scatter_mode=scatter)
    assert_trajectory_parity(te, to)
    # untouched parameters stay bit-identical
    assert np.array_equal(xe[2][7], Q.pt[7]) and np.array_equal(xe[2][8], Q.pt[8])
    assert np.array_equal(xe[0][Nc - 1], Q.pose[Nc - 1])


def test_evaluation_failure_is_reported(ba):
    """A point on a camera's principal plane (Pz = 0) makes the evaluation non-finite: FAILURE, not usable."""
    P = ba.synth("tiny")
    c, p = int(P.obs_cam[0]), int(P.obs_pt[0])
    P.pose[c] = 0.0
    P.pt[p] = [1.0, 2.0, 0.0]
    with ba.Engine(intrinsics_free=0) as eng:
        eng.set_problem(P)
        s = eng.solve()
    assert s.termination == 2 and not s.usable


def test_multi_gpu_matches_single_gpu():
    """Sharded NCCL solve == single-GPU solve (needs >= 2 GPUs on the box; skipped otherwise). tools/multi_gpu_check.py under torchrun."""
    import os
    import subprocess
    import sys
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29517", os.path.join(root, "tools", "multi_gpu_check.py"), "ladybug"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MISMATCH" not in r.stdout and "OK" in r.stdout


@pytest.mark.parametrize("policy", [0, 1])
@pytest.mark.parametrize("shape,fi,model", [("ladybug", 0, 0), ("ladybug", 1, 0), ("trafalgar", 0, 0), ("ladybug", 1, 1)])
def test_pose_factored_matvec_matches_explicit(ba, monkeypatch, shape, fi, model, policy):
    """K5a in its pose-factored form (Jc = Jp K(X) G_c: streams the point columns only) against the same engine with the
    explicit pose columns (BA_POSE_FACTORED=0, read at ba_create): the same operator up to rounding, so with a tight PCG tolerance the two
    Levenberg-Marquardt runs agree far inside the parity bar -- for both pose parametrisations (Ceres: angle-axis + t; g2o: left twist),
    with constant cameras, free intrinsics and both camera models."""
    if policy == 1 and (fi or model):
        pytest.skip("g2o policy: pinhole model with constant intrinsics only")
    if model == 1:
        from test_snavely_bal import snavely_problem
        P = snavely_problem(ba, shape, seed=2)
    else:
        P = ba.synth(shape)
    P.cam_fixed[:] = 0
    P.cam_fixed[[0, 3]] = 1
    opts = dict(intrinsics_free=fi, model=model, lm_policy=policy, linear_solver=1, pcg_rel_tolerance=1e-12, pcg_max_iterations=3000, max_num_iterations=6)
    out = []
    for factored in (1, 0):
        monkeypatch.setenv("BA_POSE_FACTORED", str(factored))
        with ba.Engine(**opts) as eng:
            eng.set_problem(P)
            s = eng.solve()
            out.append((s, eng.trajectory(), eng.get_state(), eng.execution_info()))
    (sf, tf, xf, inf), (se, te, xe, ine) = out
    assert inf["pose_factored"] == 1 and ine["pose_factored"] == 0 and inf["scatter_mode"] == ine["scatter_mode"] == 4
    assert len(tf) == len(te) and sf.termination == se.termination
    for a, b in zip(tf, te):
        assert a.step_is_successful == b.step_is_successful
        assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost), (a.iteration, a.cost, b.cost)
        assert abs(a.pcg_iterations - b.pcg_iterations) <= max(8, 0.08 * b.pcg_iterations), (a.iteration, a.pcg_iterations, b.pcg_iterations)
    assert rel_err(xf[0], xe[0]) <= 1e-8 and rel_err(xf[2], xe[2]) <= 1e-8 and rel_err(xf[1], xe[1]) <= 1e-8
    assert np.array_equal(xf[0][[0, 3]], P.pose[[0, 3]])     # constant poses: G_c = 0, untouched


@pytest.mark.parametrize("name", ["tiny", "ladybug", "trafalgar"])
def test_g2o_policy_trajectory_parity(ba, oracle, name):
    """g2o policy (Levenberg lambda schedule, uniform damping, SE3 exp-map pose update): engine against the oracle in the same policy."""
    P = ba.synth(name)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, intrinsics_free=0, lm_policy=1, max_num_iterations=10)
    assert se.usable and so.usable and len(te) == len(to) and se.termination == so.termination
    assert_trajectory_parity(te, to)
    for a, b in zip(te, to):
        assert a.relative_decrease == pytest.approx(b.relative_decrease, rel=1e-3, abs=1e-7)   # (chi2 - chi2+) cancels near convergence
    assert rel_err(xe[0], xo[0]) <= STATE_RTOL and rel_err(xe[2], xo[2]) <= STATE_RTOL
    assert np.array_equal(xe[1], P.intr)


def test_linear_solver_resolution(ba):
    """AUTO takes the direct solve where it applies (SURVEY 8 f-2 regime), PCG elsewhere; EXACT refuses what it cannot do."""
    P = ba.synth("ladybug")
    for fi in (0, 1):
        with ba.Engine(intrinsics_free=fi) as eng:
            eng.set_problem(P)
            assert eng.linear_solver_in_use() == 0
    with ba.Engine(intrinsics_free=0, linear_solver=1) as eng:
        eng.set_problem(P)
        assert eng.linear_solver_in_use() == 1
    big = ba.synth(num_cameras=400, num_points=3000, num_observations=12000, seed=1)      # 2400 reduced unknowns > 2048
    with ba.Engine(intrinsics_free=0) as eng:
        eng.set_problem(big)
        assert eng.linear_solver_in_use() == 1
    with ba.Engine(intrinsics_free=0, linear_solver=0) as eng:                              # forced: allowed up to 8192
        eng.set_problem(big)
        assert eng.linear_solver_in_use() == 0
    Nc = P.sizes[0]
    per_cam = ba.Problem(pose=P.pose, intr=np.repeat(P.intr, Nc, axis=0), cam_group=np.arange(Nc, dtype=np.int32), cam_fixed=P.cam_fixed, pt=P.pt,
                         obs_cam=P.obs_cam, obs_pt=P.obs_pt, obs_uv=P.obs_uv)
    with ba.Engine(intrinsics_free=1) as eng:
        eng.set_problem(per_cam)
        assert eng.linear_solver_in_use() == 1                                              # free intrinsics, one group per camera
    with ba.Engine(intrinsics_free=1, linear_solver=0) as eng:
        with pytest.raises(ba.BAError, match="BA_LINEAR_EXACT"):
            eng.set_problem(per_cam)


@pytest.mark.parametrize("fi", [0, 1])
def test_direct_solve_matches_oracle_tightly(ba, oracle, fi):
    """The dense path is the oracle's algorithm (exact Schur complement): agreement is at rounding level, not PCG-tolerance level."""
    P = ba.synth("ladybug", seed=4)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, intrinsics_free=fi, max_num_iterations=8, engine_solver=0)
    assert len(te) == len(to)
    for a, b in zip(te, to):
        assert a.step_is_successful == b.step_is_successful and abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
        assert a.pcg_iterations == 0
    assert rel_err(xe[2], xo[2]) <= 1e-8


def test_direct_solve_survives_a_singular_system(ba):
    """No damping on a problem whose scale gauge is free: S is singular to rounding.  Whatever the factorisation meets (a non-positive
    pivot: invalid step, radius shrinks; or a tiny pivot: a wild step that is rejected) the solve ends cleanly with finite costs."""
    P = ba.synth("tiny")
    with ba.Engine(intrinsics_free=0, linear_solver=0, min_lm_diagonal=0.0, max_lm_diagonal=0.0, max_num_iterations=3) as eng:
        eng.set_problem(P)
        s = eng.solve()
        t = eng.trajectory()
    assert s.num_iterations >= 1 and all(np.isfinite(r.cost) for r in t)


@pytest.mark.parametrize("fi", [0, 1])
def test_schur_jacobi_preconditioner(ba, oracle, fi):
    """preconditioner = SCHUR_JACOBI (block diagonal of S): same trajectory as the exact-solve oracle, the same PCG iteration counts as the
    oracle's own Schur-Jacobi PCG (to within rounding of the stop test), and fewer iterations than block-Jacobi on B."""
    P = ba.synth("ladybug")
    common = dict(intrinsics_free=fi, max_num_iterations=6)
    (se, te, xe), (so, to, xo) = solve_both(ba, oracle, P, engine_solver=1, preconditioner=1, **common)
    assert_trajectory_parity(te, to)
    with oracle.create(oracle.default_options(num_threads=4, linear_solver=1, preconditioner=1, **common)) as h:
        h.set_problem(P)
        h.solve()
        tp = h.trajectory()
    with ba.Engine(linear_solver=1, preconditioner=0, **common) as eng:
        eng.set_problem(P)
        sj = eng.solve()
    for a, b in zip(te[1:], tp[1:]):
        # the 1e-12 stop test sits on a plateau of the residual: a few iterations either way
        assert abs(a.pcg_iterations - b.pcg_iterations) <= max(4, 0.05 * b.pcg_iterations), (a.iteration, a.pcg_iterations, b.pcg_iterations)
    assert se.total_pcg_iterations <= sj.total_pcg_iterations, (se.total_pcg_iterations, sj.total_pcg_iterations)
