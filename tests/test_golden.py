"""Golden-vector tests (tests/golden/, made by tests/golden/make_golden.py).  CPU: the oracle against the independent
OpenCV / finite-difference / SciPy vectors and against its own frozen trajectories.  GPU: the CUDA engine against the same files."""
import glob
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_problem(ba, d):
    return ba.Problem(**{k: d[k] for k in ("pose", "intr", "cam_group", "cam_fixed", "pt", "obs_cam", "obs_pt", "obs_uv")})


def check_independent(ba, make_handle):
    d = np.load(os.path.join(GOLD, "tiny_independent.npz"))
    P = load_problem(ba, d)
    for kind in (0, 1, 2):
        with make_handle(intrinsics_free=1, robust_kind=kind) as h:
            h.set_problem(P)
            e = h.evaluate()
        assert e["cost"] == pytest.approx(float(d[f"cost_kind{kind}"]), rel=1e-10)
        if kind == 0:
            assert np.abs(e["r"] - d["r_cv2"]).max() <= 1e-8                       # pixels, vs cv2.projectPoints
            free = (P.cam_fixed == 0)[P.obs_cam][:, None, None]
            for got, ref in ((e["Jc"], d["Jc_fd"] * free), (e["Jp"], d["Jp_fd"]), (e["Ji"], d["Ji_fd"])):
                assert np.abs(got - ref).max() <= 1e-6 * np.abs(ref).max()            # central differences: ~1e-7 relative
    for fi in (0, 1):
        with make_handle(intrinsics_free=fi, max_num_iterations=1) as h:
            h.set_problem(P)
            h.solve()
            pose, intr, pt = h.get_state()
        sp, st = d[f"step_pose_fi{fi}"], d[f"step_pt_fi{fi}"]
        scale = max(np.abs(sp).max(), np.abs(st).max())
        assert np.abs((pose - P.pose) - sp).max() <= 2e-5 * scale                   # step from the FD Jacobian: FD-limited
        assert np.abs((pt - P.pt) - st).max() <= 2e-5 * scale
        if fi:
            assert np.abs((intr - P.intr) - d["step_intr_fi1"]).max() <= 2e-5 * np.abs(d["step_intr_fi1"]).max()


def check_trajectories(ba, make_handle, rtol):
    files = sorted(glob.glob(os.path.join(GOLD, "*_oracle_trajectory.npz")))
    assert len(files) >= 4
    for f in files:
        name, fi = os.path.basename(f).split("_")[0], int(os.path.basename(f).split("_fi")[1][0])
        d = np.load(f)
        P = ba.synth(name)
        assert int(P.obs_cam.astype(np.int64).sum() * 31 + P.obs_pt.astype(np.int64).sum()) == int(d["obs_checksum"])   # generator is bit-stable
        with make_handle(intrinsics_free=fi, max_num_iterations=15) as h:
            h.set_problem(P)
            s = h.solve()
            t = h.trajectory()
            pose, intr, pt = h.get_state()
        assert len(t) == len(d["cost"]) and s.termination == int(d["termination"]), f
        assert np.array_equal([r.step_is_successful for r in t], d["accepted"]), f
        assert np.allclose([r.cost for r in t], d["cost"], rtol=rtol, atol=0), f
        assert np.allclose([r.radius for r in t], d["radius"], rtol=1e-5), f
        for got, ref in ((pose, d["final_pose"]), (intr, d["final_intr"]), (pt[:64], d["final_pt_head"])):
            assert np.linalg.norm(got - ref) <= 1e-5 * np.linalg.norm(ref), f


def test_oracle_vs_independent_golden(ba, oracle):
    check_independent(ba, lambda **kw: oracle.create(oracle.default_options(**kw)))


def test_oracle_vs_frozen_trajectories(ba, oracle):
    check_trajectories(ba, lambda **kw: oracle.create(oracle.default_options(num_threads=8, **kw)), rtol=1e-9)


@pytest.mark.gpu
def test_engine_vs_independent_golden(ba):
    check_independent(ba, lambda **kw: ba.Engine(**kw))


@pytest.mark.gpu
def test_engine_vs_frozen_trajectories(ba):
    """chi2 trajectory within 1e-6 relative, same accept/reject flags, final state within 1e-5 (north-star bars)."""
    check_trajectories(ba, lambda **kw: ba.Engine(**kw), rtol=1e-6)
