"""The identity behind the pose-factored Schur product (csrc/ba_rows.cuh: k_pose_factors, k_schur_rows_f):

    Jc = Jp . K(X) . G_c,   K(X) = [ -[X]x | I ],   G_c = [ J_r(aa) 0 ; 0 R^T ]        (Ceres policy: angle-axis + translation)
                                                    G_c = [ R^T 0 ; -[R^T t]x R^T  R^T ] (g2o policy: left twist)

checked on the CPU against the EXPLICIT pose columns of the oracle (which restates the reference's cost functors and is itself pinned to
them, tests/test_reference_functors.py).  G_c is formed here in numpy from the same closed forms the kernel uses; the GPU side of the
same statement is tests/test_engine_gpu.py::test_pose_factored_matvec_matches_explicit."""
import numpy as np
import pytest


def skew(v):
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def rotation(aa):
    th2 = float(aa @ aa)
    if th2 <= 2.220446049250313e-16:          # Ceres' first-order branch (AngleAxisRotatePoint)
        return np.eye(3) + skew(aa), 0.0, 0.0, True
    th = np.sqrt(th2)
    a, b = np.sin(th) / th, (1.0 - np.cos(th)) / th2
    return np.cos(th) * np.eye(3) + a * skew(aa) + b * np.outer(aa, aa), a, b, False


def pose_factor(pose, left_twist):
    aa, t = pose[:3], pose[3:]
    R, a, b, first_order = rotation(aa)
    G = np.zeros((6, 6))
    G[3:, 3:] = R.T
    if left_twist:
        G[:3, :3] = R.T
        G[3:, :3] = -skew(R.T @ t) @ R.T
    else:
        th2 = float(aa @ aa)
        if first_order:
            Jr = np.eye(3)
        else:
            cc = (1.0 - a) / th2 if th2 >= 0.25 else sum((-1.0) ** k * th2 ** k / float(np.prod(np.arange(1, 2 * k + 4))) for k in range(8))
            Jr = np.eye(3) - b * skew(aa) + cc * (np.outer(aa, aa) - th2 * np.eye(3))
        G[:3, :3] = Jr
    return G


@pytest.mark.parametrize("policy", [0, 1])
@pytest.mark.parametrize("shape", ["tiny", "ladybug"])
def test_pose_columns_factor_through_the_point_columns(ba, oracle, shape, policy):
    P = ba.synth(shape)
    P.cam_fixed[:] = 0
    with oracle.create(oracle.default_options(intrinsics_free=0, robust_kind=1, robust_a=4.0, lm_policy=policy)) as h:
        h.set_problem(P)
        e = h.evaluate()
    Jc, Jp = e["Jc"], e["Jp"]
    G = np.stack([pose_factor(p, policy == 1) for p in P.pose])
    worst = 0.0
    for o in range(len(P.obs_cam)):
        X = P.pt[P.obs_pt[o]]
        K = np.hstack([-skew(X), np.eye(3)])
        want = Jc[o]
        got = Jp[o] @ K @ G[P.obs_cam[o]]
        worst = max(worst, np.abs(got - want).max() / max(np.abs(want).max(), 1e-300))
    assert worst <= 1e-11, worst


def test_snavely_model_factors_the_same_way(ba, oracle):
    """The factorisation does not depend on the camera model: both reach the pose only through P = R X + t."""
    cv2 = pytest.importorskip("cv2")   # the Snavely problem builder uses cv2.Rodrigues
    del cv2
    from test_snavely_bal import snavely_problem
    P = snavely_problem(ba, "tiny", seed=3)
    P.cam_fixed[:] = 0
    with oracle.create(oracle.default_options(model=1, intrinsics_free=1, robust_kind=0)) as h:
        h.set_problem(P)
        e = h.evaluate()
    for o in range(len(P.obs_cam)):
        X = P.pt[P.obs_pt[o]]
        got = e["Jp"][o] @ np.hstack([-skew(X), np.eye(3)]) @ pose_factor(P.pose[P.obs_cam[o]], False)
        assert np.abs(got - e["Jc"][o]).max() <= 1e-11 * np.abs(e["Jc"][o]).max()


def test_first_order_branch_is_exact_only_for_a_zero_rotation(ba, oracle):
    """A camera on Ceres' first-order branch (|aa|^2 <= eps): the stored derivative is -[X]x, the factored one -(I + [aa]x)[X]x.  Equal for
    aa = 0 exactly; for 0 < |aa| <= 1.5e-8 they differ by O(|aa|) relative -- the documented limit of the factored form."""
    P = ba.synth("tiny")
    P.cam_fixed[:] = 0
    P.pose[0, :3] = 0.0
    P.pose[1, :3] = [1e-9, -2e-9, 5e-10]
    with oracle.create(oracle.default_options(intrinsics_free=0, robust_kind=0)) as h:
        h.set_problem(P)
        e = h.evaluate()
    err = {0: 0.0, 1: 0.0}
    for o in np.flatnonzero(P.obs_cam <= 1):
        c = int(P.obs_cam[o])
        X = P.pt[P.obs_pt[o]]
        got = e["Jp"][o] @ np.hstack([-skew(X), np.eye(3)]) @ pose_factor(P.pose[c], False)
        err[c] = max(err[c], np.abs(got - e["Jc"][o]).max() / np.abs(e["Jc"][o]).max())
    assert err[0] <= 1e-13
    assert err[1] <= 1e-7
