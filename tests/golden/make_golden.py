"""Regenerates the golden vectors under tests/golden/ (run in the CPU container: python tests/golden/make_golden.py).

The reference holds no golden vectors for this path and cannot be built or imported here, so these vectors
come from two sources, both recorded per file:
  * "independent": computed WITHOUT the oracle or the engine — OpenCV (cv2.projectPoints) for residuals, central finite
    differences for Jacobians, SciPy's sparse direct solver on the full normal equations for the first LM step (tests/refmath.py);
  * "oracle": the CPU oracle's LM trajectory with the reference's configuration (exact reduced solve), first 15 LM iterations (an exact
    and a PCG reduced solve agree to ~1e-9 per step, but accept/reject decisions of a gauge-free problem become sensitive to
    that after dozens of iterations), frozen so that a later change of oracle or engine that moves the trajectory is caught.  These pin regressions, not the reference.
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(HERE))
import refmath as rm  # noqa: E402

ba = importlib.import_module("3dreconstruction_b200")
orc = ba.Library(os.path.join(ROOT, "oracle", "libba_oracle.so"))


def problem_arrays(P):
    return dict(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt, obs_cam=P.obs_cam, obs_pt=P.obs_pt, obs_uv=P.obs_uv)


def independent(name, **kw):
    P = ba.synth(name, **kw)
    r = np.array([rm.project(P.intr[P.cam_group[c]], P.pose[c], P.pt[p]) - P.obs_uv[o] for o, (c, p) in enumerate(zip(P.obs_cam, P.obs_pt))])
    J = [rm.fd_jacobian(P.intr[P.cam_group[c]], P.pose[c], P.pt[p]) for c, p in zip(P.obs_cam, P.obs_pt)]
    Ji, Jc, Jp = (np.array([j[k] for j in J]) for k in range(3))
    out = dict(problem_arrays(P), r_cv2=r, Ji_fd=Ji, Jc_fd=Jc, Jp_fd=Jp)
    for kind in (0, 1, 2):
        out[f"cost_kind{kind}"] = 0.5 * sum(rm.rho(kind, 4.0, s)[0] for s in (r * r).sum(axis=1))
    # first LM step (Huber 4, radius 1e4, Jacobi scaling) from the finite-difference-free closed forms is not independent of the
    # analytic Jacobian, so the step is computed from the FD Jacobian here: agreement is limited to ~1e-6 by FD accuracy
    for fi in (0, 1):
        w = np.sqrt([rm.rho(1, 4.0, s)[1] for s in (r * r).sum(axis=1)])
        free = (P.cam_fixed == 0)[P.obs_cam].astype(float)
        dpose, dintr, dpt = rm.lm_first_step(P, Jc * (w * free)[:, None, None], Jp * w[:, None, None], Ji * w[:, None, None], r * w[:, None],
                                             radius=1e4, intrinsics_free=bool(fi))
        out[f"step_pose_fi{fi}"], out[f"step_pt_fi{fi}"] = dpose, dpt
        if fi:
            out["step_intr_fi1"] = dintr
    return out


def oracle_trajectory(name, **opts):
    P = ba.synth(name)
    with orc.create(orc.default_options(**opts)) as h:
        h.set_problem(P)
        s = h.solve()
        t = h.trajectory()
        pose, intr, pt = h.get_state()
    return dict(cost=np.array([r.cost for r in t]), accepted=np.array([r.step_is_successful for r in t], np.int8),
                radius=np.array([r.radius for r in t]), termination=s.termination, final_pose=pose, final_intr=intr,
                final_pt_head=pt[:64], obs_checksum=np.int64(P.obs_cam.astype(np.int64).sum() * 31 + P.obs_pt.astype(np.int64).sum()))


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "tiny_independent.npz"), **independent("tiny", seed=11))
    for name, fi in (("tiny", 1), ("ladybug", 0), ("ladybug", 1), ("sequential30", 1)):
        np.savez_compressed(os.path.join(HERE, f"{name}_fi{fi}_oracle_trajectory.npz"), **oracle_trajectory(name, intrinsics_free=fi, num_threads=8, max_num_iterations=15))
    print("golden vectors written to", HERE)
