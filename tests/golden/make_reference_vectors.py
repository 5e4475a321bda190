"""Writes tests/golden/reference_functor_vectors.npz: residuals and Jacobian blocks computed by the REFERENCE's own cost functors
(oracle/_ref/libref_cost.so, built by oracle/ref_build.sh from /root/reference/src/adjuster/{BundleAdjuster.h:33-69, SnavelyReprojectionError.h})
on fixed problems.  Run in the CPU container, where /root/reference exists:  python tests/golden/make_reference_vectors.py
The vectors travel; the reference does not (tests on the GPU box read only the .npz)."""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(HERE))
import reffunctor as rf  # noqa: E402
from test_snavely_bal import snavely_problem  # noqa: E402

ba = importlib.import_module("3dreconstruction_b200")
KEYS = ("pose", "intr", "cam_group", "cam_fixed", "pt", "obs_cam", "obs_pt", "obs_uv")


def main():
    assert rf.available(), "run oracle/ref_build.sh first (needs /root/reference)"
    out = {}
    # pinhole (the reference's ReprojectCost): the 'tiny' scene, plus the reference flow's zero-initialised first pose (Taylor branch)
    P = ba.synth("tiny")
    pose = P.pose.copy(); pose[1, :3] = 0.0                      # exact zeros: theta^2 <= DBL_EPSILON branch of AngleAxisRotatePoint
    pose[2, :3] = [3e-9, -5e-9, 8e-9]                            # theta^2 ~ 1e-16 < DBL_EPSILON (rule 0: Taylor; rule 1: Rodrigues)
    P0 = ba.Problem(**{**{k: getattr(P, k) for k in KEYS}, "pose": pose})
    for k in KEYS:
        out[f"pin_{k}"] = getattr(P0, k)
    for rule in (0, 1):
        r, Ji, Jc, Jp = rf.evaluate_problem(0, P0, rule)
        out[f"pin_r_rule{rule}"], out[f"pin_Ji_rule{rule}"], out[f"pin_Jc_rule{rule}"], out[f"pin_Jp_rule{rule}"] = r, Ji, Jc, Jp
    # Snavely / BAL (the reference's SnavelyReprojectionError)
    S = snavely_problem(ba)
    for k in KEYS:
        out[f"sna_{k}"] = getattr(S, k)
    r, Ji, Jc, Jp = rf.evaluate_problem(1, S, 0)
    out["sna_r"], out["sna_Ji"], out["sna_Jc"], out["sna_Jp"] = r, Ji, Jc, Jp
    path = os.path.join(HERE, "reference_functor_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", P0.obs_cam.size, "pinhole +", S.obs_cam.size, "Snavely observations")


if __name__ == "__main__":
    main()
