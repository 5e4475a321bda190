"""ctypes access to oracle/_ref/libref_cost.so: the REFERENCE's own residual functors (ReprojectCost of src/adjuster/BundleAdjuster.h:33-69
and src/adjuster/SnavelyReprojectionError.h:11-60) compiled by oracle/ref_build.sh where /root/reference exists.  Test infrastructure."""
import ctypes as C
import os

import numpy as np

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_cost.so")


def available():
    return os.path.exists(LIB)


def _lib():
    lib = C.CDLL(LIB)
    d = C.POINTER(C.c_double)
    for name in ("ref_reproject_cost", "ref_snavely_cost"):
        f = getattr(lib, name)
        f.restype = C.c_int
        f.argtypes = [d, d, d, d, C.c_int, d, d]
    return lib


def evaluate(model, intr, ext, X, obs, taylor_rule=0):
    """model 0: ReprojectCost (intr = fx fy cx cy), 1: SnavelyReprojectionError (intr = f l1 l2).  Returns r[2], Ji[2,ni], Jc[2,6], Jp[2,3]."""
    lib = _lib()
    ni = 4 if model == 0 else 3
    a = np.ascontiguousarray(intr[:ni], np.float64); b = np.ascontiguousarray(ext, np.float64)
    c = np.ascontiguousarray(X, np.float64); o = np.ascontiguousarray(obs, np.float64)
    r = np.empty(2); J = np.empty((2, ni + 9))
    p = lambda v: v.ctypes.data_as(C.POINTER(C.c_double))
    rc = (lib.ref_reproject_cost if model == 0 else lib.ref_snavely_cost)(p(a), p(b), p(c), p(o), int(taylor_rule), p(r), p(J))
    assert rc == 0
    return r, J[:, :ni].copy(), J[:, ni:ni + 6].copy(), J[:, ni + 6:].copy()


def evaluate_problem(model, P, taylor_rule=0):
    """Per-observation residuals and Jacobian blocks of a ba.Problem by the reference functor (no robust kernel, no constant-block masking)."""
    No = P.obs_cam.size
    r = np.empty((No, 2)); Ji = np.zeros((No, 2, 4)); Jc = np.empty((No, 2, 6)); Jp = np.empty((No, 2, 3))
    ni = 4 if model == 0 else 3
    for o in range(No):
        cam, pt = int(P.obs_cam[o]), int(P.obs_pt[o])
        r[o], Ji[o][:, :ni], Jc[o], Jp[o] = evaluate(model, P.intr[P.cam_group[cam]], P.pose[cam], P.pt[pt], P.obs_uv[o], taylor_rule)
    return r, Ji, Jc, Jp
