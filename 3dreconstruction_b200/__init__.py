"""3dreconstruction_b200 — B200-native bundle-adjustment engine behind 3DReconstruction's BundleAdjuster interface.

Python here is a thin ctypes host over the C-ABI (include/ba_engine.h) for tests and bench.py; the product is
csrc/libba_engine.so (hand-written sm_100a CUDA) and host/adjuster/BundleAdjuster.h (the C++ drop-in adapter).
The directory name starts with a digit, so import it with importlib.import_module("3dreconstruction_b200").

There is NO CPU fallback: Engine() loads csrc/libba_engine.so and raises if it is missing or if no CUDA device
is usable.  Library(path) can bind any implementation of the header — tests use it to bind the CPU oracle
(oracle/libba_oracle.so) as the checker.
"""
import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import _capi
from ._capi import (BAIterRecord, BAKernelStat, BAOptions, BAProblem, BASummary, SynthProblem, SynthSpec)  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# BA_ENGINE_LIB selects another build of the SAME CUDA engine (kernel-tuning experiments: csrc/libba_engine_<variant>.so)
ENGINE_LIB = os.environ.get("BA_ENGINE_LIB") or os.path.join(_HERE, "csrc", "libba_engine.so")
SYNTH_LIB = os.path.join(_HERE, "host", "libba_synth.so")


class BAError(RuntimeError):
    pass


def _as(arr, dtype, shape=None):
    a = np.ascontiguousarray(arr, dtype=dtype)
    if shape is not None:
        a = a.reshape(shape)
    return a


@dataclass
class Problem:
    """Canonical-indexing BA problem: observations sorted by point id."""
    pose: np.ndarray       # [Nc,6] angle-axis | t
    intr: np.ndarray       # [Ng,4]
    cam_group: np.ndarray  # [Nc] int32
    cam_fixed: np.ndarray  # [Nc] uint8
    pt: np.ndarray         # [Np,3]
    obs_cam: np.ndarray    # [No] int32
    obs_pt: np.ndarray     # [No] int32, non-decreasing
    obs_uv: np.ndarray     # [No,2]
    pose_true: np.ndarray = None
    pt_true: np.ndarray = None
    extent: float = 0.0

    def __post_init__(self):
        self.pose = _as(self.pose, np.float64).reshape(-1, 6)
        self.intr = _as(self.intr, np.float64).reshape(-1, 4)
        self.cam_group = _as(self.cam_group, np.int32).reshape(-1)
        self.cam_fixed = _as(self.cam_fixed, np.uint8).reshape(-1)
        self.pt = _as(self.pt, np.float64).reshape(-1, 3)
        self.obs_cam = _as(self.obs_cam, np.int32).reshape(-1)
        self.obs_pt = _as(self.obs_pt, np.int32).reshape(-1)
        self.obs_uv = _as(self.obs_uv, np.float64).reshape(-1, 2)

    @property
    def sizes(self):
        return self.pose.shape[0], self.intr.shape[0], self.pt.shape[0], self.obs_cam.shape[0]

    def c_struct(self):
        Nc, Ng, Np, No = self.sizes
        p = BAProblem()
        p.num_cameras, p.num_groups, p.num_points, p.num_observations = Nc, Ng, Np, No
        p.pose = self.pose.ctypes.data_as(_capi.c_double_p)
        p.intr = self.intr.ctypes.data_as(_capi.c_double_p)
        p.cam_group = self.cam_group.ctypes.data_as(_capi.c_int32_p)
        p.cam_fixed = self.cam_fixed.ctypes.data_as(_capi.c_uint8_p)
        p.pt = self.pt.ctypes.data_as(_capi.c_double_p)
        p.obs_cam = self.obs_cam.ctypes.data_as(_capi.c_int32_p)
        p.obs_pt = self.obs_pt.ctypes.data_as(_capi.c_int32_p)
        p.obs_uv = self.obs_uv.ctypes.data_as(_capi.c_double_p)
        return p


class Library:
    """One loaded implementation of include/ba_engine.h."""

    def __init__(self, path):
        if not os.path.exists(path):
            raise BAError(f"{path} not found — build it first (python -c 'import __graft_entry__ as g; g.build()')")
        self.path = path
        self.lib = _capi.bind(C.CDLL(path, mode=C.RTLD_LOCAL), _capi.ENGINE_SYMBOLS)

    def backend(self):
        return self.lib.ba_backend().decode()

    def io_bytes(self):
        a, b = C.c_int64(0), C.c_int64(0)
        self.lib.ba_get_io_bytes(C.byref(a), C.byref(b))
        return a.value, b.value

    def nccl_unique_id(self):
        buf = C.create_string_buffer(128)
        rc = self.lib.ba_nccl_unique_id(buf)
        if rc != 0:
            raise BAError(f"ba_nccl_unique_id failed ({rc}): {self.lib.ba_last_error(None).decode()}")
        return buf.raw

    def shard_points(self, obs_pt, num_points, world_size):
        """Point ranges [shard_pt[r], shard_pt[r+1]) of the multi-GPU sharding (host-only; needs no device)."""
        obs_pt = _as(obs_pt, np.int32)
        out = np.zeros(world_size + 1, np.int32)
        rc = self.lib.ba_shard_points(int(num_points), int(obs_pt.size), obs_pt.ctypes.data_as(_capi.c_int32_p), int(world_size),
                                      out.ctypes.data_as(_capi.c_int32_p))
        if rc != 0:
            raise BAError(f"ba_shard_points failed ({rc})")
        return out

    def default_options(self, **kw):
        o = BAOptions()
        rc = self.lib.ba_options_default(C.byref(o))
        if rc != 0:
            raise BAError(f"ba_options_default failed: {rc}")
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(f"ba_options has no field {k}")
            setattr(o, k, v)
        return o

    def create(self, options=None, **kw):
        return Handle(self, options if options is not None else self.default_options(**kw))


class Handle:
    """An opaque ba_engine* plus numpy-friendly wrappers of every C-ABI call."""

    def __init__(self, library, options):
        self.library, self.lib = library, library.lib
        self.options = options
        self._keep = []
        h = C.c_void_p()
        rc = self.lib.ba_create(C.byref(options), C.byref(h))
        if rc != 0:
            raise BAError(f"ba_create failed ({rc}): {self.lib.ba_last_error(None).decode()}")
        self.h = h
        self.sizes = None

    def _check(self, rc, what):
        if rc != 0:
            raise BAError(f"{what} failed ({rc}): {self.lib.ba_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.ba_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_problem(self, problem: Problem):
        cs = problem.c_struct()
        self._check(self.lib.ba_set_problem(self.h, C.byref(cs)), "ba_set_problem")
        self.sizes = problem.sizes

    def append(self, pose=None, cam_group=None, cam_fixed=None, intr=None, pt=None, obs_cam=None, obs_pt=None, obs_uv=None):
        """ba_append: grow the problem held by the handle (new poses / groups / points after the existing ones, observations in any
        order); the state already in the handle is kept."""
        pose = _as(np.zeros((0, 6)) if pose is None else pose, np.float64).reshape(-1, 6)
        nc = pose.shape[0]
        cam_group = _as(np.zeros(nc) if cam_group is None else cam_group, np.int32).reshape(-1)
        cam_fixed = _as(np.zeros(nc) if cam_fixed is None else cam_fixed, np.uint8).reshape(-1)
        intr = _as(np.zeros((0, 4)) if intr is None else intr, np.float64).reshape(-1, 4)
        pt = _as(np.zeros((0, 3)) if pt is None else pt, np.float64).reshape(-1, 3)
        obs_cam = _as(np.zeros(0) if obs_cam is None else obs_cam, np.int32).reshape(-1)
        obs_pt = _as(np.zeros(0) if obs_pt is None else obs_pt, np.int32).reshape(-1)
        obs_uv = _as(np.zeros((0, 2)) if obs_uv is None else obs_uv, np.float64).reshape(-1, 2)
        if not (cam_group.size == cam_fixed.size == nc and obs_cam.size == obs_pt.size == obs_uv.shape[0]):
            raise ValueError("inconsistent delta sizes")
        d, i, u = _capi.c_double_p, _capi.c_int32_p, _capi.c_uint8_p
        D = _capi.BADelta(nc, pose.ctypes.data_as(d), cam_group.ctypes.data_as(i), cam_fixed.ctypes.data_as(u), intr.shape[0], intr.ctypes.data_as(d),
                          pt.shape[0], pt.ctypes.data_as(d), obs_cam.size, obs_cam.ctypes.data_as(i), obs_pt.ctypes.data_as(i), obs_uv.ctypes.data_as(d))
        self._check(self.lib.ba_append(self.h, C.byref(D)), "ba_append")
        self.sizes = (self.sizes[0] + nc, self.sizes[1] + intr.shape[0], self.sizes[2] + pt.shape[0], self.sizes[3] + obs_cam.size)

    def update_parameters(self, kind, first, values):
        """ba_update_parameters: kind 0 poses (n x 6), 1 intrinsics (n x 4), 2 points (n x 3), ids [first, first + n)."""
        w = (6, 4, 3)[kind]
        v = _as(values, np.float64).reshape(-1, w)
        self._check(self.lib.ba_update_parameters(self.h, kind, first, v.shape[0], v.ctypes.data_as(_capi.c_double_p)), "ba_update_parameters")

    def observations(self):
        """(obs_cam, obs_pt, obs_uv) in the handle's order."""
        No = self.sizes[3]
        oc, op, uv = np.empty(No, np.int32), np.empty(No, np.int32), np.empty((No, 2))
        n = C.c_int32(0)
        self._check(self.lib.ba_get_observations(self.h, oc.ctypes.data_as(_capi.c_int32_p), op.ctypes.data_as(_capi.c_int32_p),
                                                 uv.ctypes.data_as(_capi.c_double_p), No, C.byref(n)), "ba_get_observations")
        assert n.value == No
        return oc, op, uv

    def reset_state(self):
        self._check(self.lib.ba_reset_state(self.h), "ba_reset_state")

    def solve(self):
        s = BASummary()
        self._check(self.lib.ba_solve(self.h, C.byref(s)), "ba_solve")
        return s

    def get_state(self):
        Nc, Ng, Np, _ = self.sizes
        pose, intr, pt = np.empty((Nc, 6)), np.empty((Ng, 4)), np.empty((Np, 3))
        self._check(self.lib.ba_get_state(self.h, pose.ctypes.data_as(_capi.c_double_p), intr.ctypes.data_as(_capi.c_double_p),
                                          pt.ctypes.data_as(_capi.c_double_p)), "ba_get_state")
        return pose, intr, pt

    def trajectory(self):
        n = C.c_int32(0)
        self._check(self.lib.ba_get_trajectory(self.h, None, 0, C.byref(n)), "ba_get_trajectory")
        buf = (BAIterRecord * max(1, n.value))()
        self._check(self.lib.ba_get_trajectory(self.h, buf, n.value, C.byref(n)), "ba_get_trajectory")
        return [buf[i] for i in range(n.value)]

    def structure(self, pairs=True):
        Nc, _, Np, No = self.sizes
        perm, cam_off, pt_off = np.empty(No, np.int32), np.empty(Nc + 1, np.int32), np.empty(Np + 1, np.int32)
        n = C.c_int64(0)
        i32 = _capi.c_int32_p
        self._check(self.lib.ba_get_structure(self.h, perm.ctypes.data_as(i32), cam_off.ctypes.data_as(i32), pt_off.ctypes.data_as(i32),
                                              None, 0, C.byref(n) if pairs else None), "ba_get_structure")
        sp = np.empty((max(0, n.value), 2), np.int32)
        if pairs and n.value:
            self._check(self.lib.ba_get_structure(self.h, None, None, None, sp.ctypes.data_as(i32), n.value, C.byref(n)), "ba_get_structure")
        return perm, cam_off, pt_off, sp

    def evaluate(self, allow_failure=False):
        Nc, Ng, Np, No = self.sizes
        cost = C.c_double(0)
        r, Jc, Jp, Ji = np.empty((No, 2)), np.empty((No, 2, 6)), np.empty((No, 2, 3)), np.empty((No, 2, 4))
        g = np.empty(6 * Nc + 4 * Ng + 3 * Np)
        d = _capi.c_double_p
        rc = self.lib.ba_evaluate(self.h, C.byref(cost), r.ctypes.data_as(d), Jc.ctypes.data_as(d), Jp.ctypes.data_as(d),
                                  Ji.ctypes.data_as(d), g.ctypes.data_as(d))
        if not allow_failure:
            self._check(rc, "ba_evaluate")
        return dict(rc=rc, cost=cost.value, r=r, Jc=Jc, Jp=Jp, Ji=Ji, g=g)

    def linear_solver_in_use(self):
        """0 (EXACT: dense Schur complement + Cholesky) or 1 (PCG): what linear_solver = AUTO resolved to for the current problem."""
        v = C.c_int32(0)
        self._check(self.lib.ba_linear_solver_in_use(self.h, C.byref(v)), "ba_linear_solver_in_use")
        return v.value

    def execution_info(self):
        """dict(linear_solver, scatter_mode, allreduce (0 one GPU, 1 NCCL, 2 NVLink peer memory), host_syncs of the last solve, tiles, rows,
        pose_factored (1: the PCG matvec streams the point columns only, ba_rows.cuh k_schur_rows_f))."""
        buf = (C.c_int64 * 8)()
        n = C.c_int32(0)
        self._check(self.lib.ba_get_execution_info(self.h, buf, 8, C.byref(n)), "ba_get_execution_info")
        return dict(zip(("linear_solver", "scatter_mode", "allreduce", "host_syncs", "tiles", "rows", "pose_factored", "tail_cluster"), [int(v) for v in buf]))

    def set_profiling(self, mask, stride=1):
        """mask: False/0 off, True/1 every kernel, else a bit mask over kernel_stats() order; stride: time every n-th launch of
        the PCG-loop kernels only (events between launches serialise the chain's programmatic dependent launches)."""
        self._check(self.lib.ba_set_profiling_stride(self.h, int(stride)), "ba_set_profiling_stride")
        self._check(self.lib.ba_set_profiling(self.h, int(mask)), "ba_set_profiling")

    def kernel_stats(self):
        n = C.c_int32(0)
        buf = (BAKernelStat * 64)()
        self._check(self.lib.ba_get_kernel_stats(self.h, buf, 64, C.byref(n)), "ba_get_kernel_stats")
        return [dict(name=buf[i].name.decode(), launches=buf[i].launches, total_ms=buf[i].total_ms, bytes_per_launch=buf[i].bytes_per_launch)
                for i in range(min(64, n.value))]


_engine_library = None


def engine_library():
    """The CUDA engine library. Raises (never falls back) when it has not been built."""
    global _engine_library
    if _engine_library is None:
        _engine_library = Library(ENGINE_LIB)
        if _engine_library.backend() != "cuda-sm100a":
            raise BAError(f"{ENGINE_LIB} is not the CUDA engine (backend {_engine_library.backend()!r})")
    return _engine_library


def Engine(**options):
    """Create a CUDA engine handle. Keyword arguments override fields of ba_options."""
    lib = engine_library()
    return lib.create(lib.default_options(**options))


# ---------------------------------------------------------------------------------------------------------
# synthetic problems (include/ba_synth.h)
_synth = None


def _synth_lib():
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_LIB):
            raise BAError(f"{SYNTH_LIB} not found — build it first")
        _synth = _capi.bind(C.CDLL(SYNTH_LIB), _capi.SYNTH_SYMBOLS)
    return _synth


def synth_spec(name=None, **kw):
    lib = _synth_lib()
    s = SynthSpec()
    if name is None:
        lib.ba_synth_spec_default(C.byref(s))
    elif lib.ba_synth_named(name.encode(), C.byref(s)) != 0:
        raise BAError(f"unknown synthetic shape {name!r}")
    for k, v in kw.items():
        if not hasattr(s, k):
            raise AttributeError(f"ba_synth_spec has no field {k}")
        setattr(s, k, v)
    return s


def _problem_from_c(p):
    Nc, Ng, Np, No = p.num_cameras, p.num_groups, p.num_points, p.num_observations

    def arr(ptr, n, dt):
        return np.ctypeslib.as_array(ptr, shape=(max(n, 1),))[:n].astype(dt, copy=True)

    return Problem(pose=arr(p.pose, Nc * 6, np.float64), intr=arr(p.intr, Ng * 4, np.float64), cam_group=arr(p.cam_group, Nc, np.int32),
                   cam_fixed=arr(p.cam_fixed, Nc, np.uint8), pt=arr(p.pt, Np * 3, np.float64), obs_cam=arr(p.obs_cam, No, np.int32),
                   obs_pt=arr(p.obs_pt, No, np.int32), obs_uv=arr(p.obs_uv, No * 2, np.float64),
                   pose_true=arr(p.pose_true, Nc * 6, np.float64).reshape(Nc, 6), pt_true=arr(p.pt_true, Np * 3, np.float64).reshape(Np, 3),
                   extent=p.extent)


def read_bal(path):
    """Read a 'Bundle Adjustment in the Large' text file (Snavely model: solve with model=MODEL_SNAVELY)."""
    lib = _synth_lib()
    pp = C.POINTER(SynthProblem)()
    rc = lib.ba_bal_read(os.fsencode(path), C.byref(pp))
    if rc != 0:
        raise BAError(f"ba_bal_read({path!r}) failed ({rc})")
    try:
        return _problem_from_c(pp.contents)
    finally:
        lib.ba_synth_free(pp)


def write_bal(path, P):
    Nc, _, Np, No = P.sizes
    d, i = _capi.c_double_p, _capi.c_int32_p
    rc = _synth_lib().ba_bal_write(os.fsencode(path), Nc, Np, No, P.pose.ctypes.data_as(d), P.intr.ctypes.data_as(d), P.cam_group.ctypes.data_as(i),
                                   P.pt.ctypes.data_as(d), P.obs_cam.ctypes.data_as(i), P.obs_pt.ctypes.data_as(i), P.obs_uv.ctypes.data_as(d))
    if rc != 0:
        raise BAError(f"ba_bal_write failed ({rc})")


def write_ply(path, pt, pose):
    """ASCII PLY of points and camera centres (the reference's writeToPLYFile / exportToPly)."""
    pt, pose = _as(pt, np.float64).reshape(-1, 3), _as(pose, np.float64).reshape(-1, 6)
    rc = _synth_lib().ba_write_ply(os.fsencode(path), pt.shape[0], pt.ctypes.data_as(_capi.c_double_p), pose.shape[0], pose.ctypes.data_as(_capi.c_double_p))
    if rc != 0:
        raise BAError(f"ba_write_ply failed ({rc})")


def synth(name=None, **kw):
    """Generate a deterministic synthetic problem: synth('ladybug'), synth('tiny', seed=7), synth(num_cameras=..)."""
    lib = _synth_lib()
    s = synth_spec(name, **kw)
    pp = C.POINTER(SynthProblem)()
    rc = lib.ba_synth_generate(C.byref(s), C.byref(pp))
    if rc != 0:
        raise BAError(f"ba_synth_generate failed ({rc})")
    try:
        p = pp.contents
        Nc, Ng, Np, No = p.num_cameras, p.num_groups, p.num_points, p.num_observations

        def arr(ptr, n, dt):
            return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True)

        return Problem(pose=arr(p.pose, Nc * 6, np.float64), intr=arr(p.intr, Ng * 4, np.float64), cam_group=arr(p.cam_group, Nc, np.int32),
                       cam_fixed=arr(p.cam_fixed, Nc, np.uint8), pt=arr(p.pt, Np * 3, np.float64), obs_cam=arr(p.obs_cam, No, np.int32),
                       obs_pt=arr(p.obs_pt, No, np.int32), obs_uv=arr(p.obs_uv, No * 2, np.float64),
                       pose_true=arr(p.pose_true, Nc * 6, np.float64).reshape(Nc, 6), pt_true=arr(p.pt_true, Np * 3, np.float64).reshape(Np, 3),
                       extent=p.extent)
    finally:
        lib.ba_synth_free(pp)
