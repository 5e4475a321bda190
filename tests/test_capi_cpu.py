"""CPU checks of the drop-in boundary: both libraries load, export every symbol include/*.h declares, agree on the
option defaults, validate arguments identically — and the CUDA engine refuses to run (loudly) without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header):
    """Function names declared in a C header: identifiers followed by '(' at the start of a prototype line."""
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:const\s+)?(?:int|void|char\s*\*|const char\s*\*)\s*\*?\s*(ba_[a-z0-9_]+)\s*\(", txt, flags=re.M)))


def test_header_symbols_are_exported_by_engine_and_oracle(ba):
    names = declared_symbols("ba_engine.h")
    assert len(names) >= 17 and "ba_solve" in names and "ba_shard_points" in names
    for path in (ba.ENGINE_LIB, os.path.join(ROOT, "oracle", "libba_oracle.so")):
        lib = C.CDLL(path)
        for n in names:
            assert hasattr(lib, n), f"{os.path.basename(path)} does not export {n}"
    assert sorted(n for n, _, _ in ba._capi.ENGINE_SYMBOLS) == names     # the ctypes mirror binds exactly the header


def test_synth_header_symbols(ba):
    names = declared_symbols("ba_synth.h")
    lib = C.CDLL(ba.SYNTH_LIB)
    assert names and all(hasattr(lib, n) for n in names)
    assert sorted(n for n, _, _ in ba._capi.SYNTH_SYMBOLS) == names


def test_struct_sizes_match_the_header(ba, tmp_path):
    """sizeof() of every boundary struct as the C compiler sees it == the ctypes mirror."""
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "ba_engine.h"\n#include "ba_synth.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(ba_options),sizeof(ba_problem),sizeof(ba_summary),sizeof(ba_iter_record),sizeof(ba_kernel_stat),"
                   "sizeof(ba_synth_spec),sizeof(ba_synth_problem));return 0;}\n")
    exe = tmp_path / "sz"
    import subprocess
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    k = ba._capi
    want = [C.sizeof(t) for t in (k.BAOptions, k.BAProblem, k.BASummary, k.BAIterRecord, k.BAKernelStat, k.SynthSpec, k.SynthProblem)]
    assert got == want


def test_backends(ba, oracle):
    assert ba.engine_library().backend() == "cuda-sm100a"
    assert oracle.backend() == "cpu-oracle"


def test_option_defaults_are_the_reference_configuration(ba, oracle):
    """Ceres 2.2 defaults + BundleAdjuster() (src/adjuster/BundleAdjuster.h:109,167-174), identical in both libraries
    except the reduced-system solver (oracle: exact, as the reference; engine: PCG)."""
    e, o = ba.engine_library().default_options(), oracle.default_options()
    for name, _ in ba._capi.BAOptions._fields_:
        if name in ("linear_solver", "reserved", "nccl_unique_id", "stream"):
            continue
        assert getattr(e, name) == getattr(o, name), name
    assert (e.linear_solver, o.linear_solver) == (2, 0)      # engine AUTO, oracle EXACT
    assert (o.robust_kind, o.robust_a, o.intrinsics_free, o.num_threads) == (1, 4.0, 1, 1)
    assert (o.max_num_iterations, o.initial_trust_region_radius, o.min_relative_decrease) == (50, 1e4, 1e-3)
    assert (o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance) == (1e-6, 1e-10, 1e-8)
    assert (o.min_lm_diagonal, o.max_lm_diagonal, o.max_num_consecutive_invalid_steps, o.jacobi_scaling) == (1e-6, 1e32, 5, 1)


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="a GPU is present")
def test_engine_has_no_cpu_fallback(ba):
    """Without a CUDA device ba_create fails with BA_ERR_CUDA and says so; nothing is solved on the CPU."""
    lib = ba.engine_library()
    with pytest.raises(ba.BAError) as ei:
        lib.create(lib.default_options())
    assert "(-3)" in str(ei.value) and "no CPU fallback" in str(ei.value)


def test_engine_rejects_unsupported_options_before_touching_a_device(ba):
    lib = ba.engine_library()
    for bad in (dict(linear_solver=7), dict(abi_version=99), dict(world_size=2), dict(rank=3), dict(preconditioner=1), dict(preconditioner=5)):   # SCHUR_JACOBI is an opt-in (BA_EXPERIMENTAL_SCHUR_JACOBI=1)
        with pytest.raises(ba.BAError):
            lib.create(lib.default_options(**bad))


def test_argument_validation(ba, oracle):
    P = ba.synth("tiny")
    with oracle.create(oracle.default_options()) as h:
        with pytest.raises(ba.BAError):
            h.solve()                                        # BA_ERR_NO_PROBLEM
        Q = ba.Problem(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt, obs_cam=P.obs_cam,
                       obs_pt=P.obs_pt[::-1].copy(), obs_uv=P.obs_uv)
        with pytest.raises(ba.BAError, match="sorted by point"):
            h.set_problem(Q)
        Q = ba.Problem(pose=P.pose, intr=P.intr, cam_group=P.cam_group, cam_fixed=P.cam_fixed, pt=P.pt, obs_cam=P.obs_cam + 100,
                       obs_pt=P.obs_pt, obs_uv=P.obs_uv)
        with pytest.raises(ba.BAError, match="out of range"):
            h.set_problem(Q)


def test_empty_problem(ba, oracle):
    """No observations at all: cost 0, converges immediately, state untouched (edge case: empty input)."""
    P = ba.Problem(pose=np.zeros((2, 6)), intr=[[1, 1, 0, 0]], cam_group=[0, 0], cam_fixed=[0, 1], pt=np.ones((3, 3)),
                   obs_cam=np.zeros(0, np.int32), obs_pt=np.zeros(0, np.int32), obs_uv=np.zeros((0, 2)))
    with oracle.create(oracle.default_options()) as h:
        h.set_problem(P)
        s = h.solve()
        pose, intr, pt = h.get_state()
    assert s.usable and s.initial_cost == 0.0 and s.final_cost == 0.0
    assert np.array_equal(pt, P.pt) and np.array_equal(pose, P.pose)


def test_shard_points_partition(ba):
    """ba_shard_points (host-only): contiguous, complete, cut at point boundaries, balanced to within one track."""
    lib = ba.engine_library()
    P = ba.synth("ladybug")
    Np, No = P.sizes[2], P.sizes[3]
    counts = np.bincount(P.obs_pt, minlength=Np)
    for W in (1, 2, 3, 4, 8):
        sp = lib.shard_points(P.obs_pt, Np, W)
        assert sp[0] == 0 and sp[-1] == Np and np.all(np.diff(sp) >= 0)
        per = [counts[sp[r]:sp[r + 1]].sum() for r in range(W)]
        assert sum(per) == No
        assert max(per) - min(per) <= 2 * counts.max()
    # more ranks than points: trailing shards are empty, nothing is lost
    sp = lib.shard_points(np.array([0, 0, 1], np.int32), 2, 8)
    assert sp[0] == 0 and sp[-1] == 2 and np.all(np.diff(sp) >= 0)
    with pytest.raises(ba.BAError):
        lib.shard_points(np.array([1, 0], np.int32), 2, 2)   # not sorted


def test_profiling_entry_points(ba, oracle):
    """ba_set_profiling / ba_set_profiling_stride / ba_get_kernel_stats exist on both libraries; the stride is validated (>= 1)."""
    with oracle.create(oracle.default_options()) as h:
        h.set_profiling(1 << 7, stride=16)          # the call bench.py makes
        h.set_profiling(0)
        assert h.kernel_stats() == []               # the CPU oracle has no kernels to report
        with pytest.raises(ba.BAError):
            h.set_profiling(1, stride=0)
