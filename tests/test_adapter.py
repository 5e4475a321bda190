"""The C++ drop-in adapter (3dreconstruction_b200/host/adjuster/BundleAdjuster.h): `BundleAdjuster adj; adj(world);` over a
WorldStructure built the way SequentialActuator builds it, compared with a direct C-ABI solve of the same problem, plus KAT-10
(pose write-back: correct exponential map vs the reference's Euler-ZYX bug).
CPU: the driver is linked against the oracle library (same header; checker use).  GPU: against libba_engine.so."""
import os
import subprocess

import cv2
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "adapter_main.cpp")


def build_driver(tmp_path, libdir, libname):
    exe = str(tmp_path / f"adapter_{libname}")
    host = os.path.join(ROOT, "3dreconstruction_b200", "host")
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-Wall", SRC, "-o", exe, f"-L{libdir}", f"-l{libname}", f"-L{host}", "-lba_synth",
                           f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{host}", "-Wl,-rpath,/usr/local/cuda/lib64"])
    return exe


def run_driver(exe, tmp_path, shape, euler_bug=0, fi=1, max_it=50):
    out = str(tmp_path / "world.txt")
    txt = subprocess.run([exe, shape, out, str(euler_bug), str(fi), str(max_it)], check=True, capture_output=True, text=True).stdout
    d = dict(pose=[], pt=[], iter=[])
    for line in open(out):
        k, *v = line.split()
        if k in ("summary", "intr"):
            d[k] = [float(x) for x in v]
        else:
            d[k].append([float(x) for x in v])
    return {k: np.array(v) for k, v in d.items()}, txt


def reference_flow_problem(ba, shape):
    """The problem the adapter must derive from the world: pose 0 all-zero and free, pose 1 constant (F8a/b)."""
    P = ba.synth(shape)
    assert not P.pose[0].any() and P.cam_fixed[1] == 1 and P.cam_fixed.sum() == 1
    return P


def rot(aa):
    return cv2.Rodrigues(np.asarray(aa, np.float64).reshape(3, 1))[0]


def check_against_direct_solve(ba, lib_handle, exe, tmp_path, fi):
    P = reference_flow_problem(ba, "sequential30")
    w, txt = run_driver(exe, tmp_path, "sequential30", fi=fi, max_it=10)
    with lib_handle(intrinsics_free=fi, max_num_iterations=10) as h:
        h.set_problem(P)
        h.solve()
        t = h.trajectory()
        pose, intr, pt = h.get_state()
    assert "Bundle Adjustment statistics" in txt and f"#views: {P.sizes[0]}" in txt and f"#residuals: {2 * P.sizes[3]}" in txt
    assert int(w["summary"][2]) == len(t) and np.array_equal(w["iter"][:, 1], [r.step_is_successful for r in t])
    assert np.allclose(w["iter"][:, 2], [r.cost for r in t], rtol=1e-6)
    assert np.linalg.norm(w["pt"] - pt) <= 1e-5 * np.linalg.norm(pt)
    assert np.allclose(w["intr"], intr[0], rtol=1e-5)
    for c in range(P.sizes[0]):
        assert np.abs(w["pose"][c, :9].reshape(3, 3) - rot(pose[c, :3])).max() <= 1e-5
        assert np.allclose(w["pose"][c, 9:], pose[c, 3:], rtol=1e-5, atol=1e-7)
    assert np.array_equal(w["pose"][1, 9:], P.pose[1, 3:])                      # the constant pose is written back unchanged


def check_writeback(ba, exe, tmp_path):
    """KAT-10 with zero LM iterations: the write-back alone."""
    P = reference_flow_problem(ba, "sequential30")
    w, _ = run_driver(exe, tmp_path, "sequential30", euler_bug=0, max_it=0)
    for c in range(P.sizes[0]):
        assert np.abs(w["pose"][c, :9].reshape(3, 3) - rot(P.pose[c, :3])).max() <= 2e-15    # correct_angle_axis round-trips
    assert np.array_equal(w["pt"], P.pt)
    w, _ = run_driver(exe, tmp_path, "sequential30", euler_bug=1, max_it=0)
    worst = 0.0
    for c in range(P.sizes[0]):
        a0, a1, a2 = P.pose[c, :3]
        Rz, Ry, Rx = rot([0, 0, a0]), rot([0, a1, 0]), rot([a2, 0, 0])
        assert np.abs(w["pose"][c, :9].reshape(3, 3) - Rz @ Ry @ Rx).max() <= 2e-15               # reference_euler_bug: Rz(a0) Ry(a1) Rx(a2)
        worst = max(worst, np.abs(Rz @ Ry @ Rx - rot(P.pose[c, :3])).max())
    assert worst > 1e-3                                                                           # and that is NOT the same rotation


def test_euler_bug_magnitude_kat10():
    """at aa = (3,-5,8)e-4 the reference's write-back is off by 0.0405 degrees."""
    aa = np.array([3, -5, 8]) * 1e-4
    R = rot([0, 0, aa[0]]) @ rot([0, aa[1], 0]) @ rot([aa[2], 0, 0])
    err = np.degrees(np.linalg.norm(cv2.Rodrigues(R @ rot(aa).T)[0]))
    assert err == pytest.approx(0.0405, abs=5e-4)


@pytest.mark.parametrize("fi", [1, 0])
def test_adapter_matches_direct_solve_cpu(ba, oracle, tmp_path, fi):
    exe = build_driver(tmp_path, os.path.join(ROOT, "oracle"), "ba_oracle")
    check_against_direct_solve(ba, lambda **kw: oracle.create(oracle.default_options(num_threads=4, **kw)), exe, tmp_path, fi)


def test_adapter_writeback_cpu(ba, tmp_path):
    check_writeback(ba, build_driver(tmp_path, os.path.join(ROOT, "oracle"), "ba_oracle"), tmp_path)


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="a GPU is present")
def test_adapter_without_gpu_fails_loudly_and_leaves_world_untouched(ba, tmp_path):
    exe = build_driver(tmp_path, os.path.join(ROOT, "3dreconstruction_b200", "csrc"), "ba_engine")
    P = reference_flow_problem(ba, "sequential30")
    w, txt = run_driver(exe, tmp_path, "sequential30")
    assert "Bundle Adjustment failed." in txt and "no CPU fallback" in txt
    assert np.array_equal(w["pt"], P.pt)


@pytest.mark.gpu
@pytest.mark.parametrize("fi", [1, 0])
def test_adapter_matches_direct_solve_gpu(ba, tmp_path, fi):
    exe = build_driver(tmp_path, os.path.join(ROOT, "3dreconstruction_b200", "csrc"), "ba_engine")
    check_against_direct_solve(ba, lambda **kw: ba.Engine(**kw), exe, tmp_path, fi)
    check_writeback(ba, exe, tmp_path)


# ---- IncrementalBundleAdjuster: one adjuster kept alive over a growing map == a fresh BundleAdjuster per call -------------------
INC_SRC = os.path.join(ROOT, "tests", "cpp", "incremental_main.cpp")


def build_incremental(tmp_path, libdir, libname):
    exe = str(tmp_path / f"incremental_{libname}")
    host = os.path.join(ROOT, "3dreconstruction_b200", "host")
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-Wall", INC_SRC, "-o", exe, f"-L{libdir}", f"-l{libname}", f"-L{host}", "-lba_synth",
                           f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{host}", "-Wl,-rpath,/usr/local/cuda/lib64"])
    return exe


def run_incremental(exe, tmp_path, mode, scenario, cuts=(8, 14, 20, 26, 30), max_it=6):
    out = str(tmp_path / f"inc_{mode}_{scenario}.txt")
    subprocess.run([exe, "sequential30", out, mode, str(scenario), str(max_it)] + [str(c) for c in cuts], check=True, capture_output=True, text=True)
    d = dict(stage=[], pose=[], pt=[])
    for line in open(out):
        k, *v = line.split()
        if k in ("calls", "intr"):
            d[k] = [float(x) for x in v]
        else:
            d[k].append([float(x) for x in v])
    return {k: np.array(v) for k, v in d.items()}


def check_incremental(exe, tmp_path):
    for scenario, (appended, rebuilt) in {0: (4, 1), 1: (4, 1), 2: (3, 2)}.items():
        a = run_incremental(exe, tmp_path, "fresh", scenario)
        b = run_incremental(exe, tmp_path, "incremental", scenario)
        assert list(b["calls"]) == [appended, rebuilt], scenario        # growth (and an edited point) ride ba_append; a lost observation forces a rebuild
        assert a["stage"][:, 1].all() and np.array_equal(a["stage"], b["stage"]), scenario      # usable, same iterations, bit-identical costs
        for k in ("pose", "pt", "intr"):
            assert np.array_equal(a[k], b[k]), (scenario, k)
        assert (a["stage"][1:, 3] > a["stage"][:-1, 4]).all()           # every stage really added residuals to a converged map


def test_incremental_adjuster_matches_fresh_adjuster_cpu(tmp_path):
    check_incremental(build_incremental(tmp_path, os.path.join(ROOT, "oracle"), "ba_oracle"), tmp_path)


@pytest.mark.gpu
def test_incremental_adjuster_matches_fresh_adjuster_gpu(tmp_path):
    check_incremental(build_incremental(tmp_path, os.path.join(ROOT, "3dreconstruction_b200", "csrc"), "ba_engine"), tmp_path)
