"""ba_cost_noalloc (include/ba_engine.h): the entry point that allocates nothing — its scratch lives in a file-scope static device array;
no counterpart in the reference.  Collected last on purpose.

GPU: called through the C-ABI with device arrays, checked against the existing entry point (ba_evaluate's cost of a handle holding the same
problem) on calls in a row that share the static array — same problem twice, a different problem in between, both residual models.
The inputs and outputs are torch CUDA tensors (the test may allocate; the entry point must not)."""
import ctypes as C
import importlib

import numpy as np
import pytest

COST_RTOL = 1e-12     # the tolerance tests/test_engine_gpu.py::test_evaluate_parity uses for the cost


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def cost_noalloc(lib, torch, P, model=0, robust_kind=1, robust_a=4.0, stream=None):
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(np.ascontiguousarray(getattr(P, k))).to(dev) for k in ("pose", "intr", "cam_group", "pt", "obs_cam", "obs_pt", "obs_uv")}
    cost = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    bad = torch.full((1,), 7, dtype=torch.int32, device=dev)
    Nc, Ng, Np, No = P.sizes
    s = C.c_void_p(stream.cuda_stream) if stream is not None else None
    if stream is not None:
        stream.wait_stream(torch.cuda.current_stream())
    rc = lib.lib.ba_cost_noalloc(0, s, model, robust_kind, robust_a, 0, Nc, Ng, No, _ptr(t["pose"]), _ptr(t["intr"]), _ptr(t["cam_group"]), _ptr(t["pt"]),
                                 _ptr(t["obs_cam"]), _ptr(t["obs_pt"]), _ptr(t["obs_uv"]), _ptr(cost), _ptr(bad))
    assert rc == 0, lib.lib.ba_last_error(None).decode()
    torch.cuda.synchronize()
    return float(cost.item()), int(bad.item())


def cost_by_handle(ba, P, **opts):
    with ba.Engine(intrinsics_free=0, **opts) as eng:
        eng.set_problem(P)
        return eng.evaluate()["cost"]


@pytest.mark.gpu
def test_noalloc_cost_matches_the_handle_entry_point(ba, engine_lib):
    torch = importlib.import_module("torch")
    A, B = ba.synth("ladybug"), ba.synth("trafalgar", seed=3)
    refA, refB = cost_by_handle(ba, A), cost_by_handle(ba, B)
    mem0 = torch.cuda.memory_allocated()
    calls = [(A, refA), (A, refA), (B, refB), (A, refA)]          # in a row on the shared static array: repeat, other sizes, back
    for P, ref in calls:
        got, bad = cost_noalloc(engine_lib, torch, P)
        assert bad == 0 and abs(got - ref) <= COST_RTOL * abs(ref), (got, ref)
    # on a non-default stream, and the other robust kernels
    st = torch.cuda.Stream()
    for kind in (0, 2, 1):
        got, bad = cost_noalloc(engine_lib, torch, A, robust_kind=kind, stream=st)
        ref = cost_by_handle(ba, A, robust_kind=kind)
        assert bad == 0 and abs(got - ref) <= COST_RTOL * abs(ref), (kind, got, ref)
    assert torch.cuda.memory_allocated() == mem0                  # nothing of ours lives in torch's allocator either


@pytest.mark.gpu
def test_noalloc_snavely_model_and_bad_flag(ba, engine_lib):
    torch = importlib.import_module("torch")
    from test_snavely_bal import snavely_problem
    P = snavely_problem(ba, "ladybug", seed=2)
    ref = cost_by_handle(ba, P, model=1)
    for _ in range(2):
        got, bad = cost_noalloc(engine_lib, torch, P, model=1)
        assert bad == 0 and abs(got - ref) <= COST_RTOL * abs(ref), (got, ref)
    Q = ba.synth("tiny")
    Q.pt[0] = np.nan
    got, bad = cost_noalloc(engine_lib, torch, Q)
    assert bad == 1
    got, bad = cost_noalloc(engine_lib, torch, ba.synth("tiny"))      # the flag does not stick to the next call
    assert bad == 0 and np.isfinite(got)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(ba, engine_lib):
    """CPU: the size check comes before any device use; the limit is the header's BA_NOALLOC_MAX_CAMERAS."""
    import os, re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "ba_engine.h")).read()
    limit = int(re.search(r"#define BA_NOALLOC_MAX_CAMERAS (\d+)", hdr).group(1))
    one = C.c_void_p(16)
    args = lambda nc: (0, None, 0, 1, 4.0, 0, nc, 1, 10, one, one, one, one, one, one, one, one, one)
    rc = engine_lib.lib.ba_cost_noalloc(*args(limit + 1))
    assert rc == -1
    msg = engine_lib.lib.ba_last_error(None).decode()
    assert "BA_NOALLOC_MAX_CAMERAS" in msg and str(limit) in msg
    assert engine_lib.lib.ba_cost_noalloc(0, None, 0, 1, 4.0, 0, 4, 1, 10, one, one, one, one, one, one, one, None, one) == -1   # null output
