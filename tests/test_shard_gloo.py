"""world_size-2 (and 3) CPU test of the multi-GPU decomposition over torch.distributed/gloo.

Each rank takes the point range ba_shard_points gives it, forms ITS partial of every camera-side quantity the engine
all-reduces (Hcc | g_c, the reduced right-hand side b, one implicit Schur product y = S x) from the Jacobian blocks of its
own observations only, all-reduces with gloo exactly where the engine calls ncclAllReduce, and rank 0 compares with the
single-process result.  Point-side quantities are never communicated."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def camera_side(P, J, lo, hi, x, damp_pt):
    """Partial sums over observations of points [lo,hi): Hcc[Nc,6,6], gc[Nc,6], b[Nc,6], y[Nc,6] (= this shard's part of S x
    without the camera damping term)."""
    Nc = P.pose.shape[0]
    Jc, Jp, r = J["Jc"], J["Jp"], J["r"]
    Hcc, gc, b, y = np.zeros((Nc, 6, 6)), np.zeros((Nc, 6)), np.zeros((Nc, 6)), np.zeros((Nc, 6))
    o_lo, o_hi = np.searchsorted(P.obs_pt, lo, "left"), np.searchsorted(P.obs_pt, hi, "left")
    for p in range(lo, hi):
        a, e = np.searchsorted(P.obs_pt, p, "left"), np.searchsorted(P.obs_pt, p, "right")
        if a == e:
            continue
        Hpp = sum(Jp[o].T @ Jp[o] for o in range(a, e)) + np.diag(damp_pt[p])
        gp = sum(Jp[o].T @ r[o] for o in range(a, e))
        Hinv = np.linalg.inv(Hpp)
        vp = Hinv @ gp
        w = sum(Jp[o].T @ (Jc[o] @ x[P.obs_cam[o]]) for o in range(a, e))
        v = Hinv @ w
        for o in range(a, e):
            c = P.obs_cam[o]
            b[c] += Jc[o].T @ (r[o] - Jp[o] @ vp)
            y[c] += Jc[o].T @ (Jc[o] @ x[c] - Jp[o] @ v)
    for o in range(o_lo, o_hi):
        c = P.obs_cam[o]
        Hcc[c] += Jc[o].T @ Jc[o]
        gc[c] += Jc[o].T @ r[o]
    return Hcc, gc, b, y


def worker(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    ba = importlib.import_module("3dreconstruction_b200")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        P = ba.synth("tiny", seed=21)
        orc = ba.Library(os.path.join(ROOT, "oracle", "libba_oracle.so"))
        with orc.create(orc.default_options(intrinsics_free=0)) as h:
            h.set_problem(P)
            J = h.evaluate()
        Nc, _, Np, No = P.sizes
        sp = ba.engine_library().shard_points(P.obs_pt, Np, world)
        rng = np.random.default_rng(5)                     # same seed on every rank: x is replicated, like the PCG vectors
        x = rng.normal(size=(Nc, 6))
        damp = np.full((Np, 3), 1e-3)
        parts = camera_side(P, J, int(sp[rank]), int(sp[rank + 1]), x, damp)
        # (1) one all-reduce of [Hcc | gc] per evaluation, (2) b after K4, (3) y after each K5
        red1 = torch.from_numpy(np.concatenate([parts[0].ravel(), parts[1].ravel()]))
        dist.all_reduce(red1)
        b = torch.from_numpy(parts[2].copy()); dist.all_reduce(b)
        y = torch.from_numpy(parts[3].copy()); dist.all_reduce(y)
        # every rank holds bit-identical replicated results (so every rank takes the same LM decisions)
        gathered = [torch.zeros_like(y) for _ in range(world)]
        dist.all_gather(gathered, y)
        same = all(torch.equal(g, gathered[0]) for g in gathered)
        if rank == 0:
            full = camera_side(P, J, 0, Np, x, damp)
            np.savez(out_path, same=same, sp=sp, red1=red1.numpy(), b=b.numpy(), y=y.numpy(),
                     full1=np.concatenate([full[0].ravel(), full[1].ravel()]), fullb=full[2], fully=full[3])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_camera_sums_equal_single_process(world, tmp_path, _built):
    import torch.multiprocessing as mp
    out = str(tmp_path / "out.npz")
    mp.spawn(worker, args=(world, free_port(), out), nprocs=world, join=True)
    d = np.load(out)
    assert bool(d["same"])
    assert d["sp"][0] == 0 and len(d["sp"]) == world + 1
    for got, ref in ((d["red1"], d["full1"]), (d["b"], d["fullb"]), (d["y"], d["fully"])):
        assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()
