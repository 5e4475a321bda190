"""Multi-GPU parity check, launched with torchrun (one rank per GPU): the sharded NCCL solve must reproduce the single-GPU
trajectory (chi2 within 1e-6 relative, same accept/reject flags) and every rank must end with the same full state.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py [shape ...]"""
import ctypes
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = ba.engine_library()
    ok = True
    for shape in (sys.argv[1:] or ["ladybug", "trafalgar"]):
        for fi, p2p in ((0, 2), (1, 2), (0, 1)):   # p2p 2: peer-memory all-reduce required; 1: NCCL only
            idt = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(lib.nccl_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            idbuf = ctypes.create_string_buffer(bytes(idt.cpu().numpy().tobytes()), 128)
            P = ba.synth(shape)
            common = dict(intrinsics_free=fi, max_num_iterations=10, device=local)
            with lib.create(lib.default_options(rank=rank, world_size=world, p2p_allreduce=p2p, nccl_unique_id=ctypes.cast(idbuf, ctypes.c_void_p), **common)) as eng:
                eng.set_problem(P)
                s = eng.solve()
                t = eng.trajectory()
                pose, intr, pt = eng.get_state()
            # every rank holds the same replicated result
            chk = torch.tensor([s.final_cost, float(np.abs(pose).sum()), float(np.abs(pt).sum())], dtype=torch.float64, device=dev)
            lo, hi = chk.clone(), chk.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            same = bool(torch.equal(lo, hi))
            if rank == 0:
                with lib.create(lib.default_options(**common)) as one:
                    one.set_problem(P)
                    s1 = one.solve()
                    t1 = one.trajectory()
                    pose1, intr1, pt1 = one.get_state()
                good = same and len(t) == len(t1) and all(a.step_is_successful == b.step_is_successful and abs(a.cost - b.cost) <= 1e-6 * abs(b.cost) for a, b in zip(t, t1))
                good = good and np.linalg.norm(pt - pt1) <= 1e-5 * np.linalg.norm(pt1) and np.linalg.norm(pose - pose1) <= 1e-5 * np.linalg.norm(pose1)
                print(f"{shape} fi={fi} p2p={p2p} world={world}: {'OK' if good else 'MISMATCH'} iters {len(t)} vs {len(t1)} final cost {s.final_cost:.12e} vs {s1.final_cost:.12e} "
                      f"ranks identical: {same}; device time {s.solve_seconds*1e3:.1f} ms vs {s1.solve_seconds*1e3:.1f} ms single", flush=True)
                ok = ok and good
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(int(flag.item() != 0))


if __name__ == "__main__":
    main()
