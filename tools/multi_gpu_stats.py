"""Per-kernel event timings of a sharded solve, launched with torchrun (one rank per GPU): where does a PCG iteration go at N > 1?
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29513 tools/multi_gpu_stats.py [shape] [lm_iterations] [p2p]"""
import ctypes, importlib, os, sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
    lm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    p2p = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = ba.engine_library()
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(lib.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    idbuf = ctypes.create_string_buffer(bytes(idt.cpu().numpy().tobytes()), 128)
    P = ba.synth(shape)
    with lib.create(lib.default_options(rank=rank, world_size=world, p2p_allreduce=p2p, nccl_unique_id=ctypes.cast(idbuf, ctypes.c_void_p),
                                        intrinsics_free=0, max_num_iterations=lm, device=local)) as eng:
        eng.set_problem(P)
        eng.solve()
        eng.reset_state()
        s = eng.solve()
        plain = s.solve_seconds
        eng.reset_state(); eng.set_profiling(True)
        s2 = eng.solve()
        for r in range(world):
            dist.barrier()
            if r == rank and r in (0, world - 1):
                print(f"rank {rank}/{world} {shape} p2p={p2p}: plain {plain*1e3:.1f} ms, profiled {s2.solve_seconds*1e3:.1f} ms, pcg {s.total_pcg_iterations}, "
                      f"{plain*1e6/max(s.total_pcg_iterations,1):.1f} us per PCG iteration (all kernels)", flush=True)
                for k in eng.kernel_stats():
                    if k["launches"]:
                        print("   %-24s n=%6d avg %8.1f us  total %8.2f ms" % (k["name"], k["launches"], 1e3 * k["total_ms"] / k["launches"], k["total_ms"]), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
