#!/usr/bin/env python
"""bench.py — Venice-shape bundle adjustment: LM iterations/s, time-to-converge and HBM roofline fraction on B200.

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA engine); N>1 is launched by torchrun
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU solver path (oracle port) on host cores

One "step" = one full BA solve (Levenberg-Marquardt from the same initial guess to termination with the reference's
Ceres options) of the Venice-shape synthetic problem (1778 cameras, 993 923 points, 5 001 946 observations).
`value` = LM iterations per second with the problem resident in HBM when the timed region starts;
`e2e`   = the same metric through the C-ABI with HOST buffers: ba_set_problem (H2D) + ba_solve + ba_get_state (D2H) every step.
The CPU oracle (oracle/) is used here only for the `cpu_baseline` leg and for `--impl reference`.
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "LM iterations/sec and time-to-converge (Venice-shape BA), % HBM roofline"
UNIT = "LM iterations/s"
KID_K5 = 7  # index of K5_schur_matvec in ba_get_kernel_stats


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def solver_options(args, throughput=False):
    """The reference's configuration (Ceres defaults + BundleAdjuster.h:167-174) with the north star's 2x6/2x3 problem form.
    Parity mode: PCG to a fixed 1e-12 (reproduces the exact-solve trajectory).  Throughput mode: Ceres' inexact-Newton forcing
    sequence (eta = 0.1), judged on final chi2 only."""
    o = dict(intrinsics_free=0, robust_kind=1, robust_a=4.0, max_num_iterations=args.max_lm_iterations, linear_solver=1,
             pcg_rel_tolerance=args.pcg_tol, pcg_max_iterations=2000, preconditioner=args.preconditioner)
    if throughput:
        o.update(pcg_forcing=1, eta=0.1, pcg_max_iterations=500)
    return o


def workload_name(P, args):
    Nc, Ng, Np, No = P.sizes
    return f"{args.workload}-shape synthetic BA: {Nc} cameras, {Np} points, {No} observations, pinhole, intrinsics fixed, Huber(4)"


def load_golden(args):
    """Frozen oracle trajectory of this workload (tests/golden/make_golden_large.py), or None.  Read-only data: the oracle is not run."""
    import numpy as np
    name = {"venice": "venice_exact", "final": "final_pcg"}.get(args.workload)
    path = os.path.join(ROOT, "tests", "golden", f"{name}_oracle_trajectory_large.npz") if name else None
    if not path or not os.path.exists(path):
        return None, None
    return np.load(path), os.path.basename(path)


def parity_against_golden(gold, gname, traj, P):
    """chi2 per LM iteration / accept flags of the LAST timed solve against the frozen oracle trajectory (as far as the file goes)."""
    import numpy as np
    n = min(len(traj), len(gold["cost"]))
    cost = np.array([r.cost for r in traj[:n]])
    rel = np.abs(cost - gold["cost"][:n]) / gold["cost"][:n]
    flags = bool(np.array_equal([r.step_is_successful for r in traj[:n]], gold["accepted"][:n]))
    same_inputs = int(P.obs_cam.astype(np.int64).sum() * 31 + P.obs_pt.astype(np.int64).sum()) == int(gold["obs_checksum"])
    return {"golden": gname, "iterations_compared": int(n - 1), "chi2_rel_err_max": float(rel.max()), "chi2_rel_err_last": float(rel[-1]),
            "accept_flags_equal": flags, "same_inputs": bool(same_inputs), "ok": bool(same_inputs and flags and rel.max() <= 1e-6)}


def cpu_solve_sample(ba, P, args, lm_iterations, threads, linear_solver):
    """Time the CPU oracle (port of the reference's solver path) on a bounded sample: the first `lm_iterations` LM iterations."""
    orc = ba.Library(os.path.join(ROOT, "oracle", "libba_oracle.so"))
    opts = solver_options(args)
    opts.update(max_num_iterations=lm_iterations, num_threads=threads, linear_solver=linear_solver, preconditioner=0)
    with orc.create(orc.default_options(**opts)) as h:
        h.set_problem(P)
        t0 = time.perf_counter()
        s = h.solve()
        dt = time.perf_counter() - t0
    its = max(s.num_iterations - 1, 1)
    return its, dt, s


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU solver path (oracle port; Ceres itself cannot be built here) on all host threads."""
    if rank != 0:
        return
    ba = importlib.import_module("3dreconstruction_b200")
    P = ba.synth(args.workload)
    threads = os.cpu_count() or 1
    times, its_total = [], 0
    for i in range(args.warmup + args.steps):
        its, dt, s = cpu_solve_sample(ba, P, args, args.cpu_lm_iterations, threads, args.cpu_linear_solver)
        if i >= args.warmup:
            times.append(dt); its_total += its
    total = sum(times)
    value = its_total / total
    sample = (f"first {args.cpu_lm_iterations} LM iteration(s) of the {args.workload} solve per step "
              f"({'exact dense Cholesky on S (the reference: SPARSE_SCHUR)' if args.cpu_linear_solver == 0 else 'block-Jacobi PCG on the implicit Schur complement, rel tol %g (the algorithm of the GPU arm)' % args.pcg_tol}), OpenMP")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / max(len(times), 1), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": workload_name(P, args), "step": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def pinned_problem(ba, P):
    """Host copies of the inputs in page-locked memory (what the e2e leg uploads every step)."""
    import torch
    def pin(a):
        return torch.from_numpy(a).pin_memory().numpy()
    return ba.Problem(pose=pin(P.pose), intr=pin(P.intr), cam_group=pin(P.cam_group), cam_fixed=pin(P.cam_fixed), pt=pin(P.pt),
                      obs_cam=pin(P.obs_cam), obs_pt=pin(P.obs_pt), obs_uv=pin(P.obs_uv))


def run_ours(args, rank, world, local_rank):
    import ctypes
    import torch
    import torch.distributed as dist
    ba = importlib.import_module("3dreconstruction_b200")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    lib = ba.engine_library()   # raises if the CUDA extension is missing: no fallback
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line there
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    stream = torch.cuda.current_stream()
    keep = []

    def make_engine(opts):
        """One engine handle per rank; every handle of a multi-GPU run gets its own NCCL communicator (id from rank 0)."""
        opts = dict(opts, device=local_rank, rank=rank, world_size=world, stream=ctypes.c_void_p(stream.cuda_stream), p2p_allreduce=args.p2p)
        if world > 1:
            idt = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(lib.nccl_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            idbuf = ctypes.create_string_buffer(bytes(idt.cpu().numpy().tobytes()), 128)
            keep.append(idbuf)
            opts["nccl_unique_id"] = ctypes.cast(idbuf, ctypes.c_void_p)
        return lib.create(lib.default_options(**opts))

    P = ba.synth(args.workload)
    Ph = pinned_problem(ba, P)
    gold, gname = load_golden(args)

    # ================================ parity mode (PCG to 1e-12): the headline `value` ================================
    eng = make_engine(solver_options(args))
    eng.set_problem(Ph)
    for _ in range(args.warmup):
        eng.reset_state()
        eng.solve()

    # ---- timed region 1: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    eng.set_profiling(1 << KID_K5, stride=args.k5_stride)     # CUDA events around the dominant kernel only, every k5_stride-th launch
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if rank == 0:
        sampler.start()
    ev0.record(stream)
    lm_its = pcg_its = launches = syncs = 0
    for _ in range(args.steps):
        eng.reset_state()
        s = eng.solve()
        lm_its += s.num_iterations - 1; pcg_its += s.total_pcg_iterations; launches += s.kernel_launches
        syncs += eng.execution_info()["host_syncs"]
    ev1.record(stream)
    barrier()
    t_dev = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
    clocks = sampler.stop() if rank == 0 else None
    stats = {k["name"]: k for k in eng.kernel_stats()}
    eng.set_profiling(0)
    final = s
    traj = eng.trajectory()
    info = eng.execution_info()

    # ---- timed region 2: end to end through the C-ABI with host buffers (page-locked, then pageable as the C++ adapter's std::vectors) ----
    def e2e_leg(prob, steps):
        h0, d0 = lib.io_bytes()
        barrier()
        t0 = time.perf_counter()
        its = 0
        for _ in range(steps):
            eng.set_problem(prob)            # host -> device: parameters, observations and the index structure built from them
            s2 = eng.solve()
            eng.get_state()                  # device -> host: the write-back the adapter needs
            its += s2.num_iterations - 1
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        h1, d1 = lib.io_bytes()
        return its / dt, dt / steps, (h1 - h0) // steps, (d1 - d0) // steps

    e2e_v, e2e_s, h2d, d2h = e2e_leg(Ph, args.steps)
    e2e_pg_v, e2e_pg_s, _, _ = e2e_leg(P, 1)
    eng.close()

    # ================================ throughput mode (forcing sequence) and time-to-converge ================================
    # time until chi2 reaches the oracle's final chi2 (within 1e-6).  Target = the frozen exact-solve oracle trajectory's final
    # chi2 * (1 + 1e-6) when the file covers the whole solve (Venice); otherwise the parity-mode solve's own final chi2 * (1 + 1e-6).
    ref_final_cost = float(gold["cost"][-1]) if (gold is not None and len(gold["cost"]) == 51) else float(final.final_cost)
    target = 2.0 * ref_final_cost * (1 + 1e-6)
    topts = dict(solver_options(args, throughput=True), max_num_iterations=4 * args.max_lm_iterations, target_chi2=target)
    eng2 = make_engine(topts)
    eng2.set_problem(Ph)
    eng2.solve()                             # warm-up
    barrier()
    ev0.record(stream)
    t_its = t_pcg = t_launch = 0
    for _ in range(args.steps):
        eng2.reset_state()
        st = eng2.solve()
        t_its += st.num_iterations - 1; t_pcg += st.total_pcg_iterations; t_launch += st.kernel_launches
    ev1.record(stream)
    barrier()
    t_thr = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng2.set_problem(Ph)
        eng2.solve()
        eng2.get_state()
    barrier()
    t_thr_e2e = max_over_ranks(time.perf_counter() - t0)
    # parity-mode time until the same target (the first LM iteration of the timed trajectory at or below it)
    hit = next((r.iteration for r in traj if 2.0 * r.cost <= target), None)
    eng2.close()

    if rank == 0:
        peak_gbs, peak_src = load_peaks()
        k5 = stats["K5_schur_matvec"]
        k5_avg_s = (k5["total_ms"] / max(k5["launches"], 1)) * 1e-3
        achieved = k5["bytes_per_launch"] / k5_avg_s / 1e9 if k5_avg_s > 0 else 0.0
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "k5_dram_bytes_per_launch.json")
        if os.path.exists(tpath) and world == 1 and args.workload == "venice":
            with open(tpath) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        allreduce = {0: "none (single GPU)", 1: "NCCL", 2: "one-shot NVLink peer-memory all-reduce fused into the gather and the PCG tail"}[info["allreduce"]]
        k5_kernel = {1: "k_schur_sweep (L2 atomics)", 4: "k_schur_rows (one CTA per tile)", 6: "k_schur_stream (persistent cp.async.bulk ring)"}[info["scatter_mode"]]
        limiter = None
        survey = None
        if world == 1 and k5_avg_s > 0:
            # SURVEY 8d's formula for one PCG iteration with a materialised Jacobian (152 No + 72 Np + 384 Nc): informational -- `achieved` and
            # `frac` above use the bytes of THIS formulation (restated in BASELINE.md 5 / DESIGN.md 3), which are fewer
            sb = 152.0 * P.sizes[3] + 72.0 * P.sizes[2] + 384.0 * P.sizes[0]
            survey = {"bytes_per_launch": sb, "equivalent_gbs": sb / k5_avg_s / 1e9, "equivalent_frac": sb / k5_avg_s / 1e9 / peak_gbs,
                      "note": "what the measured launch time would mean for the SURVEY's materialised-J formulation of K5; not the roofline claim"}
        if info.get("pose_factored"):
            k5_kernel = "k_schur_rows_f (one CTA per tile, pose-factored: streams the point columns only)"
            # the contract's `bound` stays "hbm" (a byte-streaming FP64 kernel, no tensor-core work); what ncu shows saturated is stated beside it
            limiter = ("load/store data pipe (ncu: l1tex__data_pipe_lsu_wavefronts ~90 % of peak, DRAM ~40 %): with 64 B per observation the per-observation "
                       "gathers, the shared-memory camera sort and the shuffles of the fold cost more pipe cycles than the streams; see DESIGN.md 3.K5f")
        line = {
            "metric": METRIC, "value": lm_its / t_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(P, args),
                       "step": "one full BA solve: LM with the reference's Ceres options (max 50 iterations, tolerances 1e-6/1e-8/1e-10), "
                               f"PCG on the implicit Schur complement to rel. residual {args.pcg_tol:g} (trajectory-parity mode), "
                               f"{'Schur-Jacobi' if args.preconditioner else 'block-Jacobi'} preconditioner",
                       "l2": "inputs larger than L2 (J = 640 MB per sweep vs 126 MB L2)",
                       "parallelism": (f"points sharded over {world} GPUs; per-PCG-iteration all-reduce of the matvec: {allreduce}; other reductions NCCL") if world > 1 else "single GPU"},
            "e2e": {"value": e2e_v, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_s,
                    "host_buffers": "page-locked", "pageable_value": e2e_pg_v, "pageable_ms_per_step": 1e3 * e2e_pg_s},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": f"K5_schur_matvec: {k5_kernel}, one PCG iteration's J sweep", "achieved": achieved, "peak": peak_gbs,
                         "unit": "GB/s", "frac": achieved / peak_gbs, "traffic": traffic, "peak_source": peak_src, "limiter": limiter,
                         "survey_formula": survey,
                         "bytes_per_launch": k5["bytes_per_launch"], "launches": k5["launches"], "avg_us": k5_avg_s * 1e6,
                         "timed_launches": f"every {args.k5_stride}th of the timed region's K5 launches that did work is bracketed by CUDA events on the engine's stream"},
            "solve": {"lm_iterations_per_step": lm_its / args.steps, "pcg_iterations_per_step": pcg_its / args.steps, "ms_per_solve": 1e3 * t_dev / args.steps,
                      "host_syncs_per_step": syncs / args.steps, "termination": final.message.decode(), "initial_chi2": 2 * final.initial_cost,
                      "final_chi2": 2 * final.final_cost, "allreduce": allreduce, "tiles": info["tiles"], "partial_rows": info["rows"]},
            "time_to_converge": {
                "definition": "device time of one solve until chi2 <= target (SURVEY 8d); target = final chi2 of "
                              + ("the frozen exact-solve oracle trajectory (" + gname + ")" if (gold is not None and len(gold["cost"]) == 51) else "the parity-mode solve") + " * (1 + 1e-6)",
                "target_chi2": target,
                "throughput_mode": {"pcg": "Ceres forcing sequence, eta 0.1 (Q-criterion)", "seconds": t_thr / args.steps, "lm_iterations": t_its / args.steps,
                                    "pcg_iterations": t_pcg / args.steps, "lm_iterations_per_s": t_its / t_thr, "termination": st.message.decode(),
                                    "final_chi2": 2 * st.final_cost, "reached_target": bool(st.termination == 3), "e2e_seconds": t_thr_e2e / args.steps,
                                    "gpu_launches_per_solve": t_launch / args.steps},
                "parity_mode": {"pcg": f"fixed rel. tolerance {args.pcg_tol:g}", "lm_iterations_to_target": hit,
                                "seconds": (t_dev / args.steps) * (hit / max(final.num_iterations - 1, 1)) if hit is not None else None,
                                "note": "the solve's device time scaled to the first LM iteration at or below the target"}},
        }
        if gold is not None:
            line["parity"] = parity_against_golden(gold, gname, traj, P)
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            its, dt, so = cpu_solve_sample(ba, P, args, args.cpu_lm_iterations, threads, args.cpu_linear_solver)
            ok = abs(so.final_cost - traj[its].cost) <= 1e-6 * abs(so.final_cost) if len(traj) > its else None
            line["cpu_baseline"] = {"value": its / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"first {its} LM iteration(s) of the same solve on the CPU oracle "
                                              f"({'exact dense Cholesky' if args.cpu_linear_solver == 0 else 'PCG rel tol %g' % args.pcg_tol}), {dt:.1f} s",
                                    "chi2_matches_gpu_trajectory": ok}
            if args.cpu_reference_config and args.workload != "final":
                its2, dt2, so2 = cpu_solve_sample(ba, P, args, 1, threads, 0)
                line["cpu_baseline"]["reference_configuration"] = {
                    "value": its2 / dt2, "unit": UNIT, "cores": threads,
                    "sample": f"first LM iteration with the exact factorisation of the reduced system the reference configures (SPARSE_SCHUR, BundleAdjuster.h:172-173), {dt2:.1f} s",
                    "chi2_matches_gpu_trajectory": abs(so2.final_cost - traj[1].cost) <= 1e-6 * abs(so2.final_cost)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="venice")
    ap.add_argument("--max-lm-iterations", type=int, default=50)
    ap.add_argument("--pcg-tol", type=float, default=1e-12)
    ap.add_argument("--cpu-lm-iterations", type=int, default=1)
    ap.add_argument("--cpu-linear-solver", type=int, default=1, help="0 exact dense Cholesky (the reference's SPARSE_SCHUR semantics), 1 PCG")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-reference-config", type=int, default=1, help="1: also time one LM iteration of the oracle with the exact reduced solve (the reference's configuration)")
    ap.add_argument("--preconditioner", type=int, default=0, help="0 block-Jacobi (blockdiag of Hcc + Dc^2), 1 Schur-Jacobi (blockdiag of S)")
    ap.add_argument("--k5-stride", type=int, default=16, help="time every n-th K5 launch with CUDA events (an event between two launches serialises the chain's dependent launches)")
    ap.add_argument("--p2p", type=int, default=0, help="multi-GPU matvec all-reduce: 0 auto (NVLink peer memory if mappable, else NCCL), 1 NCCL only, 2 peer memory required")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus and not (world == 1 and args.gpus == 1):
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torchrun for multi-GPU", file=sys.stderr)
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
