"""Scratch GPU probe: solve a named shape on the engine, print trajectory summary and per-kernel roofline numbers."""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")
PEAK = 6553.0e9
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(_pk):
    PEAK = float(json.load(open(_pk))["hbm_gbs"]) * 1e9
name = sys.argv[1] if len(sys.argv) > 1 else "venice"
opts = {}
for a in sys.argv[2:]:
    k, v = a.split("=")
    opts[k] = float(v) if "." in v or "e" in v else int(v)
t = time.time(); P = ba.synth(name); print(name, P.sizes, "gen %.2fs" % (time.time() - t), flush=True)
base = dict(intrinsics_free=0)
base.update(opts)
with ba.Engine(**base) as eng:
    t = time.time(); eng.set_problem(P); print("set_problem %.3fs" % (time.time() - t), flush=True)
    s = eng.solve()  # warm-up
    print("warm: term", s.termination, s.message.decode(), "iters", s.num_iterations, "pcg", s.total_pcg_iterations, "cost %.9e -> %.9e" % (s.initial_cost, s.final_cost), "dev %.1f ms" % (s.solve_seconds * 1e3), flush=True)
    eng.reset_state()
    t = time.time(); s = eng.solve(); wall = time.time() - t
    print("plain: dev %.1f ms wall %.1f ms  LM it/s %.2f  pcg %d  launches %d" % (s.solve_seconds * 1e3, wall * 1e3, (s.num_iterations - 1) / s.solve_seconds, s.total_pcg_iterations, s.kernel_launches), eng.execution_info(), flush=True)
    eng.reset_state(); eng.set_profiling(True)
    s = eng.solve()
    print("profiled: dev %.1f ms" % (s.solve_seconds * 1e3))
    tot = sum(k["total_ms"] for k in eng.kernel_stats())
    for k in eng.kernel_stats():
        if k["launches"] == 0: continue
        ms = k["total_ms"] / k["launches"]
        gbs = k["bytes_per_launch"] / (ms * 1e-3) / 1e9 if ms > 0 else 0
        print("  %-22s n=%6d  total %9.3f ms (%5.1f%%)  avg %8.1f us  alg %8.2f MB  %7.1f GB/s  frac %.3f" % (k["name"], k["launches"], k["total_ms"], 100 * k["total_ms"] / max(tot, 1e-9), ms * 1e3, k["bytes_per_launch"] / 1e6, gbs, gbs * 1e9 / PEAK))
    for r in eng.trajectory()[:60]:
        print("   %2d v%d s%d cost %.9e cand %.9e rho %.3f rad %.2e pcg %d rel %.1e" % (r.iteration, r.step_is_valid, r.step_is_successful, r.cost, r.candidate_cost, r.relative_decrease, r.radius, r.pcg_iterations, r.pcg_rel_residual))
