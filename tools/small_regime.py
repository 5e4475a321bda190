"""Where the time goes on the small problems of the incremental regime: per shape, set_problem and solve wall time,
device solve time, launches, LM and PCG iterations, microseconds per launch."""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")
shapes = sys.argv[1:] or ["tiny", "sequential30", "ladybug", "trafalgar"]
for name in shapes:
    P = ba.synth(name)
    with ba.Engine(max_num_iterations=50) as e:
        for rep in range(3):
            t0 = time.perf_counter(); e.set_problem(P); t1 = time.perf_counter(); s = e.solve(); t2 = time.perf_counter()
        tr = e.trajectory()
        e.set_profiling(True); e.set_problem(P); e.solve()
        ks = {k['name']: (k['launches'], round(1e3 * k['total_ms'] / max(k['launches'], 1), 1)) for k in e.kernel_stats() if k['launches'] and (k['name'][0] == 'D' or os.environ.get('ALLK'))}
        pcg = s.total_pcg_iterations
        print(json.dumps(dict(shape=name, sizes=P.sizes, set_problem_ms=round((t1 - t0) * 1e3, 3), solve_wall_ms=round((t2 - t1) * 1e3, 3),
                              solve_dev_ms=round(s.solve_seconds * 1e3, 3), launches=s.kernel_launches, lm=len(tr) - 1, pcg=pcg,
                              us_per_launch=round(s.solve_seconds * 1e6 / max(s.kernel_launches, 1), 2), kernels_n_avg_us=ks, solver=e.linear_solver_in_use(), term=s.message.decode() if isinstance(s.message, bytes) else s.message)), flush=True)
