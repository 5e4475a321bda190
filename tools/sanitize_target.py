"""Small target for compute-sanitizer (memcheck / racecheck / synccheck): one short PCG solve and one direct solve of a small shape, the no-alloc
entry point's kernels included via the engine's own paths.  usage: compute-sanitizer --tool racecheck python tools/sanitize_target.py [shape]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")
shape = sys.argv[1] if len(sys.argv) > 1 else "tiny"
P = ba.synth(shape)
for opts in (dict(linear_solver=1, intrinsics_free=0), dict(linear_solver=1, intrinsics_free=1), dict(linear_solver=0, intrinsics_free=1), dict(linear_solver=1, intrinsics_free=0, pcg_forcing=1)):
    with ba.Engine(max_num_iterations=2, pcg_max_iterations=12, **opts) as eng:
        eng.set_problem(P)
        s = eng.solve()
        print(opts, "iterations", s.num_iterations, "pcg", s.total_pcg_iterations, "cost %.6e -> %.6e" % (s.initial_cost, s.final_cost), flush=True)
