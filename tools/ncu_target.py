"""Short single-GPU target for ncu: ONE Venice-shape solve, 2 LM iterations, PCG capped, so the launch list stays small.
usage: python tools/ncu_target.py [shape] [lm_iterations] [pcg_cap] [key=value ...]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "venice"
lm = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cap = int(sys.argv[3]) if len(sys.argv) > 3 else 24
opts = dict(intrinsics_free=0, max_num_iterations=lm, pcg_max_iterations=cap)
for a in sys.argv[4:]:
    k, v = a.split("=")
    opts[k] = float(v) if "." in v or "e" in v else int(v)
P = ba.synth(name)
with ba.Engine(**opts) as eng:
    eng.set_problem(P)
    s = eng.solve()
    print(name, P.sizes, "iters", s.num_iterations, "pcg", s.total_pcg_iterations, "launches", s.kernel_launches, "cost %.9e -> %.9e" % (s.initial_cost, s.final_cost))
