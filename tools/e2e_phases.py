"""Host-side phase times of the e2e leg (ba_set_problem / ba_solve / ba_get_state) on the Venice shape, pinned host buffers as bench.py uses.
usage: BA_TIMING=1 python tools/e2e_phases.py [shape]"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
ba = importlib.import_module("3dreconstruction_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "venice"
P = ba.synth(name)
pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
Ph = ba.Problem(pose=pin(P.pose), intr=pin(P.intr), cam_group=pin(P.cam_group), cam_fixed=pin(P.cam_fixed), pt=pin(P.pt),
                obs_cam=pin(P.obs_cam), obs_pt=pin(P.obs_pt), obs_uv=pin(P.obs_uv))
with ba.Engine(intrinsics_free=0, max_num_iterations=3) as eng:
    for rep in range(3):
        t0 = time.perf_counter(); eng.set_problem(Ph); t1 = time.perf_counter()
        s = eng.solve(); t2 = time.perf_counter()
        eng.get_state(); t3 = time.perf_counter()
        print(f"rep {rep}: set_problem {1e3*(t1-t0):.1f} ms, solve wall {1e3*(t2-t1):.1f} ms (device {1e3*s.solve_seconds:.1f}), get_state {1e3*(t3-t2):.1f} ms", flush=True)
