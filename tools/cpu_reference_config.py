"""CPU timings of the oracle in the REFERENCE'S OWN configuration (BundleAdjuster.h:167-174: num_threads = 1, SPARSE_SCHUR = exact factorisation of
the reduced camera system) and on all cores, full 50-iteration solves of the small shapes.  usage: python tools/cpu_reference_config.py [shape ...]"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ba = importlib.import_module("3dreconstruction_b200")
orc = ba.Library(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "libba_oracle.so"))
ncpu = os.cpu_count() or 1
print(f"host: {ncpu} hardware threads")
for shape in (sys.argv[1:] or ["sequential30", "ladybug", "trafalgar"]):
    P = ba.synth(shape)
    for fi in (1, 0):
        for threads in (1, ncpu):
            with orc.create(orc.default_options(intrinsics_free=fi, linear_solver=0, num_threads=threads)) as h:
                h.set_problem(P)
                t0 = time.perf_counter(); s = h.solve(); dt = time.perf_counter() - t0
            its = s.num_iterations - 1
            print(f"{shape:13s} {P.sizes} intrinsics_free={fi} threads={threads:2d}: {its} LM iterations in {dt:8.3f} s = {its/dt:8.2f} LM it/s, "
                  f"{1e3*dt/max(its,1):8.2f} ms per LM iteration, final cost {s.final_cost:.9e} ({s.message.decode()})", flush=True)
