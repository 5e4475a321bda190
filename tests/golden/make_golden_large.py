"""Frozen oracle trajectories of the headline configurations (BASELINE.json configs[3] and configs[4]).

    python tests/golden/make_golden_large.py venice_exact     # ~20-40 min on 8 host cores, 2 GB
    python tests/golden/make_golden_large.py venice_forcing   # minutes
    python tests/golden/make_golden_large.py final_pcg        # ~20 min, 8 GB

The small shapes are pinned by make_golden.py.  These are the same kind of file ("oracle": the CPU oracle's Levenberg-Marquardt
trajectory, frozen) for the two shapes the north star's acceptance sentence names, generated once in the CPU container because the
oracle needs tens of minutes on them; the GPU tests (tests/test_headline_gpu.py) hold the CUDA engine to them on the GPU box,
where the generator (bit-stable: the files carry a checksum of the observation list) re-creates the inputs.

  venice_exact   : Venice shape, the reference's configuration (exact factorisation of the reduced camera system,
                   BundleAdjuster.h:172-173), intrinsics fixed, Huber(4), 50 iterations: chi2 / accept flags / radius per iteration,
                   final poses, every 997th final point.
  venice_forcing : the same problem with the inexact-Newton linear solve (block-Jacobi PCG, Ceres' forcing sequence eta = 0.1):
                   per-iteration chi2 and PCG iteration counts; judged on final chi2 only.
  final_pcg      : Final-13682 shape, PCG to 1e-12 (a dense S would be 54 GB), first 6 LM iterations.
"""
import importlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
ba = importlib.import_module("3dreconstruction_b200")
orc = ba.Library(os.path.join(ROOT, "oracle", "libba_oracle.so"))

PT_STRIDE = 997


def run(shape, **opts):
    P = ba.synth(shape)
    t0 = time.time()
    with orc.create(orc.default_options(intrinsics_free=0, robust_kind=1, robust_a=4.0, num_threads=os.cpu_count() or 1, **opts)) as h:
        h.set_problem(P)
        s = h.solve()
        t = h.trajectory()
        pose, intr, pt = h.get_state()
    print(f"{shape} {opts}: {len(t) - 1} LM iterations in {time.time() - t0:.0f} s, cost {s.initial_cost:.9e} -> {s.final_cost:.9e}, "
          f"{s.total_pcg_iterations} PCG iterations, {s.message.decode()}", flush=True)
    for r in t:
        print(f"  {r.iteration:3d} ok {r.step_is_successful} cost {r.cost:.12e} cand {r.candidate_cost:.12e} mcc {r.model_cost_change:.6e} "
              f"rho {r.relative_decrease:.4f} radius {r.radius:.3e} pcg {r.pcg_iterations}", flush=True)
    return dict(sizes=np.array(P.sizes, np.int64), cost=np.array([r.cost for r in t]), candidate_cost=np.array([r.candidate_cost for r in t]),
                model_cost_change=np.array([r.model_cost_change for r in t]), relative_decrease=np.array([r.relative_decrease for r in t]),
                accepted=np.array([r.step_is_successful for r in t], np.int8), valid=np.array([r.step_is_valid for r in t], np.int8),
                radius=np.array([r.radius for r in t]), step_norm=np.array([r.step_norm for r in t]),
                pcg_iterations=np.array([r.pcg_iterations for r in t], np.int32), termination=np.int32(s.termination),
                final_pose=pose, final_pt_sample=pt[::PT_STRIDE].copy(), pt_stride=np.int32(PT_STRIDE),
                obs_checksum=np.int64(P.obs_cam.astype(np.int64).sum() * 31 + P.obs_pt.astype(np.int64).sum()))


JOBS = {
    "venice_exact": lambda: run("venice", linear_solver=0, max_num_iterations=50),
    "venice_forcing": lambda: run("venice", linear_solver=1, pcg_forcing=1, eta=0.1, pcg_max_iterations=500, max_num_iterations=50),
    "final_pcg": lambda: run("final", linear_solver=1, pcg_rel_tolerance=1e-12, pcg_max_iterations=2000, max_num_iterations=6),
}

if __name__ == "__main__":
    for job in sys.argv[1:]:
        out = JOBS[job]()
        np.savez_compressed(os.path.join(HERE, f"{job}_oracle_trajectory_large.npz"), **out)
        print("written", job, flush=True)
