#!/bin/bash
# Round-end single-GPU evidence: tests, smoke, bench line, per-kernel stats, Final-13682 on one GPU, ncu launch list + full captures.
TAG=${1:-r1f}
mkdir -p gpurun_out
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/${TAG}_pytest.txt 2>&1; tail -n 4 gpurun_out/${TAG}_pytest.txt
timeout -s KILL 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.txt 2>&1; tail -n 1 gpurun_out/${TAG}_smoke.txt
( time timeout -s KILL 400 python bench.py ) > gpurun_out/${TAG}_bench_n1.txt 2>&1; grep -E "^\{" gpurun_out/${TAG}_bench_n1.txt | tail -1 > gpurun_out/${TAG}_bench_venice_n1.json; cat gpurun_out/${TAG}_bench_venice_n1.json
timeout -s KILL 200 python tools/quick_gpu.py venice > gpurun_out/${TAG}_kernel_stats_venice.txt 2>&1; grep -E "plain|K[0-9]|PCG" gpurun_out/${TAG}_kernel_stats_venice.txt
timeout -s KILL 300 python tools/quick_gpu.py final max_num_iterations=10 > gpurun_out/${TAG}_kernel_stats_final13682.txt 2>&1; grep -E "plain|K[0-9]|PCG|warm" gpurun_out/${TAG}_kernel_stats_final13682.txt
KERNELS="k_schur_rows k_camera_gather k_pcg_tail k1_residual_jacobian k2c_camera_blocks k_schur_sweep" timeout -s KILL 900 bash tools/gpu_profile.sh $TAG > gpurun_out/${TAG}_profile.log 2>&1; tail -n 12 gpurun_out/${TAG}_profile.log
