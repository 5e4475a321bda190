#!/bin/bash
# Round-end refresh after the host-side and K2c changes: tests, smoke, bench line, per-kernel stats, ncu launch list + K2c capture.
TAG=${1:-r1g}
mkdir -p gpurun_out
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/${TAG}_pytest.txt 2>&1; tail -n 6 gpurun_out/${TAG}_pytest.txt | head -3
timeout -s KILL 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.txt 2>&1; tail -n 1 gpurun_out/${TAG}_smoke.txt
( time timeout -s KILL 400 python bench.py ) > gpurun_out/${TAG}_bench_n1.txt 2>&1; grep -E "^\{" gpurun_out/${TAG}_bench_n1.txt | tail -1 > gpurun_out/${TAG}_bench_venice_n1.json; cut -c1-700 gpurun_out/${TAG}_bench_venice_n1.json
timeout -s KILL 200 python tools/quick_gpu.py venice > gpurun_out/${TAG}_kernel_stats_venice.txt 2>&1; grep -E "plain|K[0-9]|PCG" gpurun_out/${TAG}_kernel_stats_venice.txt
KERNELS="k2c_camera_blocks" timeout -s KILL 300 bash tools/gpu_profile.sh $TAG > gpurun_out/${TAG}_profile.log 2>&1; tail -n 4 gpurun_out/${TAG}_profile.log
