#!/bin/bash
# GPU call: A/B of the tuning variants on the Venice shape, then the GPU test-suite on the default build.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/c1_smi.txt
( time bash tools/ab.sh venice base old nopdl minb7 minb8 obshinv xsmem xsobs k1c5 k1c3 ) > gpurun_out/c1_ab.txt 2>&1
tail -5 gpurun_out/c1_ab.txt
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c1_pytest.txt 2>&1
tail -5 gpurun_out/c1_pytest.txt
