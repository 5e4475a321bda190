#!/bin/bash
mkdir -p gpurun_out
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/c9_pytest.txt 2>&1; tail -n 6 gpurun_out/c9_pytest.txt | head -3
( time bash tools/ab.sh venice base ) > gpurun_out/c9_ab.txt 2>&1
grep -E "variant|plain|rc=|K4_|K5_|PCG_|warm|rror" gpurun_out/c9_ab.txt
