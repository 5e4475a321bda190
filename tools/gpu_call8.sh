#!/bin/bash
mkdir -p gpurun_out
BA_TIMING=1 timeout -s KILL 200 python tools/e2e_phases.py > gpurun_out/c8_phases.txt 2>&1; tail -n 14 gpurun_out/c8_phases.txt
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/c8_pytest.txt 2>&1; tail -n 4 gpurun_out/c8_pytest.txt
