#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 60 tools/probe/bw_probe > gpurun_out/c7_bw_probe.txt 2>&1; cat gpurun_out/c7_bw_probe.txt
( time bash tools/ab.sh venice base k2c2 k2c2m4 rowst1 rowst2 rowld1 rowld2 ) > gpurun_out/c7_ab.txt 2>&1
grep -E "variant|plain|rc=|K2c|K5_|PCG_|rror" gpurun_out/c7_ab.txt
