#!/bin/bash
mkdir -p gpurun_out
( time bash tools/ab.sh venice base base@BA_L2_PERSIST=1 base@BA_L2_PERSIST=64 base@BA_L2_PERSIST=32 ) > gpurun_out/c6_ab.txt 2>&1
grep -E "variant|plain|rc=|K4_|K5_|PCG_|warm|rror" gpurun_out/c6_ab.txt
