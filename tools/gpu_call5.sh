#!/bin/bash
mkdir -p gpurun_out
( time bash tools/ab.sh venice base row6 row6@BA_PDL=0 ) > gpurun_out/c5_ab.txt 2>&1
grep -E "variant|plain|rc=|K4_|K5_|PCG_|warm" gpurun_out/c5_ab.txt
