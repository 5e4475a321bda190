#!/bin/bash
mkdir -p gpurun_out
( time timeout -s KILL 150 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/c4_smoke.txt 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/c4_smoke.txt
( time bash tools/ab.sh venice base g16u4 g16u6 g10u8 fuse2 base@BA_PDL=0 ) > gpurun_out/c4_ab.txt 2>&1
grep -E "variant|plain|rc=|K5_|PCG_|warm" gpurun_out/c4_ab.txt
