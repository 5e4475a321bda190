#!/bin/bash
mkdir -p gpurun_out
( time BA_FUSE=1 timeout -s KILL 150 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/c3_smoke_fuse.txt 2>&1; echo "smoke(fuse) rc=$?"; tail -3 gpurun_out/c3_smoke_fuse.txt
( time bash tools/ab.sh venice base base@BA_FUSE=1 noalloc noalloc@BA_FUSE=1 minb7@BA_FUSE=1 base@BA_FUSE=1,BA_PDL=0 ) > gpurun_out/c3_ab.txt 2>&1
grep -E "variant|plain|rc=|K5_|PCG_" gpurun_out/c3_ab.txt
( time BA_FUSE=1 timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/c3_pytest_fuse.txt 2>&1
tail -4 gpurun_out/c3_pytest_fuse.txt
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/c3_pytest.txt 2>&1
tail -4 gpurun_out/c3_pytest.txt
( AB_LM=50 AB_TIMEOUT=120 bash tools/ab.sh venice base base@BA_FUSE=1 ) > gpurun_out/c3_full.txt 2>&1
grep -E "variant|plain|warm|rc=" gpurun_out/c3_full.txt
