#!/bin/bash
# GPU call: smoke (engine vs oracle), A/B of the tuning variants on the Venice shape, then the GPU test-suite on the default build.
mkdir -p gpurun_out
( time timeout -s KILL 150 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/c2_smoke.txt 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/c2_smoke.txt
( time BA_PDL=0 timeout -s KILL 150 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/c2_smoke_nopdl.txt 2>&1; echo "smoke(nopdl) rc=$?"; tail -3 gpurun_out/c2_smoke_nopdl.txt
( time bash tools/ab.sh venice nopdl base old nojpack minb7 minb8 obshinv xsmem xsobs k1cam k1c5 ) > gpurun_out/c2_ab.txt 2>&1
grep -E "variant|plain|rc=" gpurun_out/c2_ab.txt
( time timeout -s KILL 420 python -m pytest tests -m gpu -x -q --timeout 120 ) > gpurun_out/c2_pytest.txt 2>&1
tail -5 gpurun_out/c2_pytest.txt
