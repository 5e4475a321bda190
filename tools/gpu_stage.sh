#!/bin/bash
# Runs on the B200 box under gpurun: one named stage of the round's measurement plan; everything it keeps goes to gpurun_out/<tag>_*.
#   tools/gpu_stage.sh check <tag>     smoke of the streaming kernel (own timeout), the GPU test suite, the default bench line
#   tools/gpu_stage.sh ab <tag> <shape> <variant>...   A/B of engine builds / environment knobs (tools/ab.sh), CUDA-event kernel tables
#   tools/gpu_stage.sh profile <tag>   ncu launch list + full captures (tools/gpu_profile.sh)
# Every step runs under its own timeout, so a kernel that hangs costs that step, not the GPU call.
set -u
STAGE=${1:-check}; TAG=${2:-r2}; shift 2 || true
OUT=gpurun_out; mkdir -p $OUT
case $STAGE in
check)
  timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; rc=$?
  echo "smoke rc=$rc"; tail -6 $OUT/${TAG}_smoke.log
  if [ $rc -ne 0 ]; then echo "streaming kernel failed its smoke: the rest of this stage runs the CTA-per-tile kernel (BA_SCATTER_MODE=4)"; export BA_SCATTER_MODE=4; fi
  timeout -s KILL 1500 python -m pytest tests -m gpu -q --maxfail=12 -p no:cacheprovider > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/${TAG}_pytest.log
  timeout -s KILL 600 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_venice_n1.json 2> $OUT/${TAG}_bench_venice_n1.err; echo "bench rc=$?"; tail -3 $OUT/${TAG}_bench_venice_n1.err; cat $OUT/${TAG}_bench_venice_n1.json
  ;;
ab)
  SHAPE=$1; shift
  tools/ab.sh $SHAPE "$@" 2>&1 | tee $OUT/${TAG}_ab_$SHAPE.txt
  ;;
profile)
  tools/gpu_profile.sh $TAG
  ;;
esac
