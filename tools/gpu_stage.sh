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
mg)   # tools/gpu_stage.sh mg <tag> <N> [workload ...]: multi-GPU parity check, per-rank kernel table and bench lines at N GPUs (gpurun --gpus N)
  N=$1; shift
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
  timeout -s KILL 600 $TR --master-port 29511 tools/multi_gpu_check.py ladybug trafalgar > $OUT/${TAG}_multi_gpu_check_n$N.txt 2>&1; echo "multi_gpu_check rc=$?"; grep -E "OK|MISMATCH|rror" $OUT/${TAG}_multi_gpu_check_n$N.txt | tail -12
  for W in "${@:-venice}"; do
    timeout -s KILL 400 $TR --master-port 29513 tools/multi_gpu_stats.py $W 3 0 > $OUT/${TAG}_multi_gpu_stats_${W}_n$N.txt 2>&1; echo "stats $W rc=$?"; grep -vE "^W|^\*|warn" $OUT/${TAG}_multi_gpu_stats_${W}_n$N.txt | tail -40
    ST=3; [ "$W" = "final" ] && ST=2
    timeout -s KILL 900 $TR --master-port 29515 bench.py --gpus $N --steps $ST --warmup 3 --workload $W > $OUT/${TAG}_bench_${W}_n$N.json 2> $OUT/${TAG}_bench_${W}_n$N.err; echo "bench $W rc=$?"; tail -2 $OUT/${TAG}_bench_${W}_n$N.err; cat $OUT/${TAG}_bench_${W}_n$N.json
  done
  ;;
profile)
  tools/gpu_profile.sh $TAG
  ;;
esac
