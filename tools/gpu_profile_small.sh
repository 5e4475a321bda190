#!/bin/bash
# Runs on the B200 box under gpurun: launch lists + full ncu captures of the direct-solve kernels on the small shapes
# (the reference's own regime).  usage: tools/gpu_profile_small.sh <tag>  -> gpurun_out/<tag>_small_*.{csv,ncu-rep}
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
for SHAPE in sequential30 trafalgar; do
  python tools/small_regime.py $SHAPE > $OUT/${TAG}_small_${SHAPE}_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_small_${SHAPE}_plain.log; exit 1; }
  tail -1 $OUT/${TAG}_small_${SHAPE}_plain.log | cut -c1-400
  ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/${TAG}_small_${SHAPE}_launches.csv python tools/small_regime.py $SHAPE > $OUT/${TAG}_small_ncu.log 2>&1
  echo "launch list $SHAPE rc=$?"
done
for K in k_dense_chol_df k_dense_pairs; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -f -o $OUT/${TAG}_small_$K python tools/small_regime.py trafalgar > $OUT/${TAG}_small_ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done
