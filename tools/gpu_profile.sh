#!/bin/bash
# Runs on the B200 box under gpurun: launch list + full ncu captures of the LM-iteration kernels (Venice shape, short solve).
# usage: tools/gpu_profile.sh <tag>   -> gpurun_out/<tag>_*.{csv,ncu-rep,log}
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
CMD="python tools/ncu_target.py venice 2 24"
$CMD > $OUT/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -20 $OUT/${TAG}_plain.log; exit 1; }
cat $OUT/${TAG}_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
for K in ${KERNELS:-k_schur_stream k_camera_gather k_pcg_tail k1_residual_jacobian k2c_camera_blocks k_schur_sweep k1c_candidate_cost}; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -f -o $OUT/${TAG}_$K $CMD > $OUT/${TAG}_ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done
ls -la $OUT
