#!/bin/bash
# usage: tools/gpu_call_mg.sh N  — multi-GPU parity check + bench at N GPUs
N=${1:-2}
mkdir -p gpurun_out
( time timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py ladybug trafalgar ) > gpurun_out/mg${N}_check.txt 2>&1
echo "check rc=$?"; grep -E "OK|MISMATCH|rror" gpurun_out/mg${N}_check.txt | head -20
( time timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 1 ) > gpurun_out/mg${N}_bench.txt 2>&1
echo "bench rc=$?"; grep -E "^\{" gpurun_out/mg${N}_bench.txt | tail -1
