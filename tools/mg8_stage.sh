TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for W in final venice; do for wide in 1 0; do
  BA_P2P_WIDE=$wide timeout -s KILL 300 $TR --master-port 29521 tools/multi_gpu_stats.py $W 3 0 2>&1 | grep -E "rank 0|K5_|K5b|tail" | head -4 | sed "s/^/[$W wide=$wide] /" | tee -a gpurun_out/r2i_pull_ab_n8.txt
done; done
tools/gpu_stage.sh mg r2i 8 venice final
