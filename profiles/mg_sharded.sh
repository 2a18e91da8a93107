# usage: bash profiles/mg_sharded.sh N   (under gpurun --gpus N): spatial 4K mode, sharded update beside the replicated exchanges
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for ex in ${EXCHANGES:-sharded peer nccl}; do
  timeout 240 $TR --master-port 29543 tests/run_spatial_multi_gpu.py --exchange $ex --phase-times 30 > gpurun_out/spatial_n${N}_$ex.log 2>&1
  echo "rc=$?"; tail -1 gpurun_out/spatial_n${N}_$ex.log | cut -c1-1200 | tee gpurun_out/spatial_n${N}_$ex.json
done
