# under gpurun --gpus 8: config 4 (GOP, a scene cut every 15 frames = 8 GOPs) and config 5 (4K tile-row split,
# sharded update and the NCCL all-reduce beside it) with the round's final kernels
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29542 bench.py --gpus 8 --workload gop --gop-cut 15 2>/dev/null | tail -1 > gpurun_out/gop_n8_cut15_v2.json
for ex in sharded nccl; do timeout 300 $TR --master-port 29543 tests/run_spatial_multi_gpu.py --exchange $ex 2>/dev/null | tail -1 > gpurun_out/spatial_n8_${ex}_v2.json; done
cut -c1-300 gpurun_out/gop_n8_cut15_v2.json; cut -c1-420 gpurun_out/spatial_n8_*_v2.json
