TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 bench.py --gpus 4 --steps 200 --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_n4.json
timeout 600 $TR --master-port 29542 bench.py --gpus 4 --workload gop 2>/dev/null | tail -1 > gpurun_out/gop_n4.json
for ex in nccl peer; do timeout 300 $TR --master-port 29543 tests/run_spatial_multi_gpu.py --exchange $ex 2>/dev/null | tail -1 > gpurun_out/spatial_n4_$ex.json; done
cut -c1-400 gpurun_out/bench_n4.json; cut -c1-300 gpurun_out/gop_n4.json; cat gpurun_out/spatial_n4_*.json
