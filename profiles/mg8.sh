TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 bench.py --gpus 8 --steps 200 --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_n8.json
timeout 600 $TR --master-port 29542 bench.py --gpus 8 --workload gop --gop-cut 15 2>/dev/null | tail -1 > gpurun_out/gop_n8_cut15.json
for ex in nccl peer nvls; do timeout 300 $TR --master-port 29543 tests/run_spatial_multi_gpu.py --exchange $ex 2>/dev/null | tail -1 > gpurun_out/spatial_n8_$ex.json; done
cut -c1-200 gpurun_out/bench_n8.json; cut -c1-200 gpurun_out/gop_n8_cut15.json; cut -c1-420 gpurun_out/spatial_n8_*.json
