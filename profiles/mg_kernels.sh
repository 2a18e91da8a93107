# usage: bash profiles/mg_kernels.sh N EXCHANGE  -- per-kernel device times of the spatial mode on N GPUs (torch.profiler, rank 0)
N=$1; EX=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29544 tests/run_spatial_multi_gpu.py --exchange $EX --kernel-table 30 > gpurun_out/spatial_n${N}_${EX}_kernels.log 2>&1
echo "rc=$?"; grep KERNEL gpurun_out/spatial_n${N}_${EX}_kernels.log | cut -c1-75
