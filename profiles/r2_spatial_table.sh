#!/bin/bash
N=${1:-4}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
GS2D_NO_PDL=${NOPDL:-0} timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  tests/run_spatial_multi_gpu.py --exchange sharded --iters 50 --phase-times 20 --kernel-table 20 > $O/r2_spatial_n${N}_table.txt 2>$O/r2_spatial_n${N}_err.log
grep -E "^\{|KERNEL" $O/r2_spatial_n${N}_table.txt | cut -c1-230 | head -60; tail -3 $O/r2_spatial_n${N}_err.log
