#!/bin/bash
# multi-GPU: the driver's scaling command at N ranks (bench line with the spatial and gop legs)
N=${1:-2}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
   bench.py --gpus $N --steps 20 --warmup 5 > $O/r2_mg${N}_bench.json 2>$O/r2_mg${N}_err.log ) 2>&1 | tail -3
tail -c 800 $O/r2_mg${N}_bench.json; tail -5 $O/r2_mg${N}_err.log
