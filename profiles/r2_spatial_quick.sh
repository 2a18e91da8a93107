#!/bin/bash
# N GPUs: the sharded C5 step (PDL on), 100 iterations, twice
N=${1:-4}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
for rep in 1 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$rep \
    tests/run_spatial_multi_gpu.py --exchange sharded --iters 100 > $O/r2_spatial_quick_n${N}_$rep.txt 2>$O/r2_spatial_quick_err.log
  grep -E "^\{" $O/r2_spatial_quick_n${N}_$rep.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('world','spatial_ms_per_iter','single_gpu_ms_per_iter','speedup','loss_spatial','loss_single','params_identical_across_ranks') if k in d})"
done
tail -2 $O/r2_spatial_quick_err.log
