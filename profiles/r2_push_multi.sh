#!/bin/bash
# N GPUs: per-kernel table of the sharded spatial step with the pushed reduction and with the pulled one, then the bench line
N=${1:-4}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
for mode in push pull; do
  nopush=0; [ $mode = pull ] && nopush=1
  GS2D_NO_PUSH=$nopush GS2D_NO_PDL=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    tests/run_spatial_multi_gpu.py --exchange sharded --iters 50 --phase-times 20 --kernel-table 20 > $O/r2_spatial_n${N}_${mode}_table.txt 2>$O/r2_spatial_n${N}_${mode}_err.log
  echo "== $mode"; grep -E "^\{|KERNEL" $O/r2_spatial_n${N}_${mode}_table.txt | cut -c1-260 | head -40; tail -3 $O/r2_spatial_n${N}_${mode}_err.log
done
bash profiles/r2_mg2.sh $N
