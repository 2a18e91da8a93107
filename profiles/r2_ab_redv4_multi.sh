#!/bin/bash
# N GPUs: C5 spatial step, pushed reduction with vector (the tree's build) and with scalar (-DGS2D_RED_V4=0) reductions
N=${1:-4}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
for lib in "" gsvc_b200/lib/variants/libgs2d_redv0.so; do
  tag=v4; [ -n "$lib" ] && tag=scalar
  GS2D_LIB=$lib GS2D_NO_PDL=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    tests/run_spatial_multi_gpu.py --exchange sharded --iters 50 --phase-times 20 --kernel-table 20 > $O/r2_redv4_n${N}_${tag}.txt 2>$O/r2_redv4_n${N}_${tag}_err.log
  echo "== $tag (no PDL, kernel table)"; grep -E "^\{|KERNEL rank 0" $O/r2_redv4_n${N}_${tag}.txt | cut -c1-230 | head -8
  GS2D_LIB=$lib timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
    tests/run_spatial_multi_gpu.py --exchange sharded --iters 100 > $O/r2_redv4_n${N}_${tag}_pdl.txt 2>>$O/r2_redv4_n${N}_${tag}_err.log
  echo "== $tag (PDL)"; grep -E "^\{" $O/r2_redv4_n${N}_${tag}_pdl.txt | cut -c1-330
done
