#!/bin/bash
# A/B: fully unrolled per-warp id ranking (-DGS2D_RANK_UNROLLED=1) against the tree's build
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
V=gsvc_b200/lib/variants
GS2D_LIB=$V/libgs2d_rank.so timeout 900 python -m pytest tests/test_gpu_fit.py tests/test_gpu_raster.py tests/test_gpu_edges.py -m gpu -q -x 2>&1 | tail -2 | cut -c1-200 | tee $O/r2_ab_rank_pytest.txt
: > $O/r2_ab_rank.txt
for rep in 1 2; do
for lib in "" $V/libgs2d_rank.so; do
  for f in "" "--trained-like"; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_ab_rank_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default} $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_ab_rank.txt
  done
done
done
