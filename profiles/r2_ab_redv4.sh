#!/bin/bash
# A/B: gradient rows added by 16-byte vector reductions (REDG.E.ADD.F32x4).  1 GPU: parity on both builds, bench
# of the tree's build (scalar reductions on one GPU) against -DGS2D_RED_V4=2 (vector reductions everywhere)
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
V=gsvc_b200/lib/variants
timeout 900 python -m pytest tests/test_gpu_fit.py -m gpu -q -x 2>&1 | tail -2 | cut -c1-200 | tee $O/r2_ab_redv4_pytest.txt
GS2D_LIB=$V/libgs2d_redv4.so timeout 900 python -m pytest tests/test_gpu_fit.py tests/test_gpu_raster.py -m gpu -q -x 2>&1 | tail -2 | cut -c1-200 | tee -a $O/r2_ab_redv4_pytest.txt
: > $O/r2_ab_redv4.txt
for rep in 1 2; do
for lib in "" $V/libgs2d_redv4.so; do
  for f in "" "--trained-like"; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_ab_redv4_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default} $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_ab_redv4.txt
  done
done
done
