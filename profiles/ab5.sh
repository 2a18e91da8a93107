#!/bin/bash
# round 2, A/B #1: tile-kernel fixed cost (per-warp staging, smaller batch buffer, resident CTAs, one warp per tile)
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_ab5_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/r2_ab5_pytest.log
tail -3 gpurun_out/r2_ab5_pytest.log
V=gsvc_b200/lib/variants
for lib in "" $V/libgs2d_nowarp.so $V/libgs2d_sq32.so $V/libgs2d_persist.so $V/libgs2d_persist_sq32.so $V/libgs2d_w1.so $V/libgs2d_w1_p.so; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --steps 200 2>gpurun_out/r2_ab5_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default}', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a gpurun_out/r2_ab5.txt
done
for lib in "" $V/libgs2d_nowarp.so; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --steps 200 --trained-like 2>>gpurun_out/r2_ab5_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default} trained-like', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a gpurun_out/r2_ab5.txt
done
