#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
for env in "" "GS2D_NO_SMALL=1"; do for f in "" "--trained-like"; do
  env $env timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_c8_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${env:-default} $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_c8_ab.txt
done; done
bash profiles/r2_sanitize.sh memcheck
