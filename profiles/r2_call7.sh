#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > $O/r2_c7_pytest.log 2>&1; echo "pytest rc $?" >> $O/r2_c7_pytest.log; tail -8 $O/r2_c7_pytest.log
for env in "" "GS2D_NO_SMALL=1"; do for f in "" ; do
  env $env timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_c7_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${env:-default} $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_c7_ab.txt
done; done
env timeout 300 python bench.py --no-cpu --no-extra-legs --steps 20 --warmup 5 2>>$O/r2_c7_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('20 steps', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'])" | tee -a $O/r2_c7_ab.txt
tail -3 $O/r2_c7_err.log
