#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2 | cut -c1-200 | tee $O/r2_qf_pytest.txt
for f in "" "--trained-like"; do
  timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_qf_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_qf.txt
done
timeout 300 python bench.py --impl b200-dropin --no-cpu --steps 100 2>>$O/r2_qf_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('dropin', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_qf.txt
