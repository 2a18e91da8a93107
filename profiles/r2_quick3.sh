#!/bin/bash
# GPU tests, the C2 bench (+ trained-like), and the drop-in leg with per-stage times
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > $O/r2_q_pytest.log 2>&1; echo "pytest rc $?" | tee -a $O/r2_q_pytest.log; tail -3 $O/r2_q_pytest.log | cut -c1-200
for f in "" "--trained-like"; do
  timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 $f 2>>$O/r2_q_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'rq', round(d['render_quantized']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_q_ab.txt
done
timeout 300 python bench.py --impl b200-dropin --no-cpu --steps 100 2>>$O/r2_q_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('dropin', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_q_ab.txt
