#!/bin/bash
# one GPU: the spatial emulation tests (pushed / pulled reduction), then the C2 bench for a regression check
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fit.py -m gpu -q -x > $O/r2_push_pytest.log 2>&1; echo "pytest rc $?" | tee -a $O/r2_push_pytest.log; tail -5 $O/r2_push_pytest.log | cut -c1-300
timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 2>>$O/r2_push_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee $O/r2_push_ab.txt
