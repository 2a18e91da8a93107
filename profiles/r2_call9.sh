#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
V=gsvc_b200/lib/variants
GS2D_LIB=$V/libgs2d_checked.so timeout 1200 python -m pytest tests -m gpu -q -x > $O/r2_c9_checked_pytest.log 2>&1; echo "checked-build pytest rc $?" | tee -a $O/r2_c9_checked_pytest.log; tail -4 $O/r2_c9_checked_pytest.log
GS2D_LIB=$V/libgs2d_checked.so timeout 600 python profiles/sanitizer_case.py > $O/r2_c9_checked_cases.log 2>&1; echo "checked-build cases rc $?" | tee -a $O/r2_c9_checked_cases.log; grep -c "GS2D_ASSERT" $O/r2_c9_checked_cases.log
timeout 600 python -m pytest tests/test_gpu_fit.py -m gpu -q -x -k "adan_arithmetic" > $O/r2_c9_adan.log 2>&1; tail -5 $O/r2_c9_adan.log
for lib in "" $V/libgs2d_small_noperm.so $V/libgs2d_small14.so $V/libgs2d_small12.so $V/libgs2d_small20.so; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --no-extra-legs --steps 200 2>>$O/r2_c9_err.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default}', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'psnr', round(d['final_psnr'],4), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})" 2>&1 | tee -a $O/r2_c9_ab.txt
done
