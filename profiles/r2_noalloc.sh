#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r2_noalloc_alone.log 2>&1; echo "alone rc $?" | tee -a $O/r2_noalloc_alone.log; tail -3 $O/r2_noalloc_alone.log | cut -c1-200
timeout 1500 python -m pytest tests -m gpu -q -x > $O/r2_noalloc_all.log 2>&1; echo "all rc $?" | tee -a $O/r2_noalloc_all.log; tail -4 $O/r2_noalloc_all.log | cut -c1-200
