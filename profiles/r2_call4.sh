#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
python profiles/debug/overflow_dbg.py > $O/r2_c4_dbg.txt 2>&1; cat $O/r2_c4_dbg.txt
timeout 1200 python -m pytest tests -m gpu -q -x > $O/r2_c4_pytest.log 2>&1; echo "pytest rc $?" >> $O/r2_c4_pytest.log; tail -25 $O/r2_c4_pytest.log
