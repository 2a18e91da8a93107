#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > $O/r2_c5_pytest.log 2>&1; echo "pytest rc $?" >> $O/r2_c5_pytest.log; tail -15 $O/r2_c5_pytest.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2_c5_bench.json 2>$O/r2_c5_bench_err.log ) 2>&1 | tail -3; tail -c 1500 $O/r2_c5_bench.json; tail -5 $O/r2_c5_bench_err.log
