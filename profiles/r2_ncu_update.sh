#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_fit_update --launch-skip 30 --launch-count 2 -f \
    -o $O/r2_update python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > $O/r2_ncu_update.log 2>&1
tail -3 $O/r2_ncu_update.log
