#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_fit --launch-skip 60 --launch-count 4 -f \
    -o $O/r2_h_fit python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > $O/r2_h_ncu.log 2>&1
tail -2 $O/r2_h_ncu.log
