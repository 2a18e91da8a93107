#!/bin/bash
# ncu --set full of the fit kernels with trained-like Gaussians (what a fit looks like after its first few hundred iterations)
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
python bench.py --trained-like --no-cpu --no-extra-legs --steps 40 --warmup 10 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_fit --launch-skip 60 --launch-count 2 -f \
    -o $O/r2_trained python bench.py --trained-like --no-cpu --no-extra-legs --steps 40 --warmup 10 > $O/r2_ncu_trained.log 2>&1
tail -2 $O/r2_ncu_trained.log
