#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 600 python profiles/debug/e2e_sweep.py > $O/r2_e2e_sweep.txt 2>&1; cat $O/r2_e2e_sweep.txt | cut -c1-160
