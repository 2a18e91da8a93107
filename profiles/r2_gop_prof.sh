#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 600 python profiles/gop_host_profile.py 2>&1 | grep -v Warning | cut -c1-200 | tee $O/r2_gop_host_profile.txt
