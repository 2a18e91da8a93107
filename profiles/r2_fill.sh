#!/bin/bash
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
for v in "A=0" "GS2D_PIPE_STREAMS=1" "GS2D_STREAM_FIRST_ALONE=1" "GS2D_PIPE_STREAMS=1 GS2D_STREAM_FIRST_ALONE=1"; do
  echo "== $v"; env $v timeout 600 python profiles/e2e_fill_probe.py 2>&1 | grep -v Warning
done | tee $O/r2_e2e_fill_probe.txt
