#!/bin/bash
# usage: r2_sanitize.sh memcheck|racecheck|synccheck      (one tool per gpurun call: B200_PROFILING.md)
tool=$1
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 300 python profiles/sanitizer_case.py > $O/r2_san_plain.log 2>&1 || { echo "plain run failed"; tail -20 $O/r2_san_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $tool --print-limit 30 python profiles/sanitizer_case.py > $O/r2_san_$tool.log 2>&1
echo "rc $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok$|ok," $O/r2_san_$tool.log | tail -12; grep -c "========= " $O/r2_san_$tool.log
