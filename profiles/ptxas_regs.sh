#!/bin/bash
# registers / spills / shared memory of every kernel of fit.cu (or the file given), with extra -D flags
# usage: profiles/ptxas_regs.sh [file.cu] [-DFOO=1 ...]
f=${1:-fit.cu}; shift
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr \
  -Xptxas -v "$@" -c /root/repo/gsvc_b200/csrc/$f -o /dev/null 2>&1 | \
  awk '/Compiling entry function/ {name=$5} /Used/ {print name, $0}' | sed -e "s/ptxas info    : //" | c++filt | \
  sed -E 's/\(.*\)//' 
