#!/bin/bash
# the whole GPU test-suite and the sanitizer cases on the checked build (-DGS2D_CHECKS=1: GS2D_ASSERT compiled in)
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
GS2D_LIB=gsvc_b200/lib/variants/libgs2d_checked.so timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4 | cut -c1-200 | tee $O/r2_checked_build_pytest_tail.log
GS2D_LIB=gsvc_b200/lib/variants/libgs2d_checked.so timeout 600 python profiles/sanitizer_case.py 2>&1 | grep -v Warning | tail -8 | cut -c1-200 | tee $O/r2_checked_build_cases.log
