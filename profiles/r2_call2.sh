#!/bin/bash
# round 2, call 2: GPU tests of the new operator path, drop-in benches before/after, host-time breakdown,
# ncu source-level capture of the tile kernel
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_c2_pytest.log 2>&1; echo "pytest rc $?" >> $O/r2_c2_pytest.log; tail -15 $O/r2_c2_pytest.log
timeout 300 python profiles/dropin_host_profile.py > $O/r2_c2_host_ranked.txt 2>&1
timeout 300 python profiles/dropin_host_profile.py --binning onesweep --sync --foreach-adan > $O/r2_c2_host_r1like.txt 2>&1
head -3 $O/r2_c2_host_ranked.txt $O/r2_c2_host_r1like.txt
timeout 300 python bench.py --impl b200-dropin --no-cpu --steps 200 > $O/r2_c2_bench_dropin.json 2>$O/r2_c2_err.log; tail -c 600 $O/r2_c2_bench_dropin.json
timeout 300 python bench.py --no-cpu --steps 3 --warmup 3 > $O/r2_c2_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fit_tiles -s 10 -c 2 -o $O/r2_fit_tiles python bench.py --no-cpu --steps 3 --warmup 3 > $O/r2_c2_ncu.log 2>&1
ls -la $O | tail -12
