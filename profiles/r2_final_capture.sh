#!/bin/bash
# Round-2 evidence run on one GPU (under gpurun): tests, smoke, the driver's bench command, trained-like, then --
# each only after its own command exited 0 without ncu -- the ncu launch list of the bench command and one
# `--set full` capture of the fit kernels and of the operator-path kernels.  Outputs land in gpurun_out/.
T=${1:-r2_g}
mkdir -p $GRAFT_REPO_ROOT/gpurun_out; O=$GRAFT_REPO_ROOT/gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -2 | tee $O/${T}_pytest_tail.txt
python __graft_entry__.py smoke 2>&1 | tail -2 | tee $O/${T}_smoke.txt
python bench.py --gpus 1 --steps 20 --warmup 5 > $O/${T}_bench.json 2> $O/${T}_bench.err || exit 1
python bench.py --no-cpu --no-extra-legs > $O/${T}_bench_200steps.json 2>/dev/null
python bench.py --trained-like --no-cpu --no-extra-legs > $O/${T}_bench_trained_like.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-extra-legs > $O/${T}_ncu_list.log 2>&1
python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_fit --launch-skip 60 --launch-count 4 -f \
    -o $O/${T}_fit python bench.py --no-cpu --no-extra-legs --steps 40 --warmup 10 > $O/${T}_ncu_full.log 2>&1
python bench.py --impl b200-dropin --no-cpu --steps 6 --warmup 3 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none -k regex:"k_fit|k_raster|k_project|k_adan" --launch-skip 40 --launch-count 8 -f \
    -o $O/${T}_dropin python bench.py --impl b200-dropin --no-cpu --steps 6 --warmup 3 > $O/${T}_ncu_dropin.log 2>&1
cut -c1-300 $O/${T}_bench.json
