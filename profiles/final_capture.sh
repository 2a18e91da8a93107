#!/bin/bash
# One-GPU evidence run (under gpurun): tests, smoke, the bench line, then -- each only after its own command
# exited 0 without ncu -- the ncu launch list of the same bench command and one `--set full` capture of the
# two fit kernels.  Outputs land in gpurun_out/ (copied to profiles/ by hand).
T=${1:-r1_g}
python -m pytest tests -m gpu -q 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || exit 1
python bench.py --trained-like --no-cpu > gpurun_out/${T}_bench_trained_like.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/${T}_ncu_list.log 2>&1
python bench.py --no-cpu --steps 40 --warmup 10 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_fit --launch-skip 60 --launch-count 4 -f \
    -o gpurun_out/${T}_fit python bench.py --no-cpu --steps 40 --warmup 10 > gpurun_out/${T}_ncu_full.log 2>&1
cut -c1-400 gpurun_out/${T}_bench.json
