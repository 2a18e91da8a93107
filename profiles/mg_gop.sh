# usage: bash profiles/mg_gop.sh N [CUT]   (under gpurun --gpus N): config 4 on N GPUs, GOPs sharded over the ranks,
# each rank fitting up to 4 of its GOPs concurrently
N=$1; CUT=${2:-30}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29542 bench.py --gpus $N --workload gop --gop-cut $CUT 2>/dev/null | tail -1 > gpurun_out/gop_n${N}_cut${CUT}_v3.json
python -c "
import json; d=json.load(open('gpurun_out/gop_n${N}_cut${CUT}_v3.json')); print($N, $CUT, round(d['value'],2), round(d['fit_iters_per_s']), round(d['psnr_mean'],3), d['config']['gops_per_rank'])"
