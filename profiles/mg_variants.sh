# usage: bash profiles/mg_variants.sh N  -- sharded spatial mode, in-tree build beside deliberately broken variants (what costs what)
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for v in "" ${VARIANTS}; do
  if [ -n "$v" ]; then export GS2D_LIB=gsvc_b200/lib/variants/libgs2d_$v.so; fi
  timeout 200 $TR --master-port 29545 tests/run_spatial_multi_gpu.py --exchange sharded --no-check --kernel-table 30 > gpurun_out/var_$v.log 2>&1
  echo "variant=[$v] rc=$?"; grep '^{' gpurun_out/var_$v.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(round(d['spatial_ms_per_iter']*1e3,1),'us/iter')"; grep KERNEL gpurun_out/var_$v.log | head -4 | cut -c1-60
done
