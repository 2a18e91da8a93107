#!/bin/bash
# A/B of kernel variants on the GPU box: bash profiles/ab.sh base mb12 r4 ...   (base = the in-tree build)
for v in "$@"; do
  if [ "$v" = base ]; then unset GS2D_LIB; else export GS2D_LIB=$PWD/gsvc_b200/lib/variants/libgs2d_$v.so; fi
  for f in "" "--trained-like"; do
    timeout 300 python bench.py --no-cpu --steps 200 $f 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$v', '$f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})"
  done
done
