#!/bin/bash
# like ab.sh but only the headline config; extra bench args after --
for v in "$@"; do
  if [ "$v" = base ]; then unset GS2D_LIB; else export GS2D_LIB=$PWD/gsvc_b200/lib/variants/libgs2d_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 200 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$v', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})"
done
