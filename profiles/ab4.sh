#!/bin/bash
# A/B of tile-kernel variants (profiles/build_variant.py): headline and trained-like
for lib in ${GS2D_AB_LIBS:-"" gsvc_b200/lib/variants/libgs2d_scalar.so gsvc_b200/lib/variants/libgs2d_x2r4.so}; do for f in "" "--trained-like"; do
  GS2D_LIB=$lib timeout 300 python bench.py --no-cpu --steps 200 $f 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('${lib:-default} $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})"
done; done
