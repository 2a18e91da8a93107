#!/bin/bash
# in-tree build, tile slots vs tile ranges, headline and trained-like
for m in "" "--ranges"; do for f in "" "--trained-like"; do
  timeout 300 python bench.py --no-cpu --steps 200 $m $f 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$m $f', 'value', round(d['value']), 'warm', round(d['l2_warm']['value']), 'e2e', round(d['e2e']['value']), {k:round(v['ms_per_step']*1e3,1) for k,v in d['stages'].items()})"
done; done
