#!/bin/bash
# usage: profiles/sweep.sh  -- bench matrix over env knobs (experiments only)
for fr in "0.08" "0.05,0.2" "0.08,0.3" "0.04,0.12,0.35" "0.08,0.2,0.45"; do
GSPLAT_SLICE_FRACTIONS=$fr timeout 120 python bench.py --steps 150 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fr=$fr', round(d['value'],1), round(d['e2e']['value'],1), d['frame']['duplicates'], d['frame']['depth_sorted'], {k[:3]:round(v['ms'],3) for k,v in d['stages'].items()})"
done
