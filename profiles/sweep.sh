#!/bin/bash
# usage: profiles/sweep.sh  -- bench matrix over env knobs (experiments only)
for fr in "0.0625" "0.03" "0.125"; do
for sp in 2 4; do
GSPLAT_SLICE_FRACTIONS=$fr GSPLAT_K5_SPLIT=$sp timeout 120 python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fr=$fr split=$sp', round(d['value'],1), round(d['e2e']['value'],1), d['frame']['duplicates'], {k[:3]:round(v['ms'],3) for k,v in d['stages'].items()})"
done; done
