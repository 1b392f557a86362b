#!/bin/bash
# usage: profiles/sweep.sh  -- bench matrix over env knobs (experiments only)
for fr in "0.0625" "0.04" "0.1" "0.02,0.25"; do
for sp in 1 2; do
GSPLAT_SLICE_FRACTIONS=$fr GSPLAT_K5_SPLIT=$sp timeout 120 python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fr=$fr split=$sp', round(d['value'],1), round(d['e2e']['value'],1), d['frame']['duplicates'], d['frame']['depth_sorted'], {k[:3]:round(v['ms'],3) for k,v in d['stages'].items()})"
done; done
