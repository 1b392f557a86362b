#!/bin/bash
# usage: profiles/quick.sh [label]  -- one short bench line (experiments only)
timeout 200 python bench.py --steps 150 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['value'],1), round(d['e2e']['value'],1), d['frame']['pairs'], d['frame']['blended'], {k[:3]:round(v['ms'],3) for k,v in d['stages'].items()})"
