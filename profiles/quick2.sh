#!/bin/bash
# usage: profiles/quick2.sh [label] [bench args...]  -- headline numbers of one default bench run
lab=$1; shift
timeout 300 python bench.py --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lab', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'seq', round(d['e2e']['sequential']['value'],1), 'one_slice', round(d['one_slice']['value'],1), 'lanes', d['config'].get('concurrent_frames'))"
