O=gpurun_out/r3b; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 > $O/noalloc_alone.log; cat $O/noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > $O/tests.log; cat $O/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 500 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
python -c "
import json; d=json.load(open('$O/bench_6M.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['parity']['pass'], {k:round(v['ms'],4) for k,v in d['stages'].items()})"
