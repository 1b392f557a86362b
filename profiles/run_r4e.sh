O=gpurun_out/r4e; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
python -c "
import json; d=json.load(open('$O/bench_6M.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic_source'], d['parity']['pass'], d['gpu_launches'])"
