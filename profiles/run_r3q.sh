O=gpurun_out/r3q; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
timeout 500 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
python -c "
import json; d=json.load(open('$O/bench_6M.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic_source'], d['parity']['pass'])"
