O=gpurun_out/r4f; mkdir -p $O
timeout 300 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
python -c "
import json; d=json.load(open('$O/bench_6M.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['parity']['pass'], {k:round(v['frac'],3) for k,v in d['stages'].items()})"
