O=gpurun_out/r3w; mkdir -p $O
timeout 500 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
timeout 300 python bench.py --workload synthetic-1M-1080p > $O/bench_cfg2_1M.json 2>> $O/bench.err
timeout 400 python bench.py --workload synthetic-6M-4k --steps 100 > $O/bench_cfg3_4k_views.json 2>> $O/bench.err
python -c "
import json
for f in ['bench_6M','bench_cfg2_1M','bench_cfg3_4k_views']:
    d=json.load(open('$O/'+f+'.json')); print(f, d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['ms'], d['parity']['pass'])"
