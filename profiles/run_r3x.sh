O=gpurun_out/r3x; mkdir -p $O
timeout 400 python bench.py --workload synthetic-6M-4k --steps 100 > $O/bench_cfg3_4k_views.json 2>> $O/bench.err
timeout 300 python bench.py --steps 100 --no-cpu-baseline > $O/b6.json 2>> $O/bench.err
python -c "
import json
for f in ['bench_cfg3_4k_views','b6']:
    d=json.load(open('$O/'+f+'.json')); print(f, d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['ms'], d['parity']['pass'] if 'parity' in d else None)"
