O=gpurun_out/r3t; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'])"
