O=gpurun_out/r3l; mkdir -p $O
for i in 1 2; do
GSPLAT_CONCURRENT_FRAMES=1 python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>$O/err$i.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], json.dumps(d['e2e']))"
done
GSPLAT_CONCURRENT_FRAMES=1 GSPLAT_SORT_SWEEP=0 python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>$O/err3.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], json.dumps(d['e2e']))"
tail -3 $O/err1.log
