O=gpurun_out/r3v; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
bash profiles/ab_k5.sh $O/ab.txt default=$L allnarrow=$L:GSPLAT_K5_WIDE=0 allwide=$L:GSPLAT_K5_WIDE=1 default2=$L
for i in 1; do
GSPLAT_CONCURRENT_FRAMES=1 python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('one frame at a time, wide:', d['value'], d['e2e']['value'], d['e2e']['sequential']['value'])"
GSPLAT_K5_WIDE=0 GSPLAT_CONCURRENT_FRAMES=1 python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('one frame at a time, narrow:', d['value'], d['e2e']['value'], d['e2e']['sequential']['value'])"
done
