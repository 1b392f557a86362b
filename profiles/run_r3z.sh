O=gpurun_out/r3z; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
grep -q passed $O/tests.log || exit 1
bash profiles/ab_k5.sh $O/ab.txt default=$L k1plain=$L:GSPLAT_K1_DENSE=0 alldense=$L:GSPLAT_K1_DENSE=1
timeout 500 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || tail -5 $O/bench.err
GSPLAT_CONCURRENT_FRAMES=1 python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('one frame at a time:', d['value'], d['e2e']['value'], d['e2e']['sequential']['value'])"
