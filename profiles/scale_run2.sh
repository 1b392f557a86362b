#!/bin/bash
# Reduced multi-GPU pass (GPU-minutes are charged N x):  profiles/scale_run2.sh N "views multiview bands4k bands1080 tests"
N=${1:-2}; WHAT=${2:-"views multiview bands4k"}; O=gpurun_out/scale2; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
[ "$N" = 1 ] && TR=python
for w in $WHAT; do case $w in
  bands4k)   timeout 300 python bench.py --mode bands --gpus $N --workload synthetic-6M-4k --steps 120 --warmup 5 > $O/bands_4k_n$N.json 2> $O/err_bands4k_n$N.log;;
  bands1080) timeout 300 python bench.py --mode bands --gpus $N --workload synthetic-6M-1080p --steps 120 --warmup 5 > $O/bands_1080p_n$N.json 2> $O/err_bands1080_n$N.log;;
  multiview) timeout 400 $TR bench.py --gpus $N --mode multiview --workload synthetic-3M-1080p --steps 5 --warmup 3 > $O/multiview_n$N.json 2> $O/err_mv_n$N.log;;
  views)     timeout 400 $TR bench.py --gpus $N --steps 200 --warmup 5 --no-cpu-baseline > $O/views_6M_n$N.json 2> $O/err_views_n$N.log;;
  tests)     timeout 300 python -m pytest tests -m gpu -q -k "group or two_devices" 2>&1 | tail -2 > $O/tests_n$N.log;;
esac; done
for f in $O/*_n$N.json; do
  python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1].split("/")[-1], d["n_gpus"], "GPUs:", round(d["value"], 1), "frames/s", "e2e", round((d.get("e2e") or {}).get("value", 0), 1), d.get("last_cuts", ""))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
cat $O/tests_n$N.log 2>/dev/null
