#!/bin/bash
# Multi-GPU measurement of ONE GPU count (run under `gpurun --gpus N`):  profiles/scale_run.sh N
#   bands     one 6 M @4K frame per step cut into tile-row bands (group API, one process)   BASELINE configs[3]
#   multiview the 64-camera batch over 3 M Gaussians, whole views per GPU (torchrun)        BASELINE configs[4]
#   views     the default bench line under torchrun (what the driver's SCALE run launches)
N=${1:-2}; O=gpurun_out/scale; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
[ "$N" = 1 ] && TR=python
timeout 300 python bench.py --mode bands --gpus $N --workload synthetic-6M-4k --steps 120 --warmup 5 > $O/bands_4k_n$N.json 2> $O/err_bands_n$N.log
timeout 300 python bench.py --mode bands --gpus $N --workload synthetic-6M-1080p --steps 120 --warmup 5 > $O/bands_1080p_n$N.json 2>> $O/err_bands_n$N.log
timeout 400 $TR bench.py --gpus $N --mode multiview --workload synthetic-3M-1080p --steps 5 --warmup 3 > $O/multiview_n$N.json 2> $O/err_mv_n$N.log
timeout 400 $TR bench.py --gpus $N --steps 200 --warmup 5 --no-cpu-baseline > $O/views_6M_n$N.json 2> $O/err_views_n$N.log
[ "$N" -gt 1 ] && timeout 300 python -m pytest tests -m gpu -q -k "group or two_devices" 2>&1 | tail -2 > $O/tests_n$N.log
for f in $O/bands_4k_n$N.json $O/bands_1080p_n$N.json $O/multiview_n$N.json $O/views_6M_n$N.json; do
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
