#!/bin/bash
# A/B of K5's staging batch (entries per thread and barrier pair): variant libraries built with
# -DGSPLAT_BLEND_BATCH=N as gaussian_splat_ipu_b200/libgsplat_b200_bN.so are swapped in for the
# default library on the box (the snapshot is scratch), one bench run each.
cd gaussian_splat_ipu_b200 && cp libgsplat_b200.so libgsplat_b200_b2.so && cd ..
mkdir -p gpurun_out; : > gpurun_out/ab_batch.txt
for b in 2 1 3 4 2; do
  [ -f gaussian_splat_ipu_b200/libgsplat_b200_b$b.so ] || continue
  cp gaussian_splat_ipu_b200/libgsplat_b200_b$b.so gaussian_splat_ipu_b200/libgsplat_b200.so
  timeout 120 python bench.py --no-cpu-baseline --steps 400 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('batch', $b, 'fps', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'k5_ms', round(d['roofline']['ms'], 4))" | tee -a gpurun_out/ab_batch.txt
done
cp gaussian_splat_ipu_b200/libgsplat_b200_b2.so gaussian_splat_ipu_b200/libgsplat_b200.so
