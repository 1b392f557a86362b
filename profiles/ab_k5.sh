#!/usr/bin/env bash
# A/B of K5 builds on the GPU box: frames/s (3 lanes) and the K5 stage time of the 6M @1080p bench frame.
#   profiles/ab_k5.sh out.txt  name=libpath[:ENV=VAL] ...
out=$1; shift
mkdir -p "$(dirname "$out")"; : > "$out"
for spec in "$@"; do
  name=${spec%%=*}; rest=${spec#*=}; lib=${rest%%:*}; envs=""
  [ "$rest" != "$lib" ] && envs=${rest#*:}
  line=$(env GSPLAT_LIB="$lib" $envs python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null)
  python - "$name" "$line" >> "$out" <<'PY'
import json, sys
d = json.loads(sys.argv[2])
st = d["stages"]
print(f"{sys.argv[1]:>14}: {d['value']:8.1f} frames/s  e2e {d['e2e']['value']:8.1f}  one-slice {d['one_slice']['value']:7.1f}  "
      f"K5 {st['k5_blend']['ms']:.4f} ms (frac {st['k5_blend']['frac']:.3f})  K1 {st['k1_project']['ms']:.4f}  "
      f"sortA {st['k3a_sort_depth']['ms']:.4f}  K2 {st['k2_duplicate']['ms']:.4f}  sortB {st['k3b_sort_tile']['ms']:.4f}  launches/frame {d['gpu_launches'] / d['steps']:.0f}")
PY
done
cat "$out"
