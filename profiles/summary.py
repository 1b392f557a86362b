"""Markdown table of the committed bench lines of one round:  python profiles/summary.py r2"""
import glob
import json
import os
import sys

R = sys.argv[1] if len(sys.argv) > 1 else "r2"
here = os.path.dirname(os.path.abspath(__file__))
rows = []
for f in sorted(glob.glob(os.path.join(here, f"bench_*_{R}.json")) + glob.glob(os.path.join(here, f"bench_*_{R}.jsonl"))):
    for line in open(f):
        line = line.strip()
        if not line.startswith("{"):
            continue
        try:
            d = json.loads(line)
        except ValueError:
            try:
                d = json.load(open(f))
            except ValueError:
                continue
            rows.append((os.path.basename(f), d))
            break
        rows.append((os.path.basename(f), d))
print("| file | workload | mode / impl | GPUs | frames/s | e2e frames/s | CPU reference frames/s (cores) | parity | K5 frac |")
print("|---|---|---|---|---|---|---|---|---|")
for f, d in rows:
    cfg = d.get("config", {})
    cpu = d.get("cpu_baseline") or {}
    par = d.get("parity") or {}
    ptxt = "" if not par else ("pass" if par.get("pass") else "FAIL") + f" (max|d| {par.get('fb_max_abs', 0):.1e})"
    extra = ""
    if cpu.get("o1_as_shipped"):
        extra = f"; O1 as shipped {cpu['o1_as_shipped']['value']:.2f}; 1 thread {cpu['o2_one_thread']['value']:.2f}"
    roof = d.get("roofline") or {}
    print(f"| `{f}` | {cfg.get('workload', '')} | {d.get('impl', d.get('mode', 'views'))} | {d.get('n_gpus', '')} | {d.get('value', 0):.1f} | "
          f"{(d.get('e2e') or {}).get('value', 0):.1f} | {('%.2f (%d)' % (cpu['value'], cpu['cores'])) if cpu.get('value') else ''}{extra} | {ptxt} | "
          f"{('%.3f' % roof['frac']) if roof.get('frac') else ''} |")
