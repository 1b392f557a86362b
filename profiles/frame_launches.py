"""Prints the kernel sequence of one frame from an ncu launch list (gpu__time_duration.sum csv):
    python profiles/frame_launches.py gpurun_out/launches_x.csv"""
import csv, re, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
h = rows[0]; ik, iv = h.index('Kernel Name'), h.index('Metric Value')
seq = [(re.sub(r'\(.*', '', r[ik]).replace('void ', '').replace('gsplat::', ''), float(r[iv].replace(',', '')) / 1e3)
       for r in rows[1:] if len(r) > iv]
idx = [i for i, (k, _) in enumerate(seq) if k.startswith('project_kernel')]
a, b = idx[-3], idx[-2]
for k, v in seq[a:b]:
    print(f"{k:45s} {v:8.1f}")
print(f"{'frame total (ncu-serialised)':45s} {sum(v for _, v in seq[a:b]):8.1f}   launches {b - a}")
