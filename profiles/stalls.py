"""Top stall sites of one kernel from an ncu report (source page, SASS level):
   python profiles/stalls.py gpurun_out/prof_X.ncu-rep <kernel regex> [launch index] [top N]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
# the output holds one table per launch, each starting with a "Kernel Name" line
tables, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}; tables.append(cur)
    elif cur is not None and row:
        cur["rows"].append(row)
t = tables[idx]
hdr = t["rows"][0]; rows = t["rows"][1:]
i_src, i_samp = hdr.index("Source"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[i_samp] or 0) for r in rows)
print(f"# {t['name'][:100]}  launch {idx}: {tot} samples")
agg = {}
for r in rows:
    for i in stall_cols:
        agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
print("stall totals:", ", ".join(f"{k[6:]}={100*v/max(tot,1):.0f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(rows, key=lambda r: -int(r[i_samp] or 0))[:top]:
    why = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{100*int(r[i_samp])/max(tot,1):5.1f}%  {r[i_src].strip()[:70]:70s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
