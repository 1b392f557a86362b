"""Per-SASS-instruction execution counts of one kernel launch from an ncu report (source page):
   python profiles/sass_counts.py report.ncu-rep [launch index] > listing.txt
columns: line, warp instructions executed, share, average active threads, stall samples, instruction"""
import csv, subprocess, sys
rep = sys.argv[1]
idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
tables, hdr = [], None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Kernel Name":
        tables.append([]); continue
    if r and r[0] == "Address":
        hdr = r; continue
    if hdr and r and tables:
        tables[-1].append(r)
rows = tables[idx]
iS, iI, iT, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Avg. Threads Executed"), hdr.index("# Samples")
tot = sum(int(r[iI]) for r in rows)
print(f"# launch {idx}: {tot} warp instructions, {sum(int(r[iSm]) for r in rows)} samples, {len(rows)} SASS lines")
for k, r in enumerate(rows):
    print(f"{k:4d} {int(r[iI]):10d} {100 * int(r[iI]) / tot:5.2f}% thr={r[iT]:>5s} smp={r[iSm]:>5s}  {r[iS].strip()}")
