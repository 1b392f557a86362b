"""Opcode mix / hottest SASS ranges of one kernel launch from an ncu report:
   python profiles/sass_mix.py rep <kernel regex> [launch index]"""
import csv, io, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
tables, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}; tables.append(cur)
    elif cur is not None and row:
        cur["rows"].append(row)
t = tables[idx]; hdr = t["rows"][0]; rows = t["rows"][1:]
isrc, iex, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot = sum(int(r[iex] or 0) for r in rows); ts = sum(int(r[isamp] or 0) for r in rows)
print(f"# {t['name'][:90]} launch {idx}: {tot} warp instructions, {ts} samples, {len(rows)} SASS lines")
mix = collections.Counter(); smix = collections.Counter()
for r in rows:
    op = r[isrc].strip().split()
    op = op[1] if op and op[0].startswith("@") else (op[0] if op else "?")
    op = op.split(".")[0]
    mix[op] += int(r[iex] or 0); smix[op] += int(r[isamp] or 0)
for op, c in mix.most_common(18):
    print(f"{op:10s} {100*c/tot:5.1f}% instr  {100*smix[op]/max(ts,1):5.1f}% samples")
