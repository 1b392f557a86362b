"""Static evidence from the built library (no GPU needed): registers / stack / static shared memory of
every kernel (cuobjdump -res-usage) and, for the kernels with asynchronous-copy, mbarrier or look-back
machinery, how often the telling SASS mnemonics occur (cuobjdump -sass).
    python profiles/sass_evidence.py > profiles/sass_evidence_rX.txt"""
import collections
import os
import re
import subprocess

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gaussian_splat_ipu_b200", "libgsplat_b200.so")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
rows, cur = [], None
for line in res.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+)", line)
    if m and cur:
        rows.append((cur, int(m.group(1)), int(m.group(2)), int(m.group(3))))
        cur = None
names = demangle([r[0] for r in rows])
print("# libgsplat_b200.so, sm_100a: per-kernel resources (ptxas) and telling SASS mnemonics\n")
print(f"{'kernel':78s} {'regs':>4s} {'stack':>5s} {'static smem':>11s}")
for fn, reg, stack, sh in sorted(rows, key=lambda r: names[r[0]]):
    short = re.sub(r"\(.*", "", names[fn]).replace("void ", "").replace("gsplat::", "")
    print(f"{short:78s} {reg:4d} {stack:5d} {sh:11d}")

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
TELL = ["LDGSTS", "ARRIVES.LDGSTSBAR", "SYNCS.PHASECHK", "SYNCS.ARRIVE", "SYNCS.EXCH", "UBLKCP", "NANOSLEEP", "DFMA", "DMUL", "DADD",
        "F2F.F64.F32", "F2F.F32.F64", "ATOMG", "ATOMS", "LD.E.64.STRONG.GPU", "ST.E.64.STRONG.GPU", "HMMA", "UTCHMMA", "UTMALDG", "FFMA", "FMUL", "FADD"]
counts, fn = collections.defaultdict(collections.Counter), None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and fn:
        op = m.group(1)
        for t in TELL:
            if op == t or op.startswith(t + ".") or (t in ("LD.E.64.STRONG.GPU", "ST.E.64.STRONG.GPU") and op.startswith(t)):
                counts[fn][t] += 1
        counts[fn]["(all)"] += 1
names2 = demangle(list(counts))
print("\n# SASS mnemonics (static counts) of the kernels that stage through shared memory or wait on one another")
for fn in sorted(counts, key=lambda f: names2[f]):
    short = re.sub(r"\(.*", "", names2[fn]).replace("void ", "").replace("gsplat::", "")
    if not re.search(r"blend_kernel|rs_sweep_kernel<7>|rs_hist_all|duplicate_kernel<0>|select_kernel<0>|project_kernel|cov3", short):
        continue
    c = counts[fn]
    print(f"{short:44s} " + "  ".join(f"{t}:{c[t]}" for t in TELL if c[t]) + f"  (all:{c['(all)']})")
print("\n# no tensor-core instruction (HMMA / UTCHMMA) and no TMA bulk copy (UBLKCP / UTMALDG) in the default build: nothing on this path is a dense\n"
      "# contraction, and the 48-byte record gather is served by LDGSTS + mbarrier (DESIGN.md section 4; -DGSPLAT_BLEND_TMA=1 builds the UBLKCP form)")
