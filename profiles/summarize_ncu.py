"""Summarises ncu outputs brought back from the GPU box into small text files under profiles/.
  python profiles/summarize_ncu.py launches gpurun_out/launches_X.csv  > profiles/launches_X.txt
  python profiles/summarize_ncu.py report   gpurun_out/prof_X.ncu-rep  > profiles/prof_X.txt
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    hdr = rows[0]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv:
            continue
        name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("gsplat::", "")
        ns = float(r[iv].replace(",", ""))
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot/1e6:.3f} ms total (ncu-serialised, cold cache)")
    print(f"{'kernel':40s} {'launches':>8s} {'total us':>12s} {'avg us':>10s} {'share':>7s}")
    for k, (c, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {c:8d} {ns/1e3:12.1f} {ns/1e3/c:10.1f} {100*ns/tot:6.1f}%")


KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
        "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_xu.sum",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warp_issue_stalled_barrier_per_warp_active.pct",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__cycles_active.avg",
        "smsp__cycles_active.avg", "smsp__thread_inst_executed_per_inst_executed.ratio"]


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"\n## {d.get('Kernel Name', '?')[:110]}  grid {d.get('launch__grid_size')} block {d.get('launch__block_size')}")
        for k in KEYS:
            if k in d:
                print(f"{k:70s} {d[k]:>18s} {units[hdr.index(k)]}")
        rd = float(d.get("dram__bytes_read.sum", "0").replace(",", "") or 0)
        wr = float(d.get("dram__bytes_write.sum", "0").replace(",", "") or 0)
        u = units[hdr.index("dram__bytes_read.sum")] if "dram__bytes_read.sum" in hdr else ""
        print(f"{'traffic = dram read + write':70s} {rd + wr:18.3f} {u}")


if __name__ == "__main__":
    {"launches": launches, "report": report}[sys.argv[1]](sys.argv[2])
