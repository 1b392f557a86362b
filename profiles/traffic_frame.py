"""DRAM traffic of ONE frame from an `ncu --set full` report that covers at least one whole frame
(project_kernel .. next project_kernel): per kernel name and per pipeline stage,
dram__bytes_read.sum + dram__bytes_write.sum summed over the frame's launches.
   python profiles/traffic_frame.py gpurun_out/prof_frame_rX.ncu-rep > profiles/traffic_rX.json
bench.py reads the newest profiles/traffic_r*.json to fill `roofline.traffic` for the stage it reports."""
import csv, io, json, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ik, ir, iw, it = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
# hardware-side view of each kernel (bench.py copies the dominant stage's into roofline.ncu)
PIPE = {"fma_pipe_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "fp64_pipe_pct": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "issue_active_pct": "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "threads_per_inst": "smsp__thread_inst_executed_per_inst_executed.ratio",
        "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram_throughput_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "tensor_pipe_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "registers": "launch__registers_per_thread"}
ipipe = {k: hdr.index(v) for k, v in PIPE.items() if v in hdr}
mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tmult = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
seq = []
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("gsplat::", "")
    b = float(r[ir].replace(",", "")) * mult[units[ir]] + float(r[iw].replace(",", "")) * mult[units[iw]]
    extra = {}
    for k, i in ipipe.items():
        try:
            extra[k] = float(r[i].replace(",", ""))
        except ValueError:
            pass
    seq.append((name, b, float(r[it].replace(",", "")) * tmult[units[it]], extra))
starts = [i for i, s in enumerate(seq) if s[0].startswith("project_kernel")]
# the last complete frame of the capture (project_kernel .. next project_kernel, or .. the end of the
# capture when it stops after a frame's last blend_kernel)
a, b = (starts[-2], starts[-1]) if len(starts) > 1 else (starts[0], len(seq))
frame = seq[a:b]

STAGES = [("k1_project", r"project_kernel|tile_count|depth_quantile"), ("k3a_sort_depth", r"select_kernel|live_sat|rs_\w+<[7854]>"),
          ("k2_duplicate", r"duplicate_kernel"), ("k3b_sort_tile", r"rs_\w+<6>"), ("k4_ranges", r"tile_ranges"),
          ("k5_blend", r"blend_kernel|tile_order"), ("post_bgr8", r"bgr8")]
# rs_scan_kernel carries no digit width: it belongs to the sort whose kernels surround it
kern, stages, last_sort = {}, {}, "k3a_sort_depth"
pipes_acc = {}
for idx, (name, by, us, extra) in enumerate(frame):
    k = kern.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "us": 0.0})
    k["launches"] += 1; k["dram_bytes"] += by; k["us"] += us
    st = next((s for s, pat in STAGES if re.search(pat, name)), None)
    if name.startswith("rs_hist_all"):   # all-digit histogram of a single-sweep sort: it belongs to the sweeps behind it
        nxt = next((n for n, _, _, _ in frame[idx + 1:] if n.startswith("rs_sweep")), "")
        st = next((s for s, pat in STAGES if re.search(pat, nxt)), None)
    if name.startswith("rs_scan"):
        st = last_sort
    elif st in ("k3a_sort_depth", "k3b_sort_tile") and name.startswith("rs_"):
        last_sort = st
    if name.startswith("rs_hist"):      # the histogram launch precedes its scan
        last_sort = st
    s = stages.setdefault(st or "other", {"launches": 0, "dram_bytes": 0.0, "us": 0.0})
    s["launches"] += 1; s["dram_bytes"] += by; s["us"] += us
    pa = pipes_acc.setdefault(st or "other", {})       # duration-weighted mean of each metric over the stage's launches
    for k, v in extra.items():
        acc = pa.setdefault(k, [0.0, 0.0])
        acc[0] += v * us; acc[1] += us
pipes = {st: {k: round(a / w, 2) for k, (a, w) in d.items() if w > 0} for st, d in pipes_acc.items()}
print(json.dumps({"source": sys.argv[1], "frame_launches": len(frame), "stages": stages, "pipes": pipes, "kernels": kern}, indent=1))
