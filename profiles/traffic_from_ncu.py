"""DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) per launch of every kernel in an
`ncu --set full` report, averaged per kernel name and summed per pipeline stage of ONE frame:
   python profiles/traffic_from_ncu.py gpurun_out/prof_frame_rX.ncu-rep > profiles/traffic_rX.json
bench.py reads the newest profiles/traffic_r*.json to fill `roofline.traffic` (bytes per launch /
per stage and frame, same unit as `achieved`'s numerator)."""
import csv, io, json, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ik, ir, iw, it = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tmult = {"ns": 1e-3, "us": 1.0, "ms": 1e3}
per = {}
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("gsplat::", "")
    b = float(r[ir].replace(",", "")) * mult[units[ir]] + float(r[iw].replace(",", "")) * mult[units[iw]]
    per.setdefault(name, []).append((b, float(r[it].replace(",", "")) * tmult[units[it]]))
kern = {k: {"launches": len(v), "dram_bytes_per_launch": sum(x[0] for x in v) / len(v),
            "us_per_launch": sum(x[1] for x in v) / len(v)} for k, v in per.items()}
print(json.dumps({"source": sys.argv[1], "kernels": kern}, indent=1))
