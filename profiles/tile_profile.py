"""Per-tile load picture of K5 for one view of a bench workload (run on the GPU box):
    python profiles/tile_profile.py [workload] [view]
Prints the distribution of per-tile list lengths, visited pairs and CTA cycles (counting K5)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gaussian_splat_ipu_b200.splatter import B200Splatter  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
view = int(sys.argv[2]) if len(sys.argv) > 2 else 3
scn = bench.make_scene(wl)
g, proj, W, H, views = scn.g, scn.proj, scn.W, scn.H, scn.views
s = B200Splatter(g, W, H, collect_timings=True)
s.updateProjection(proj); s.updateFocalLengths(bench.FOV, 1.0)
for _ in range(3):
    s.updateModelView(views[(view) % len(views)]); s.execute(sync=True)
print(wl, "view", view, "timings", {k: round(v, 4) for k, v in s.timings().items()})
pairs, blended, cyc, prs = s.profile_tiles()
cnt = s.getIPUHistogram().astype(np.int64)
st = s.stats()
print("N", st.num_gaussians, "V", st.num_drawn, "D", st.num_duplicates, "pairs", pairs, "blended", blended)
order = np.argsort(-cyc.astype(np.int64))
print("tiles", len(cnt), "nonempty", int((cnt > 0).sum()), "count pct 50/90/99/max",
      [int(np.percentile(cnt, p)) for p in (50, 90, 99, 100)])
print("cycles pct 50/90/99/max", [int(np.percentile(cyc, p)) for p in (50, 90, 99, 100)], "sum", int(cyc.sum()))
print("top tiles by cycles: (tile, row, col, count, visited pairs, visited/px, cycles)")
for t in order[:12]:
    print("  ", int(t), int(t) // s.tiles_across, int(t) % s.tiles_across, int(cnt[t]), int(prs[t]),
          round(int(prs[t]) / (s.tile_w * s.tile_h), 1), int(cyc[t]))
if len(sys.argv) > 3:       # raw per-tile arrays for fitting a cost model (band balancing)
    np.savez_compressed(sys.argv[3], counts=cnt, cycles=cyc, pairs=prs, tiles_across=s.tiles_across, timings=np.array(list(s.timings().values())))
