"""Which tiles stay live into the second depth slice, and how much of its work is theirs."""
import sys; sys.path.insert(0, '.')
import numpy as np, bench
from gaussian_splat_ipu_b200.splatter import B200Splatter
scn = bench.make_scene("synthetic-6M-1080p")
g, proj, W, H, views = scn.g, scn.proj, scn.W, scn.H, scn.views
s = B200Splatter(g, W, H, debug_records=True, depth_slices=2)
s.updateProjection(proj); s.updateFocalLengths(bench.FOV, 1.0); s.updateModelView(views[(5) % len(views)]); s.execute(sync=True)
off, lst = s.tile_lists(); lens = np.diff(off); counts = s.getIPUHistogram().astype(np.int64); st = s.stats()
live = (lens == counts) & (counts > 0)
print("binned", st.num_duplicates, "ref", st.num_reference_duplicates, "depth_sorted", st.num_depth_sorted)
print("tiles nonempty", int((counts > 0).sum()), "complete lists (live to the end)", int(live.sum()))
for thr in (2000, 5000, 10000, 20000, 50000):
    m = live & (counts > thr)
    print(f"  live tiles with count > {thr}: {int(m.sum())}, their entries {int(lens[m].sum())} ({lens[m].sum() / max(st.num_duplicates,1):.2%} of binned)")
