import sys; sys.path.insert(0,'.')
import numpy as np, bench
from gaussian_splat_ipu_b200.splatter import B200Splatter
for wl in ("synthetic-6M-1080p","synthetic-1M-1080p","synthetic-6M-4k"):
    scn = bench.make_scene(wl)
g, proj, W, H, views = scn.g, scn.proj, scn.W, scn.H, scn.views
    s = B200Splatter(g, W, H)
    s.updateProjection(proj); s.updateFocalLengths(bench.FOV, 1.0)
    for k in (0, 5):
        s.updateModelView(views[(k) % len(views)]); s.execute(sync=True)
        pairs, culled = s.count_pairs(); st = s.stats()
        print(wl, k, "D", st.num_duplicates, "V", st.num_drawn, "staged entries", culled, culled/st.num_duplicates, "(early-stop tiles never stage all entries)")
    s.close()
