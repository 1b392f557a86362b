"""Debug aid (GPU box): render a PLY orbit, report every frame whose framebuffer is not bit-identical to the
oracle's, and save the GPU framebuffer of the worst frame for offline tracing."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pyoracle
from gaussian_splat_ipu_b200 import scene as sc
from gaussian_splat_ipu_b200.splatter import B200Splatter

name = sys.argv[1] if len(sys.argv) > 1 else "point_cloud_10.ply"
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/debug_fb.npz"
W, H = 1280, 720
g = sc.gaussians_from_ply(sc.load_ply(os.path.join(pyoracle.REF_DATA, name)))
proj = sc.app_projection(g, sc.DEFAULT_FOV, W, H)
port = pyoracle.Oracle("port")
s = B200Splatter(g, W, H, debug_records=True, depth_slices=1)
s.updateProjection(proj); s.updateFocalLengths(sc.DEFAULT_FOV, div)
worst = (0.0, None)
for k in range(32):
    v = sc.orbit_view(k, 32)
    f = port.frame(g, v, proj, sc.DEFAULT_FOV, div, W, H)
    s.updateModelView(v); s.execute(sync=True)
    fb = s.framebuffer_tiled().copy()
    nd = np.flatnonzero(fb.view(np.uint32) != f.fb.view(np.uint32))
    if nd.size:
        d = np.abs(fb[nd] - f.fb[nd])
        print(f"frame {k}: {nd.size} floats differ, max {d.max():.3e}; first: "
              + ", ".join(f"[{i} t{i//4//640} p{i//4%640} c{i%4} gpu {fb[i]!r} ref {f.fb[i]!r}]" for i in nd[:4]))
        if d.max() > worst[0]:
            worst = (float(d.max()), (k, fb))
if worst[1]:
    np.savez_compressed(out, k=worst[1][0], fb=worst[1][1], name=name, div=div)
    print("saved frame", worst[1][0], "to", out)
