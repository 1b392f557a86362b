"""Generates tests/golden/*.npz from the REFERENCE's own code (oracle/_ref/libsplat_ref.so, built
by oracle/build_ref.sh from /root/reference).  Run in the authoring container only:

    python tests/golden/make_golden.py

Each fixture holds the inputs (a subsample of a shipped PLY, camera, fxy, geometry) and the
reference's outputs for them: projected records, tile rectangles, per-tile lists/counts, and the
tile-major float framebuffer of a small image.  The reference publishes no golden vectors for
this path, so these are the pin for the C restatement and for the CUDA path."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402
from gaussian_splat_ipu_b200 import scene as sc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    po.build_ref()
    ref = po.Oracle("reference")
    rng = np.random.default_rng(20261104)
    for name, W, H, div, nsub in (("point_cloud_10", 320, 180, 1.0, 3000), ("point_cloud_9", 320, 180, 0.1, 3000),
                                  ("point_cloud_12", 640, 360, 1.0, 4000)):
        g = sc.gaussians_from_ply(sc.load_ply(os.path.join(po.REF_DATA, name + ".ply")))
        proj = sc.app_projection(g, sc.DEFAULT_FOV, W, H)       # frustum from the FULL scene bbox
        idx = np.sort(rng.choice(g.shape[0], nsub, replace=False))
        sub = g[idx].copy()
        sub[:, 15] = np.arange(1, nsub + 1, dtype=np.float32)
        sub[::97, 15] = 0.0                                      # a few empty slots (gid <= 0)
        view = sc.orbit_view(3)
        f = ref.frame(sub, view, proj, sc.DEFAULT_FOV, div, W, H)
        np.savez_compressed(os.path.join(OUT, f"{name}_{W}x{H}_div{div}.npz"), g16=sub, view=view, proj=proj,
                            fov=np.float32(sc.DEFAULT_FOV), scale_div=np.float32(div), W=W, H=H, rec=f.rec,
                            rect=f.rect, counts=f.counts, offsets=f.offsets, lists=f.lists, fb=f.fb)
        print(name, "drawn", int(f.drawn.sum()), "D", len(f.lists), "max/tile", int(f.counts.max()))
    # per-function known answers: ComputeConicOpacity and getTileBounds
    covs = rng.normal(size=(256, 3)).astype(np.float32) * 4
    covs[:, 0] = np.abs(covs[:, 0]) + 0.3
    covs[:, 2] = np.abs(covs[:, 2]) + 0.3
    covs[0] = (1.0, 1.0, 1.0)   # det == 0
    ops = rng.uniform(-6, 20, 256).astype(np.float32)
    conic = np.stack([ref.conic_opacity(c, float(o)) for c, o in zip(covs, ops)])
    tb = np.stack([ref.tile_bounds(1280, 720, 40, 36, t) for t in (0, 1, 3, 39, 40, 120, 1439)])
    np.savez_compressed(os.path.join(OUT, "kat_functions.npz"), covs=covs, ops=ops, conic=conic,
                        tile_ids=np.array([0, 1, 3, 39, 40, 120, 1439]), tile_bounds=tb)


if __name__ == "__main__":
    main()
