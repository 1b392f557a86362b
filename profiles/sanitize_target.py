"""What compute-sanitizer wraps (one tool per gpurun call): __graft_entry__.smoke() — production
context, one-slice parity context, injected lists — then a 3-lane streamed orbit with a duplicate
buffer small enough that the first frames overflow and are re-run, and a band change per frame."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402
from gaussian_splat_ipu_b200 import scene as sc  # noqa: E402
from gaussian_splat_ipu_b200.splatter import B200Splatter  # noqa: E402

entry.smoke()
W, H = 640, 360
g = sc.synth_scene(80000, 0xB2000000)
g[:, 12:15] += 1.5                      # bigger splats: tiles saturate, the depth slices do something
bb = sc.bounds(g[:, 0:3])
view = sc.look_at_bounding_box(bb, (0.0, 1.0, 1.0), 1.0)
proj = sc.app_projection(g, sc.DEFAULT_FOV, W, H)
s = B200Splatter(g, W, H, concurrent_frames=3, dup_capacity=16384, depth_slices=2)
s.updateProjection(proj); s.updateFocalLengths(sc.DEFAULT_FOV, 1.0)
tickets, chk = [], 0
for k in range(12):
    s.updateModelView(sc.rotate(view, 0.3 * k, (0.0, 1.0, 0.0)))
    if k >= 6:
        s.set_band(k % 5, 18 - k % 4)
    tickets.append(s.executeStreamed())
    if len(tickets) > 4:
        img, cnt = s.acquireFrame(tickets.pop(0)); chk += int(img[::9, ::9].sum()) + int(cnt.sum())
for t in tickets:
    img, cnt = s.acquireFrame(t); chk += int(img[::9, ::9].sum()) + int(cnt.sum())
s.close()
print("sanitize target ok", chk)
