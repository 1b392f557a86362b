"""A handful of frames of a bench workload, one frame at a time (one lane, fixed slicing): the
command ncu wraps.   python profiles/frames.py [workload] [frames] [slices]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gaussian_splat_ipu_b200.splatter import B200Splatter  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 6
slices = int(sys.argv[3]) if len(sys.argv) > 3 else 2
scn = bench.make_scene(wl)
s = B200Splatter(scn.g, scn.W, scn.H, concurrent_frames=1, depth_slices=slices)
s.updateProjection(scn.proj); s.updateFocalLengths(bench.FOV, scn.scale_div)
for k in range(frames):
    s.updateModelView(scn.views[(5 + k) % len(scn.views)]); s.execute(sync=True)
st = s.stats()
print("frames", frames, "N", st.num_gaussians, "D", st.num_duplicates, "slices", st.depth_slices)
