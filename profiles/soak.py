"""Soak: thousands of streamed frames through the default context (frames in flight, depth slices,
frame graphs) on the bench scene; every delivered frame's image and counts must equal the ones a
single-lane, one-slice context renders for the same camera.   python profiles/soak.py [frames]"""
import sys, time, zlib
sys.path.insert(0, '.')
import numpy as np, bench
from gaussian_splat_ipu_b200.splatter import B200Splatter
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 1920
scn = bench.make_scene(bench.DEFAULT_WORKLOAD)
g, proj, W, H, views = scn.g, scn.proj, scn.W, scn.H, scn.views
views = [views[(k) % len(views)] for k in range(bench.ORBIT_VIEWS)]
ref = B200Splatter(g, W, H, concurrent_frames=1, depth_slices=1)
ref.updateProjection(proj); ref.updateFocalLengths(bench.FOV, 1.0)
want = []
for v in views:
    ref.updateModelView(v); ref.execute()
    want.append((zlib.crc32(ref.getFrameBuffer().tobytes()), zlib.crc32(ref.getIPUHistogram().tobytes())))
ref.close()
s = B200Splatter(g, W, H)
s.updateProjection(proj); s.updateFocalLengths(bench.FOV, 1.0)
depth, tickets, got, bad = 2 * s.concurrent_frames, [], 0, 0
t0 = time.perf_counter()
def take():
    global got, bad
    img, cnt = s.acquireFrame(tickets[got])
    if (zlib.crc32(img.tobytes()), zlib.crc32(cnt.tobytes())) != want[got % len(views)]:
        bad += 1
    got += 1
for k in range(frames):
    s.updateModelView(views[k % len(views)]); tickets.append(s.executeStreamed())
    if len(tickets) - got > depth: take()
while got < frames: take()
dt = time.perf_counter() - t0
print(f"soak: {frames} frames, lanes {s.concurrent_frames}, {frames / dt:.0f} frames/s incl. crc of every frame, mismatches {bad}")
s.close()
sys.exit(1 if bad else 0)
