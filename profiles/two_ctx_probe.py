"""Throughput of ONE GPU when frames alternate between two (or more) contexts, i.e. two frame graphs
in flight on different streams: the frame is a chain of short latency-bound launches, so two
independent chains interleave.   python profiles/two_ctx_probe.py [workload] [contexts]"""
import sys, time
sys.path.insert(0, '.')
import numpy as np, bench
from gaussian_splat_ipu_b200.splatter import B200Splatter
wl = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
scn = bench.make_scene(wl)
g, proj, W, H, views = scn.g, scn.proj, scn.W, scn.H, scn.views
views = [views[(k) % len(views)] for k in range(bench.ORBIT_VIEWS)]
for nctx in ([int(sys.argv[2])] if len(sys.argv) > 2 else [1, 2, 3]):
    ss = [B200Splatter(g, W, H) for _ in range(nctx)]
    for s in ss:
        s.updateProjection(proj); s.updateFocalLengths(bench.FOV, 1.0)
    def run(K):
        for k in range(K):
            s = ss[k % nctx]; s.updateModelView(views[k % 64]); s.execute()
        for s in ss: s.sync()
    run(60 * nctx)
    K = 600
    t0 = time.perf_counter(); run(K); dt = time.perf_counter() - t0
    # streamed delivery too, at several depths of frames in flight
    res = []
    for depth in (nctx, 2 * nctx, 3 * nctx - 1):
        tick = []
        t0 = time.perf_counter()
        for k in range(K):
            s = ss[k % nctx]; s.updateModelView(views[k % 64]); tick.append((s, s.executeStreamed()))
            if len(tick) > depth:
                a, t = tick.pop(0); a.acquireFrame(t)
        for a, t in tick: a.acquireFrame(t)
        res.append((depth, round(K / (time.perf_counter() - t0), 1)))
    dt2 = K / res[1][1]
    print(f"{wl} contexts {nctx}: {K / dt:.1f} frames/s device-resident, {res} frames/s streamed to host at (frames in flight)")
    for s in ss: s.close()
