"""CPU tests of the multi-GPU host logic: band balancing, view assignment, and the
band gather over a world_size-2 gloo group (the N>1 path without GPUs).  The oracle supplies real
per-tile counts."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import load_golden
from gaussian_splat_ipu_b200 import sharding as sh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_balance_bands_properties():
    rng = np.random.default_rng(1)
    for rows, n in ((54, 8), (108, 8), (36, 2), (9, 4), (3, 8), (1, 1), (54, 1)):
        cost = rng.gamma(0.5, 1000.0, rows) + 1.0
        cuts = sh.balance_bands(cost, n)
        assert len(cuts) == n + 1 and cuts[0] == 0 and cuts[-1] == rows
        assert all(b >= a for a, b in zip(cuts[:-1], cuts[1:]))
        if n <= rows:
            assert all(b > a for a, b in zip(cuts[:-1], cuts[1:])), "every band needs a row"
    # equal costs -> equal bands
    assert sh.balance_bands(np.ones(64), 8) == list(range(0, 65, 8))
    # one hot row dominates: it gets a band of its own
    cost = np.ones(20); cost[7] = 1000.0
    cuts = sh.balance_bands(cost, 4)
    assert any(a == 7 and b == 8 for a, b in zip(cuts[:-1], cuts[1:])) or sh.band_imbalance(cost, cuts) < 4.0
    with pytest.raises(ValueError):
        sh.balance_bands(np.ones(4), 0)


def test_balance_bands_on_oracle_counts(port):
    """Counts of a real frame (in-cloud style imbalance): balanced bands beat equal-height bands."""
    z = load_golden("point_cloud_12_640x360_div1.0.npz")
    TA = int(z["W"]) // 32
    cost = sh.row_costs(z["counts"], TA)
    rows = cost.size
    for n in (2, 4):
        bal = sh.balance_bands(cost, n)
        even = [round(rows * i / n) for i in range(n + 1)]
        assert sh.band_imbalance(cost, bal) <= sh.band_imbalance(cost, even) + 1e-9
        assert sh.band_imbalance(cost, bal) < 1.6


def test_assign_views():
    costs = [5, 1, 1, 1, 9, 2, 2, 3]
    a = sh.assign_views(costs, 3)
    assert sorted(i for v in a for i in v) == list(range(8))
    loads = [sum(costs[i] for i in v) for v in a]
    assert max(loads) <= 9           # the heaviest view alone bounds the makespan here
    assert a == sh.assign_views(costs, 3)   # deterministic
    assert sh.assign_views([], 2) == [[], []]
    assert sh.round_robin_views(7, 1, 3) == [1, 4]
    # weak-scaling orbit: equal costs -> equal counts
    a = sh.assign_views(np.ones(64), 8)
    assert all(len(v) == 8 for v in a)


def test_band_slices():
    cuts = [0, 3, 3, 9]
    sl = sh.band_slices(cuts, 20, 64, 3)
    assert sl == [(0, 3 * 20 * 192), (3 * 20 * 192, 0), (3 * 20 * 192, 6 * 20 * 192)]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from gaussian_splat_ipu_b200 import sharding as shw
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        W, H, th = 64, 180, 20
        rows = H // th
        # every rank derives the same cuts from the same (all-reduced) counts
        counts = torch.zeros(rows * 2, dtype=torch.int32)
        local = np.arange(rows * 2, dtype=np.int32) * (rank + 1)
        band0 = [0, 4, rows]
        t0, t1 = band0[rank] * 2, band0[rank + 1] * 2
        counts[t0:t1] = torch.from_numpy(local[t0:t1])
        shw.allgather_tile_counts(counts)
        cuts = shw.balance_bands(shw.row_costs(counts.numpy(), 2, tile_pixels=640), world)
        # each rank "renders" only its band of a known image
        truth = (np.arange(H * W * 3) % 251).astype(np.uint8)
        img = torch.zeros(H * W * 3, dtype=torch.uint8)
        off, ln = shw.band_slices(cuts, th, W)[rank]
        img[off:off + ln] = torch.from_numpy(truth[off:off + ln])
        out = shw.gather_bands(img, cuts, th, W, 3, dst=0)
        ok = True
        if rank == 0:
            ok = out is not None and np.array_equal(out.numpy(), truth)
        else:
            ok = out is None
        # float tiled framebuffer bands gather the same way (channels = 4)
        fb = torch.full((H * W * 4,), float(rank + 1))
        out = shw.gather_bands(fb, cuts, th, W, 4, dst=0)
        if rank == 0:
            exp = np.concatenate([np.full(l, r + 1.0, np.float32) for r, (o, l) in enumerate(shw.band_slices(cuts, th, W, 4))])
            ok = ok and np.array_equal(out.numpy(), exp)
        q.put((rank, bool(ok), cuts))
    finally:
        dist.destroy_process_group()


def test_gather_bands_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res), res
    assert res[0][2] == res[1][2], "ranks disagree on the band cuts"


def test_native_band_balancer_equals_the_python_mirror():
    """gsplat_balance_bands (what gsplat_create_group's GSPLAT_GROUP_BANDS mode cuts every frame with,
    host only) gives exactly sharding.balance_bands(sharding.row_costs(...)) — the rule the gloo
    world-2 test above exercises."""
    from gaussian_splat_ipu_b200.splatter import balance_bands_native
    rng = np.random.default_rng(8)
    for TA, TD in ((40, 36), (60, 54), (120, 108), (4, 3)):
        for n in (1, 2, 3, 4, 8):
            for k in range(6):
                counts = rng.integers(0, 4000, TA * TD).astype(np.uint32)
                if k == 0:
                    counts[:] = 0
                if k == 1:
                    counts[: TA * (TD // 2)] = 0          # everything in the lower half
                if k == 2:
                    counts[TA * (TD // 3): TA * (TD // 3 + 1)] *= 200   # one very heavy row
                want = sh.balance_bands(sh.row_costs(counts, TA, 640), n)
                assert balance_bands_native(counts, TA, TD, n, 640) == want, (TA, TD, n, k)
