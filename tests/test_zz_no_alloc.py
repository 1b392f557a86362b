"""gsplat_sort_pairs_static: the K3 device radix sort with its scratch in a file-scope device array
(no cudaMalloc in the call; include/gsplat_b200.h).  No counterpart in the reference.
Checked through the C ABI against the existing entry point
gsplat_sort_pairs (bit-exact: integer work) and against numpy's stable sort, on two calls in a row
that share the scratch array.  Collected last on purpose (file name)."""
import numpy as np
import pytest

from gaussian_splat_ipu_b200 import splatter as sp

pytestmark = pytest.mark.gpu


def _device_sort(torch, keys, vals, begin_bit, end_bit):
    dk = torch.from_numpy(keys.view(np.int32)).cuda()
    dv = torch.from_numpy(vals.view(np.int32)).cuda()
    ok = torch.empty_like(dk)
    ov = torch.empty_like(dv)
    torch.cuda.synchronize()
    sp.sort_pairs_static(0, dk.data_ptr(), dv.data_ptr(), ok.data_ptr(), ov.data_ptr(), keys.size, begin_bit, end_bit,
                         torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return ok.cpu().numpy().view(np.uint32), ov.cpu().numpy().view(np.uint32)


def _numpy_sort(keys, vals, begin_bit, end_bit):
    digit = (keys >> np.uint32(begin_bit)) & np.uint32((1 << (end_bit - begin_bit)) - 1 if end_bit - begin_bit < 32 else 0xFFFFFFFF)
    order = np.argsort(digit, kind="stable")
    return keys[order], vals[order]


def test_static_sort_matches_the_context_sort_twice_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    rng = np.random.default_rng(2024)
    s = sp.B200Splatter(sp.synth_scene(4096, 1), 256, 160)   # the existing entry point needs a context
    # two calls in a row on the shared scratch: the second is SHORTER, sorts other bits and needs a
    # different number of passes (its histogram rows and intermediate buffer hold the first call's data)
    cases = [(700_001, 0, 32), (123_457, 3, 21), (2048, 0, 12), (1, 0, 32), (sp.SORT_STATIC_MAX, 5, 17)]
    for n, b0, b1 in cases:
        keys = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        keys[: n // 3] &= np.uint32(0x0000FFFF)                 # many equal digits: stability matters
        vals = rng.permutation(n).astype(np.uint32)
        k_ref, v_ref = s.sort_pairs(keys, vals, b0, b1)
        k_new, v_new = _device_sort(torch, keys, vals, b0, b1)
        assert np.array_equal(k_new, k_ref) and np.array_equal(v_new, v_ref), (n, b0, b1)
        k_np, v_np = _numpy_sort(keys, vals, b0, b1)
        assert np.array_equal(k_new, k_np) and np.array_equal(v_new, v_np), (n, b0, b1)


def test_static_sort_rejects_what_it_cannot_hold():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    buf = torch.zeros(64, dtype=torch.int32, device="cuda")
    p = buf.data_ptr()
    with pytest.raises(sp.GsplatError, match="GSPLAT_SORT_STATIC_MAX") as ei:
        sp.sort_pairs_static(0, p, p, p, p, sp.SORT_STATIC_MAX + 1)
    assert ei.value.code == -3
    with pytest.raises(sp.GsplatError, match="distinct"):
        sp.sort_pairs_static(0, p, p, p, p, 8)
    with pytest.raises(sp.GsplatError, match="aligned"):
        sp.sort_pairs_static(0, p, p + 64, p + 4, p + 128, 8)
    with pytest.raises(sp.GsplatError, match="bit range"):
        sp.sort_pairs_static(0, p, p + 64, p + 128, p + 192, 8, 5, 5)
    sp.sort_pairs_static(0, 0, 0, 0, 0, 0)   # nothing to sort: no pointers needed
