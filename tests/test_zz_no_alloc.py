"""bfq_sort_pairs_noalloc (include/bfq_gpu.h): the allocation-free entry point, whose scratch is a fixed device array
inside the library, against the existing context-based entry point bfq_debug_sort_pairs on the same pairs -- bit-exact,
like every integer path here.  Two calls in a row: the second is smaller, uses other key bits and another pass count,
so anything the first call left in the shared scratch (status words, tile counter, histograms) would show.

Collected last on purpose.
"""
import numpy as np
import pytest

from bfqzip_b200 import lib

pytestmark = pytest.mark.gpu


def _pairs(n, bits, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << 63, size=n, dtype=np.uint64)
    if bits < 64:
        keys &= np.uint64((1 << bits) - 1)
    keys[: n // 2] = keys[rng.integers(0, 50, size=n // 2)]          # many duplicates: stability matters
    return keys, np.arange(n, dtype=np.uint32)


def _noalloc(keys, vals, bits):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n = len(keys)
    ka = torch.from_numpy(keys.view(np.int64)).to(dev); va = torch.from_numpy(vals.view(np.int32)).to(dev)
    kb = torch.empty_like(ka); vb = torch.empty_like(va)
    stream = torch.cuda.current_stream(dev)
    flag = lib.sort_pairs_noalloc(ka, kb, va, vb, n, bits, stream.cuda_stream)
    stream.synchronize()
    k, v = (kb, vb) if flag else (ka, va)
    return k.cpu().numpy().view(np.uint64), v.cpu().numpy().view(np.uint32)


def test_noalloc_sort_matches_context_sort_twice():
    g = lib.BfqGpu(2, 0, 0)
    try:
        for n, bits, seed in [(3_000_001, 48, 1), (70_001, 13, 2)]:
            keys, vals = _pairs(n, bits, seed)
            k0, v0 = g.debug_sort_pairs(keys, vals, bits)            # existing entry point (arena scratch)
            k1, v1 = _noalloc(keys, vals, bits)
            assert np.array_equal(k1, k0) and np.array_equal(v1, v0)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(v1, vals[order])
    finally:
        g.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    t = torch.zeros(16, dtype=torch.int64, device="cuda:0")
    with pytest.raises(lib.BfqError) as ei:
        lib.sort_pairs_noalloc(t, t, t, t, lib.NOALLOC_MAX_PAIRS + 1, 8)
    assert ei.value.code == -8 and "at most" in str(ei.value)
