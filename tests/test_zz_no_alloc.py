"""b2rt_sort_pairs_noalloc (-m gpu): the radix sort with its scratch in the library's static device array, for callers that
cannot call cudaMalloc.  No counterpart in the reference.  Checked through the
C ABI against the existing entry point b2rt_sort_pairs on the same inputs, two calls in a row on the shared scratch; integer
work, so the bar is bit-exact.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from cuda_raytrace_b200 import b2rt

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _pairs(rng, n, bits_):
    eff = 8 * ((bits_ + 7) // 8)
    keys = rng.integers(0, 2 ** eff, n, dtype=np.uint64).astype(np.uint32)
    if n > 1000:
        keys[: n // 3] = keys[n // 2: n // 2 + n // 3]      # duplicates: stability matters
    return keys, np.arange(n, dtype=np.uint32)


def _both(c, keys, vals, bits_):
    """(existing entry point, no-alloc entry point) on copies of the same device arrays"""
    n = len(keys)
    k0, v0 = _dev(keys.view(np.int32)), _dev(vals.view(np.int32))
    k1, v1 = _dev(keys.view(np.int32)), _dev(vals.view(np.int32))
    torch.cuda.synchronize()
    c.sort_pairs_device(n, k0.data_ptr(), v0.data_ptr(), bits_)
    c.sort_pairs_noalloc_device(n, k1.data_ptr(), v1.data_ptr(), bits_)
    c.sync()
    return (k0.cpu().numpy().view(np.uint32), v0.cpu().numpy().view(np.uint32),
            k1.cpu().numpy().view(np.uint32), v1.cpu().numpy().view(np.uint32))


@pytest.mark.parametrize("variant", ["1", "0"])       # 1 = the default (one-sweep below 2^22 pairs), 0 = three launches per pass
def test_noalloc_sort_equals_sort_pairs_on_consecutive_calls(built, monkeypatch, variant):
    monkeypatch.setenv("B2RT_SORT_VARIANT", variant)
    rng = np.random.default_rng(11)
    with b2rt.Context(0, seed=1) as c:
        c.sort_pairs_noalloc_device(0, 0, 0, 32)         # the once-per-device set-up on its own
        # back-to-back calls on the shared scratch: the largest allowed size first, then smaller / ragged ones that find the
        # previous call's histograms, look-back states and ping-pong data in the array; odd pass counts end in the scratch
        for n, bits_ in [(b2rt.SORT_NOALLOC_MAX, 32), (300_001, 24), (1_000_003, 26), (4097, 8), (2, 32), (1, 32), (70_000, 16),
                         (b2rt.SORT_NOALLOC_MAX, 32)]:
            keys, vals = _pairs(rng, n, bits_)
            k0, v0, k1, v1 = _both(c, keys, vals, bits_)
            assert np.array_equal(k1, k0) and np.array_equal(v1, v0), (n, bits_, variant)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(v1, order.astype(np.uint32)), (n, bits_, variant)
        # all keys equal: one digit chain through every tile of the look-back, twice in a row
        for _ in range(2):
            keys, vals = np.full(500_000, 0xdeadbeef, np.uint32), np.arange(500_000, dtype=np.uint32)
            k0, v0, k1, v1 = _both(c, keys, vals, 32)
            assert np.array_equal(k1, k0) and np.array_equal(v1, vals)


def test_noalloc_sort_rejects_sizes_beyond_the_static_scratch(built):
    with b2rt.Context(0, seed=1) as c:
        n = b2rt.SORT_NOALLOC_MAX + 1
        k, v = torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros(n, dtype=torch.int32, device="cuda")
        with pytest.raises(b2rt.B2rtError) as e:
            c.sort_pairs_noalloc_device(n, k.data_ptr(), v.data_ptr(), 32)
        assert e.value.code == -1 and "B2RT_SORT_NOALLOC_MAX" in str(e.value)      # B2RT_E_ARG, names the limit
        # the context stays usable
        keys, vals = _pairs(np.random.default_rng(3), 10_000, 32)
        k0, v0, k1, v1 = _both(c, keys, vals, 32)
        assert np.array_equal(k1, k0) and np.array_equal(v1, v0)


def test_noalloc_sort_allocates_no_device_memory(built):
    """cudaMemGetInfo before and after: the call leaves the free device memory where it was (a context's first
    b2rt_sort_pairs, by contrast, reserves its ping-pong arrays)"""
    with b2rt.Context(0, seed=1) as c:
        n = 1 << 20
        keys, vals = _pairs(np.random.default_rng(5), n, 32)
        k, v = _dev(keys.view(np.int32)), _dev(vals.view(np.int32))
        w = torch.arange(5000, dtype=torch.int32, device="cuda")
        c.sort_pairs_noalloc_device(5000, w.data_ptr(), w.clone().data_ptr(), 32)     # set-up + the kernels' first (lazy) load
        c.sync()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        c.sort_pairs_noalloc_device(n, k.data_ptr(), v.data_ptr(), 32)
        c.sync()
        free1 = torch.cuda.mem_get_info()[0]
        assert free1 == free0
        assert np.array_equal(v.cpu().numpy().view(np.uint32), np.argsort(keys, kind="stable").astype(np.uint32))
