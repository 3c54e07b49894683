"""hako_unique_merge_noalloc (include/hako.h): the merge for callers that may not allocate.  It has no counterpart in the reference;
its scratch is a static device array, so it is checked here -- through the C ABI -- against
the existing entry point hako_unique_merge on the same input, bit for bit (integer work: tolerance 0), on two calls in a row (the array
keeps the first call's contents) and with other calls in between.  The file is named so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ITEMS = 1 << 24      # HAKO_UNIQUE_NOALLOC_MAX_ITEMS


@pytest.fixture(scope="module")
def hk():
    import massivevoxelraytracing_b200 as h
    return h


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("needs a GPU")
    return t


def _pairs(n, distinct, seed):
    """sorted keys with runs of every length and random packed attributes (colour | emission, 8 bytes)"""
    rng = np.random.default_rng(seed)
    keys = np.sort(rng.integers(0, max(distinct, 1), n, dtype=np.uint64)) * np.uint64(3) + np.uint64(5)
    attrs = rng.integers(0, 256, (n, 8), dtype=np.uint8)
    if seed % 2:
        attrs[:, 4:] = 0                      # no emission anywhere: hasEmission must read 0
    return keys, attrs


def _existing(hk, ctx, keys, attrs):
    """the existing entry point: hako_unique_merge (host arrays in place)"""
    k = keys.copy()
    a = attrs.copy()
    U, he = C.c_uint64(), C.c_uint32()
    rc = ctx.lib.hako_unique_merge(ctx.h, None, k.ctypes.data, a.ctypes.data, k.size, C.byref(U), C.byref(he))
    assert rc == 0
    return k[:U.value], a[:U.value], int(he.value)


def _noalloc(torch, ctx, keys, attrs, bufs=None):
    """the new entry point: device pointers only; every buffer is created by the CALLER before the call"""
    n = keys.size
    dev = torch.device("cuda", 0)
    if bufs is None:
        bufs = (torch.empty(max(n, 2), dtype=torch.int64, device=dev), torch.empty((max(n, 2), 8), dtype=torch.uint8, device=dev),
                torch.full((max(n, 2),), -1, dtype=torch.int64, device=dev), torch.full((max(n, 2), 8), 255, dtype=torch.uint8, device=dev),
                torch.full((2,), -1, dtype=torch.int64, device=dev))
    k_d, a_d, ko_d, ao_d, cnt_d = bufs
    k_d[:n].copy_(torch.from_numpy(keys.view(np.int64)))
    a_d[:n].copy_(torch.from_numpy(attrs))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = ctx.lib.hako_unique_merge_noalloc(ctx.h, None, k_d.data_ptr(), a_d.data_ptr(), n, ko_d.data_ptr(), ao_d.data_ptr(), cnt_d.data_ptr())
    assert rc == 0, ctx.lib.hako_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    cnt = cnt_d.cpu().numpy().view(np.uint64)
    U = int(cnt[0])
    return ko_d[:U].cpu().numpy().view(np.uint64), ao_d[:U].cpu().numpy(), int(cnt[1] != 0), bufs


def test_noalloc_merge_equals_existing_entry_point_twice_in_a_row(hk, torch):
    ctx = hk.Context(0)
    # two calls in a row of DIFFERENT sizes (the second smaller: stale descriptors of the first lie behind its own), then the first again
    cases = [_pairs(1_000_003, 300_000, 1), _pairs(70_001, 50, 2), _pairs(1_000_003, 300_000, 1), _pairs(2049, 2049 * 4, 3), _pairs(1, 1, 4)]
    for keys, attrs in cases:
        ek, ea, ehe = _existing(hk, ctx, keys, attrs)
        for rep in range(2):
            gk, ga, ghe, _ = _noalloc(torch, ctx, keys, attrs)
            assert gk.size == ek.size
            np.testing.assert_array_equal(gk, ek)
            np.testing.assert_array_equal(ga, ea)
            assert ghe == ehe
    # caller-owned buffers reused across calls, as a serving loop would
    bufs = None
    for keys, attrs in (cases[1], cases[3], cases[1]):
        ek, ea, ehe = _existing(hk, ctx, keys, attrs)
        gk, ga, ghe, bufs = _noalloc(torch, ctx, keys, attrs, bufs if bufs and bufs[0].numel() >= keys.size else None)
        np.testing.assert_array_equal(gk, ek)
        np.testing.assert_array_equal(ga, ea)
        assert ghe == ehe
    ctx.close()


def test_noalloc_merge_limits_and_errors(hk, torch):
    ctx = hk.Context(0)
    dev = torch.device("cuda", 0)
    buf = torch.zeros(64, dtype=torch.int64, device=dev)
    cnt = torch.full((2,), 7, dtype=torch.int64, device=dev)
    p = buf.data_ptr()
    f = ctx.lib.hako_unique_merge_noalloc
    # beyond the static scratch: a clear error, nothing launched
    assert f(ctx.h, None, p, p, MAX_ITEMS + 1, p, p, cnt.data_ptr()) != 0
    msg = ctx.lib.hako_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "HAKO_UNIQUE_NOALLOC_MAX_ITEMS" in msg
    # null / misaligned arguments
    assert f(ctx.h, None, None, p, 4, p, p, cnt.data_ptr()) != 0
    assert f(ctx.h, None, p + 8, p, 4, p, p, cnt.data_ptr()) != 0
    assert f(ctx.h, None, p, p, 4, p, p, None) != 0
    torch.cuda.synchronize()
    assert cnt.cpu().tolist() == [7, 7]
    # n == 0: counts cleared, nothing else touched
    assert f(ctx.h, None, None, None, 0, None, None, cnt.data_ptr()) == 0
    torch.cuda.synchronize()
    assert cnt.cpu().tolist() == [0, 0]
    ctx.close()
