"""CPU tests of the path-tracing half of the oracle: the reference's own unit-test properties for the sampler
(unittest.cpp:66-132), pinning of every host-runnable piece against the reference itself (oracle/_ref), and sanity
of the device-only restatements (SAT build, renderPT)."""
import numpy as np
import pytest

from scenes import procedural_hdri, scene_params


def _murmur3_x86_32(data: bytes, seed: int) -> int:
    """smhasher MurmurHash3_x86_32 (the comparator of unittest.cpp:106-132), written from the published algorithm."""
    M = 0xFFFFFFFF
    h = seed
    n = len(data) // 4
    for i in range(n):
        k = int.from_bytes(data[4 * i:4 * i + 4], "little")
        k = (k * 0xcc9e2d51) & M
        k = ((k << 15) | (k >> 17)) & M
        k = (k * 0x1b873593) & M
        h ^= k
        h = ((h << 13) | (h >> 19)) & M
        h = (h * 5 + 0xe6546b64) & M
    h ^= len(data)
    h ^= h >> 16
    h = (h * 0x85ebca6b) & M
    h ^= h >> 13
    h = (h * 0xc2b2ae35) & M
    h ^= h >> 16
    return h


def test_murmur_compatibility(oracle):
    rng = np.random.default_rng(1)
    for _ in range(300):
        n = int(rng.integers(0, 6))
        v = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        seed = int(rng.integers(0, 2**32))
        assert oracle.murmur(seed, v) == _murmur3_x86_32(v.tobytes(), seed)


def test_owen_scramble_properties(oracle):
    """unittest.cpp:66-104: nested_uniform_scramble is a bijection on the top bits; scramble_f32 stays in [0,1)."""
    L = oracle.lib
    rng = np.random.default_rng(2)
    for _ in range(20):
        seed = int(rng.integers(0, 2**32))
        outs = {L.orc_nested_uniform_scramble(i, seed) & 63 for i in range(64)}
        assert len(outs) == 64
        for x in rng.random(50, dtype=np.float32):
            y = L.orc_scramble_f32(float(x), seed)
            assert 0.0 <= y < 1.0


def test_pmj_table_is_stratified(oracle):
    t = oracle.pmj_table().reshape(128, 4096, 2)
    assert t.min() >= 0.0 and t.max() < 1.0
    for seq in (0, 7, 127):
        for n in (16, 256, 4096):
            s = t[seq, :n]
            # (0,2)-sequence: every elementary interval n x 1 and 1 x n holds exactly one sample
            assert np.unique((s[:, 0] * n).astype(int)).size == n
            assert np.unique((s[:, 1] * n).astype(int)).size == n


def test_pt_pieces_equal_reference(oracle, ref_oracle):
    rng = np.random.default_rng(3)
    for _ in range(100):
        seed = int(rng.integers(0, 2**32))
        v = rng.integers(0, 2**32, 3, dtype=np.uint64).astype(np.uint32)
        assert oracle.murmur(seed, v) == ref_oracle.murmur(seed, v)
        x = int(rng.integers(0, 2**32))
        assert oracle.lib.orc_nested_uniform_scramble(x, seed) == ref_oracle.lib.orc_nested_uniform_scramble(x, seed)
        f = float(rng.random(dtype=np.float32))
        assert oracle.lib.orc_scramble_f32(f, seed) == ref_oracle.lib.orc_scramble_f32(f, seed)
    assert np.array_equal(oracle.pcg32(0, 2525, 1000), ref_oracle.pcg32(0, 2525, 1000))
    assert np.array_equal(oracle.pcg32(77, 123456789, 100), ref_oracle.pcg32(77, 123456789, 100))
    table = oracle.pmj_table()
    assert np.array_equal(table.view(np.uint32), ref_oracle.pmj_table().view(np.uint32))
    for _ in range(2000):
        si, dim, st = (int(rng.integers(0, 2**32)) for _ in range(3))
        if rng.random() < 0.5:
            si, dim = si % 4096, dim % 27
        a = oracle.pmj_sample2d(table, si, dim, st)
        b = ref_oracle.pmj_sample2d(table, si, dim, st)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    normals = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    for _ in range(500):
        a, b = rng.random(2, dtype=np.float32)
        n = normals[int(rng.integers(0, 6))]
        assert np.array_equal(oracle.sample_lambertian(float(a), float(b), n).view(np.uint32),
                              ref_oracle.sample_lambertian(float(a), float(b), n).view(np.uint32))
    w, h = 64, 32
    img = procedural_hdri(w, h)
    sats = oracle.hdri_build_sats(img, w, h)
    for _ in range(2000):
        u = [float(x) for x in rng.random(4, dtype=np.float32)]
        n = normals[int(rng.integers(0, 6))]
        aa = bool(rng.integers(0, 2))
        d0, l0, p0, _ = oracle.hdri_importance_sample(img, sats, w, h, 1.75, n, aa, u)
        d1, l1, p1, _ = ref_oracle.hdri_importance_sample(img, sats, w, h, 1.75, n, aa, u)
        assert np.array_equal(d0.view(np.uint32), d1.view(np.uint32))
        assert np.array_equal(l0.view(np.uint32), l1.view(np.uint32))
        assert np.float32(p0).view(np.uint32) == np.float32(p1).view(np.uint32)


def test_sats_are_monotone_and_normalised(oracle):
    w, h = 96, 40
    img = procedural_hdri(w, h)
    sats = oracle.hdri_build_sats(img, w, h).reshape(7, h, w)
    for k in range(7):
        s = sats[k].astype(np.int64)
        assert s[-1, -1] == 0xFFFFFFFF
        assert (np.diff(s, axis=1) >= 0).all() and (np.diff(s, axis=0) >= 0).all()
    # the +y table puts no mass on the lower hemisphere, the -y table none on the upper one
    assert sats[3][h - 1, w - 1] - sats[3][h // 2, w - 1] == 0
    assert sats[4][h // 2 - 2, w - 1] == 0
    # sampling frequency follows the table: texel counts ~ luminance * solid angle
    rng = np.random.default_rng(5)
    hist = np.zeros((h, w))
    N = 20000
    for _ in range(N):
        u = [float(x) for x in rng.random(4, dtype=np.float32)]
        _, _, _, xy = oracle.hdri_importance_sample(img, sats.reshape(-1), w, h, 1.75, (0, 1, 0), False, u)
        hist[xy[1], xy[0]] += 1
    lum = 0.2126 * img[..., 0] + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]
    th = (np.arange(h) + 0.5) / h * np.pi
    target = lum * np.sin(th)[:, None]
    target /= target.sum()
    # coarse 4x4 binning keeps the statistical noise below the tolerance
    b = lambda a: a[: h // 4 * 4, : w // 4 * 4].reshape(h // 4, 4, w // 4, 4).sum((1, 3))
    assert np.abs(b(hist / N) - b(target)).max() < 0.02


def test_pt_render_small(oracle, bunny):
    """renderPT restatement: deterministic, accumulates 16 spp per step, converges towards itself, resolve in range."""
    tris, col, emi = bunny
    grid, W, H = 32, 48, 32
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    tree = oracle.build_octree(mk, origin, float(dps), grid)
    cam = oracle.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H, focus=6.0, lens_r=0.02)
    hw, hh = 64, 32
    img = procedural_hdri(hw, hh)
    sats = oracle.hdri_build_sats(img, hw, hh)
    pmj = oracle.pmj_table()
    has_em = int(me[:, :3].any())
    assert has_em == 1
    acc = np.zeros(W * H * 4, np.float32)
    rays0 = oracle.pt_render(tree, mc, me, has_em, cam, W, H, 0, pmj, acc, img, sats, (hw, hh), max_depth=4, threads=4)
    acc1 = np.zeros(W * H * 4, np.float32)
    rays1 = oracle.pt_render(tree, mc, me, has_em, cam, W, H, 0, pmj, acc1, img, sats, (hw, hh), max_depth=4, threads=1)
    assert rays0 == rays1 and np.array_equal(acc.view(np.uint32), acc1.view(np.uint32))  # thread count does not matter
    a = acc.reshape(H, W, 4)
    assert (a[..., 3] == 16).all() and np.isfinite(a).all() and (a[..., :3] >= 0).all()
    assert rays0 > W * H * 16
    oracle.pt_render(tree, mc, me, has_em, cam, W, H, 1, pmj, acc, img, sats, (hw, hh), max_depth=4, threads=4)
    a2 = acc.reshape(H, W, 4)
    assert (a2[..., 3] == 32).all()
    m16 = acc1.reshape(H, W, 4)[..., :3] / 16
    m32 = a2[..., :3] / 32
    assert np.sqrt(((m16 - m32) ** 2).mean()) < 0.25 * max(m32.mean(), 1e-3) + 0.2
    rgba = oracle.pt_resolve(acc).reshape(H, W, 4)
    assert (rgba[..., 3] == 255).all() and rgba[..., :3].max() > 0
    # rows [y0,y1) only touch their own pixels
    acc3 = np.zeros(W * H * 4, np.float32)
    oracle.pt_render(tree, mc, me, has_em, cam, W, H, 0, pmj, acc3, img, sats, (hw, hh), max_depth=4, y0=8, y1=16, threads=2)
    a3 = acc3.reshape(H, W, 4)
    assert (a3[:8] == 0).all() and (a3[16:] == 0).all()
    assert np.array_equal(a3[8:16].view(np.uint32), acc1.reshape(H, W, 4)[8:16].view(np.uint32))
