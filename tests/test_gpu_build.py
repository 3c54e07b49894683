"""GPU parity of the build path (voxelize -> sort -> unique -> octree) against the CPU oracle, stage by stage and end to end.
Everything here is integer / byte work: the bar is bit-exact."""
import numpy as np
import pytest

from scenes import scene_params, subdivide, vertex_colors

pytestmark = pytest.mark.gpu
FLT_MAX = np.float32(3.402823466e38)


@pytest.fixture(scope="module")
def hk():
    import massivevoxelraytracing_b200 as h
    return h


@pytest.fixture(scope="module")
def ctx(hk):
    c = hk.Context(0)
    yield c
    c.close()


def _attrs_to_arrays(a):
    return a["color"], a["emission"]


def _oracle_stages(oracle, tris, col, emi, grid):
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    tree = oracle.build_octree(mk, origin, float(dps), grid)
    return origin, dps, keys, c, e, mk, mc, me, tree


# ---------------------------------------------------------------- voxelize
def _canon(keys, color, emission):
    """the dumped voxels as a sorted multiset of (key, colour, emission) records: the emission order is unspecified
    (include/hako.h; the reference's own is an atomicAdd race, voxKernel.cu:121)"""
    rec = np.empty(keys.size, np.dtype([("k", np.uint64), ("c", np.uint32), ("e", np.uint32)]))
    rec["k"] = keys
    rec["c"] = np.ascontiguousarray(color).view(np.uint32).reshape(-1)
    rec["e"] = np.ascontiguousarray(emission).view(np.uint32).reshape(-1)
    return np.sort(rec, order=("k", "c", "e"))


def _assert_same_voxels(gk, ga, keys, c, e):
    assert gk.size == keys.size
    assert np.array_equal(_canon(gk, ga["color"], ga["emission"]), _canon(keys, c, e))


@pytest.mark.parametrize("grid", [16, 64, 256, 2048])
def test_voxelize_matches_oracle(hk, ctx, oracle, bunny, grid):
    """16..256: every bunny triangle takes the warp-cooperative path (32 triangles per warp); 2048: a mix of that and the
    per-column path of large triangles"""
    tris, col, emi = bunny
    if grid == 2048:
        tris, col, emi = tris[:600], col[:600], emi[:600]
    origin, dps = scene_params(bunny[0], grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    gk, ga = ctx.voxelize(tris, col, emi, origin, float(dps), grid)
    _assert_same_voxels(gk, ga, keys, c, e)
    if grid <= 256:
        # small triangles only: the emission order is (triangle, x, y, z) == the CPU loop's order (voxRT.cpp:198-240)
        np.testing.assert_array_equal(gk, keys)
        np.testing.assert_array_equal(ga["color"], c)
        np.testing.assert_array_equal(ga["emission"], e)


def test_voxelize_paths_agree(hk, oracle, bunny, monkeypatch):
    """every execution path of the voxelizer produces the same dumped voxels, and they equal the oracle's.  Ragged triangle
    counts exercise the partial last warp."""
    tris, col, emi = bunny
    origin, dps = scene_params(tris, 512)
    out = []
    # default (record path: the counting pass keeps the voxels, the emit pass only shades) | per-column kernels only | two
    # enumerations (no records) | record slabs of 32 voxels per warp, so that most warps overflow and take the hybrid path
    for extra in ({}, {"HAKO_VOX_COLUMNS_ONLY": "1"}, {"HAKO_VOX_RECORD": "0"}, {"HAKO_VOX_SLAB_CAP": "32"}):
        import subprocess, sys, os, json
        code = ("import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
                "import massivevoxelraytracing_b200 as hk\nfrom scenes import bunny_scene, scene_params\n"
                "t, c, e = bunny_scene(); o, d = scene_params(t, 512); ctx = hk.Context(0)\n"
                "import hashlib\n"
                "for n in (1, 31, 33, 1000, t.shape[0]):\n"
                "    k, a = ctx.voxelize(t[:n], c[:n], e[:n], o, float(d), 512)\n"
                "    r = np.empty(k.size, np.dtype([('k', np.uint64), ('a', np.uint64)])); r['k'] = k; r['a'] = a.view(np.uint64).reshape(-1)\n"
                "    print(n, k.size, hashlib.sha1(np.sort(r, order=('k', 'a')).tobytes()).hexdigest())\n") % (
                    os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ, **extra))
        assert r.returncode == 0, r.stderr
        out.append(r.stdout)
    assert out[0] == out[1] == out[2] == out[3] and len(out[0].splitlines()) == 5
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), 512)
    assert int(out[0].splitlines()[-1].split()[1]) == keys.size


def test_voxelize_conservative_and_degenerate(hk, ctx, oracle, bunny, small_cases):
    tris, col, emi = bunny
    origin, dps = scene_params(tris, 64)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), 64, six=False)
    gk, ga = ctx.voxelize(tris, col, emi, origin, float(dps), 64, sixSeparating=False)
    _assert_same_voxels(gk, ga, keys, c, e)
    deg = small_cases["deg_tris"]
    for grid in (8, 32):
        for six in (1, 0):
            gk, _ = ctx.voxelize(deg, np.ones_like(deg), np.zeros_like(deg), np.zeros(3, np.float32), 1.0 / grid, grid, sixSeparating=bool(six))
            np.testing.assert_array_equal(np.sort(gk), small_cases["deg_keys_%d_%d" % (grid, six)])
    # empty input
    gk, ga = ctx.voxelize(np.zeros((0, 3, 3), np.float32), np.zeros((0, 3, 3), np.float32), np.zeros((0, 3, 3), np.float32), np.zeros(3, np.float32), 1.0, 8)
    assert gk.size == 0 and ga.size == 0


def test_voxelize_out_of_range_colours(hk, ctx, oracle):
    """(uint8_t)(c * 255 + 0.5) with colours outside [0,1] and huge values: x86 cast semantics"""
    rng = np.random.default_rng(5)
    tris = rng.random((64, 3, 3)).astype(np.float32)
    col = (rng.standard_normal((64, 3, 3)) * 3).astype(np.float32)
    emi = (rng.standard_normal((64, 3, 3)) * 1e12).astype(np.float32)
    keys, c, e = oracle.voxelize(tris, col, emi, np.zeros(3, np.float32), 1.0 / 32, 32)
    gk, ga = ctx.voxelize(tris, col, emi, np.zeros(3, np.float32), 1.0 / 32, 32)
    _assert_same_voxels(gk, ga, keys, c, e)


def test_voxelize_zero_emission_triangles(hk, ctx, oracle):
    """the shading pass skips the emission interpolation of a triangle whose nine emission values are all +-0; the bytes must still be
    the reference's for degenerate triangles (NaN / infinite barycentrics), negative zeros, and emissive triangles in the same warp"""
    rng = np.random.default_rng(6)
    T = 96
    tris = rng.random((T, 3, 3)).astype(np.float32)
    tris[::7, 1] = tris[::7, 0]                        # two equal vertices: zero area
    tris[3::11, 1] = tris[3::11, 0]
    tris[3::11, 2] = tris[3::11, 0]                    # a point
    tris[5::13, 2] = (tris[5::13, 0] + tris[5::13, 1]) * np.float32(0.5)   # collinear
    col = rng.random((T, 3, 3)).astype(np.float32)
    emi = np.zeros((T, 3, 3), np.float32)
    emi[::3] = rng.random((len(emi[::3]), 3, 3)).astype(np.float32)        # every third triangle emits
    emi[1::6] = np.float32(-0.0)                       # negative zeros count as zero
    emi[2::9, 2, 1] = np.float32(1e-30)                # one tiny non-zero value: not the fast path
    origin = np.zeros(3, np.float32)
    for grid in (32, 128):
        keys, c, e = oracle.voxelize(tris, col, emi, origin, 1.0 / grid, grid)
        gk, ga = ctx.voxelize(tris, col, emi, origin, 1.0 / grid, grid)
        _assert_same_voxels(gk, ga, keys, c, e)


# ---------------------------------------------------------------- sort
@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 100000, 1 << 20])
def test_sort_pairs_stable(hk, ctx, n):
    """tinyhipradixsort unittest.cpp:356-424,426-488 (SortPairs K64V64): result == std::stable_sort, stability asserted"""
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    if n > 100:
        keys[: n // 2] = keys[n // 2: n // 2 + n // 2] & np.uint64(0xFFFF)     # many duplicates
    vals = np.arange(n, dtype=np.uint64)
    k, v = ctx.sort_pairs(keys, vals)
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(k, keys[order])
    np.testing.assert_array_equal(v, vals[order])
    k2 = ctx.sort_pairs(keys)
    np.testing.assert_array_equal(k2, keys[order])


@pytest.mark.parametrize("begin,end", [(0, 8), (0, 24), (8, 16), (3, 30), (0, 40), (16, 64), (0, 63)])
def test_sort_bit_window(hk, ctx, begin, end):
    """unittest.cpp:248-355 StartBits: only bits [begin, end) take part; the rest of the key rides along, stably"""
    rng = np.random.default_rng(begin * 64 + end)
    n = 50000
    keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    vals = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    k, v = ctx.sort_pairs(keys, vals, begin, end)
    sub = (keys >> np.uint64(begin)) & np.uint64((1 << (end - begin)) - 1)
    order = np.argsort(sub, kind="stable")
    np.testing.assert_array_equal(k, keys[order])
    np.testing.assert_array_equal(v, vals[order])


def test_sort_extreme(hk, ctx):
    """unittest.cpp extremeCase: all-equal keys, already sorted, reversed"""
    n = 70000
    for keys in (np.full(n, 0xFFFFFFFFFFFFFFFF, np.uint64), np.zeros(n, np.uint64), np.arange(n, dtype=np.uint64), np.arange(n, dtype=np.uint64)[::-1].copy()):
        vals = np.arange(n, dtype=np.uint64)
        k, v = ctx.sort_pairs(keys, vals)
        order = np.argsort(keys, kind="stable")
        np.testing.assert_array_equal(k, keys[order])
        np.testing.assert_array_equal(v, vals[order])


def test_sort_variants_agree():
    """every selectable form of the onesweep pass (bulk-copy tiles with 32- or 64-bit look-back descriptors, ballot or match.any ranking,
    the 512 x 8 and 256 x 16 tilings, the thread-block-cluster pass with 2 / 8 tiles per cluster, the first-generation kernel) is the same
    stable sort: pairs, keys only, ragged and odd lengths, 9-bit (27- and 36-bit windows) and 8-bit digits (16-bit window)"""
    import subprocess, sys, os
    code = ("import sys, hashlib, numpy as np; sys.path.insert(0, %r)\n"
            "import massivevoxelraytracing_b200 as hk\nctx = hk.Context(0)\n"
            "for n, bits in ((0, 27), (1, 27), (3839, 27), (3841, 36), (100001, 27), (1 << 20, 36), (250007, 16)):\n"
            "    rng = np.random.default_rng(n + bits)\n"
            "    keys = rng.integers(0, 1 << bits, n, dtype=np.uint64) | (rng.integers(0, 1 << 20, n, dtype=np.uint64) << np.uint64(40))\n"
            "    if n > 100: keys[: n // 2] &= np.uint64(0xFF)\n"
            "    vals = np.arange(n, dtype=np.uint64)\n"
            "    k, v = ctx.sort_pairs(keys, vals, 0, bits)\n"
            "    k2 = ctx.sort_pairs(keys, None, 0, bits)\n"
            "    order = np.argsort(keys & np.uint64((1 << bits) - 1), kind='stable')\n"
            "    ok = bool(np.array_equal(k, keys[order]) and np.array_equal(v, vals[order]) and np.array_equal(k2, keys[order]))\n"
            "    print(n, bits, ok, hashlib.sha1(k.tobytes() + v.tobytes()).hexdigest())\n") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    outs = []
    for extra in ({}, {"HAKO_SORT_D32": "0"}, {"HAKO_SORT_RANK": "0"}, {"HAKO_SORT_VARIANT": "2"}, {"HAKO_SORT_VARIANT": "3"}, {"HAKO_SORT_VARIANT": "0"},
                  {"HAKO_SORT_VARIANT": "4", "HAKO_SORT_CLUSTER": "2"}, {"HAKO_SORT_VARIANT": "4", "HAKO_SORT_CLUSTER": "8"},
                  {"HAKO_SORT_VARIANT": "4", "HAKO_SORT_CLUSTER": "4", "HAKO_SORT_D32": "0"}):
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ, **extra))
        assert r.returncode == 0, (extra, r.stderr[-2000:])
        outs.append(r.stdout)
        assert all(line.split()[2] == "True" for line in r.stdout.splitlines()), (extra, r.stdout)
    assert all(o == outs[0] for o in outs) and len(outs[0].splitlines()) == 7


# ---------------------------------------------------------------- unique
@pytest.mark.parametrize("grid", [16, 64, 256])
def test_unique_merge_matches_oracle(hk, ctx, oracle, bunny, grid):
    tris, col, emi = bunny
    origin, dps, keys, c, e, mk, mc, me, tree = _oracle_stages(oracle, tris, col, emi, grid)
    attrs = np.zeros(keys.size, hk.VOXEL_ATTRIB)
    attrs["color"], attrs["emission"] = c, e
    sk, sv = ctx.sort_pairs(keys, attrs.view(np.uint64), 0, 24 if grid <= 256 else 32)
    uk, ua, he = ctx.unique_merge(sk, sv.view(hk.VOXEL_ATTRIB))
    np.testing.assert_array_equal(uk, mk)
    np.testing.assert_array_equal(ua["color"], mc)          # includes alpha = n / n = 1 as mergeVoxels gives (voxRT.cpp:44-47)
    np.testing.assert_array_equal(ua["emission"], me)
    assert he == bool(me[:, :3].any())


def test_unique_long_runs(hk, ctx, oracle):
    rng = np.random.default_rng(9)
    keys = np.sort(rng.integers(0, 50, 20000, dtype=np.uint64))
    attrs = np.zeros(keys.size, hk.VOXEL_ATTRIB)
    attrs["color"] = rng.integers(0, 256, (keys.size, 4), dtype=np.uint8)
    attrs["emission"] = 0
    mk, mc, me = oracle.merge_voxels(keys, attrs["color"], attrs["emission"])
    uk, ua, he = ctx.unique_merge(keys, attrs)
    np.testing.assert_array_equal(uk, mk)
    np.testing.assert_array_equal(ua["color"][:, :3], mc[:, :3])
    assert not he
    uk, ua, he = ctx.unique_merge(np.zeros(0, np.uint64), np.zeros(0, hk.VOXEL_ATTRIB))
    assert uk.size == 0


def test_unique_integer_means_are_exact_for_every_run_length(hk, ctx, oracle):
    """the integer means of a run (voxRT.cpp:38-47) for run lengths 1..300, around the 4 096 boundary below which the kernel divides by
    multiplying with ceil(2^32 / count), and far above it; channel sums at their extremes (all 255, all 0, one 255)"""
    rng = np.random.default_rng(10)
    lengths = list(range(1, 301)) + [4095, 4096, 4097, 4103, 4104, 9001, 70000]
    keys = np.repeat(np.arange(len(lengths), dtype=np.uint64) * 3 + 7, lengths)
    attrs = np.zeros(keys.size, hk.VOXEL_ATTRIB)
    attrs["color"] = rng.integers(0, 256, (keys.size, 4), dtype=np.uint8)
    attrs["emission"] = rng.integers(0, 256, (keys.size, 4), dtype=np.uint8)
    start = np.concatenate([[0], np.cumsum(lengths)])
    for i in range(0, len(lengths), 3):         # every third run saturated: sum = 255 * count, the largest dividend
        attrs["color"][start[i]:start[i + 1]] = 255
        attrs["emission"][start[i]:start[i + 1]] = 255
    for i in range(1, len(lengths), 3):         # one 255 among zeros: quotient 0 unless the run has one voxel ... 255 / count
        attrs["color"][start[i]:start[i + 1]] = 0
        attrs["color"][start[i]] = 255
    mk, mc, me = oracle.merge_voxels(keys, attrs["color"], attrs["emission"])
    uk, ua, he = ctx.unique_merge(keys, attrs)
    np.testing.assert_array_equal(uk, mk)
    np.testing.assert_array_equal(ua["color"][:, :3], mc[:, :3])
    np.testing.assert_array_equal(ua["emission"][:, :3], me[:, :3])
    assert he


# ---------------------------------------------------------------- octree
def _cmp_nodes(a, b):
    assert a.size == b.size
    for f in ("mask", "children", "psum"):
        np.testing.assert_array_equal(a[f], b[f], err_msg=f)


@pytest.mark.parametrize("grid", [2, 4, 16, 64, 256])
def test_octree_from_morton_matches_oracle(hk, ctx, oracle, bunny, grid):
    tris, col, emi = bunny
    origin, dps, keys, c, e, mk, mc, me, tree = _oracle_stages(oracle, tris, col, emi, grid)
    g = hk.IntersectorOctree(ctx).buildDAGReference(mk, origin, float(dps), grid)
    nodes, _, _ = g.export()
    _cmp_nodes(nodes, tree.nodes())                       # byte-identical node array incl. the reference's numbering
    lo, up = tree.bounds()
    np.testing.assert_array_equal(g.m_lower, lo)
    np.testing.assert_array_equal(g.m_upper, up)
    naive = oracle.build_octree(mk, origin, float(dps), grid, dag=False)
    g2 = hk.IntersectorOctree(ctx).build(mk, origin, float(dps), grid)
    nodes2, _, _ = g2.export()
    _cmp_nodes(nodes2, naive.nodes())


def test_octree_random_sparse_and_single(hk, ctx, oracle):
    rng = np.random.default_rng(11)
    org = np.zeros(3, np.float32)
    for grid, n in ((8, 1), (8, 512), (32, 700), (1024, 5000), (1 << 15, 20000)):
        bits = 3 * int(np.log2(grid))
        m = np.unique(rng.integers(0, 1 << bits, n, dtype=np.uint64)) if n < (1 << bits) else np.arange(1 << bits, dtype=np.uint64)
        tree = oracle.build_octree(m, org, 1.0 / grid, grid)
        g = hk.IntersectorOctree(ctx).buildDAGReference(m, org, 1.0 / grid, grid)
        _cmp_nodes(g.export()[0], tree.nodes())
    # empty voxel list: no nodes, every ray misses
    g = hk.IntersectorOctree(ctx).buildDAGReference(np.zeros(0, np.uint64), org, 1.0 / 8, 8)
    assert g.m_numberOfNodes == 0
    t, _, _ = g.intersect(np.zeros((4, 3), np.float32) - 1, np.ones((4, 3), np.float32))
    assert (t == FLT_MAX).all()


# ---------------------------------------------------------------- end to end
@pytest.mark.parametrize("grid,W,H", [(64, 256, 256), (256, 1024, 1024), (1024, 1920, 1080), (4096, 1920, 1080)])
def test_build_end_to_end_matches_golden(hk, ctx, oracle, bunny, golden, grid, W, H):
    """IntersectorOctreeGPU::build + render == the reference CPU path's golden values"""
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    case = golden["cases"][str(grid)]
    f = lambda a: "%016x" % oracle.fnv1a64(a)
    nodes, mk, attrs = g.export(True, True, True)
    st = ctx.build_stats()
    assert st["dumpedVoxels"] == case["D"] and st["uniqueVoxels"] == case["U"] == mk.size
    assert nodes.size == case["dag_nodes"]
    assert f(mk) == case["fnv_merged_keys"]
    assert f(np.ascontiguousarray(attrs["color"][:, :3])) == case["fnv_merged_rgb"]
    assert f(np.ascontiguousarray(attrs["emission"][:, :3])) == case["fnv_merged_emission_rgb"]
    assert f(np.ascontiguousarray(nodes["mask"])) == case["fnv_node_mask"]
    assert f(np.ascontiguousarray(nodes["children"])) == case["fnv_node_children"]
    assert f(np.ascontiguousarray(nodes["psum"])) == case["fnv_node_psum"]
    assert case["image"] == [W, H]
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    assert bytes(cam).hex() == case["camera_bytes_hex"]
    _, t, nm, vi = g.render(cam, W, H, rgba=False)
    assert int((t != FLT_MAX).sum()) == case["hits"]
    assert f(vi) == case["fnv_vindex"] and f(nm) == case["fnv_nmajor"] and f(t) == case["fnv_t"]
    assert st["hostSyncs"] <= 6 and st["kernelLaunches"] > 0


@pytest.mark.parametrize("grid", [8192, 16384, 32768])
def test_build_large_grid_subset_matches_golden(hk, ctx, oracle, bunny, golden, grid):
    """BASELINE config 4 / 5 grids: 39 / 42 / 45-bit Morton keys, 5 / 6 sort passes, 13-15 octree levels -- bit-exact against the
    reference's CPU code on the first 64 bunny triangles placed in the full mesh's frame"""
    case = golden["subset_cases"][str(grid)]
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    n = case["triangles"]
    g = hk.IntersectorOctreeGPU(ctx).build(tris[:n], col[:n], emi[:n], None, origin, float(dps), grid)
    f = lambda a: "%016x" % oracle.fnv1a64(a)
    nodes, mk, attrs = g.export(True, True, True)
    st = ctx.build_stats()
    assert (st["dumpedVoxels"], st["uniqueVoxels"], nodes.size) == (case["D"], case["U"], case["dag_nodes"])
    bits = 3 * int(np.log2(grid))
    assert st["sortPasses"] == min((bits + 7) // 8, (bits + 8) // 9)        # 9-bit digits when they save a pass
    assert f(mk) == case["fnv_merged_keys"]
    assert f(np.ascontiguousarray(attrs["color"][:, :3])) == case["fnv_merged_rgb"]
    assert f(np.ascontiguousarray(attrs["emission"][:, :3])) == case["fnv_merged_emission_rgb"]
    assert f(np.ascontiguousarray(nodes["mask"])) == case["fnv_node_mask"]
    assert f(np.ascontiguousarray(nodes["children"])) == case["fnv_node_children"]
    assert f(np.ascontiguousarray(nodes["psum"])) == case["fnv_node_psum"]
    W, H = case["image"]
    cam = hk.camera_lookat((-2.2, 1.2, 1.0), tuple(float(x) for x in tris[:n].reshape(-1, 3).mean(0)), width=W, height=H)
    assert bytes(cam).hex() == case["camera_bytes_hex"]
    _, t, nm, vi = g.render(cam, W, H, rgba=False)
    assert int((t != FLT_MAX).sum()) == case["hits"]
    assert f(vi) == case["fnv_vindex"] and f(nm) == case["fnv_nmajor"] and f(t) == case["fnv_t"]
    g.cleanUp()


def test_build_is_independent_of_the_upload_slicing(hk, ctx, monkeypatch):
    """hako_octree_build sends colours / emissions up in triangle ranges and launches the emit pass per range behind its own
    copies (HAKO_BUILD_SLICES): nodes, keys and attributes are the same bytes for 1, 2, 3 and the clamped maximum of ranges,
    and equal to the build from device-resident vertex data (no staging at all)"""
    import torch
    from scenes import dragon_standin
    tris, col, emi = dragon_standin(3, 0.15)          # 327 040 triangles, 35 MB of vertex data: the staged path, up to 4 ranges
    grid = 1024
    origin, dps = scene_params(tris, grid)
    outs = []
    for slices in ("1", "2", "3", "16"):
        monkeypatch.setenv("HAKO_BUILD_SLICES", slices)
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
        outs.append(tuple(a.copy() for a in g.export(True, True, True)))
        g.cleanUp()
    monkeypatch.delenv("HAKO_BUILD_SLICES")
    dev = [torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda() for a in (tris, col, emi)]
    g = hk.IntersectorOctreeGPU(ctx).build_device(dev[0], dev[1], dev[2], None, origin, float(dps), grid)
    outs.append(tuple(a.copy() for a in g.export(True, True, True)))
    g.cleanUp()
    assert outs[0][1].size > 1_000_000 and np.unique(outs[0][2]["color"], axis=0).shape[0] > 1000
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8))


@pytest.mark.parametrize("grid", [2048, 8192, 16384])
def test_build_sweep_full_mesh_invariants(hk, ctx, bunny, golden, grid):
    """BASELINE config 4: the whole bunny at 2^11 / 2^13 / 2^14 (up to 554 M dumped voxels).  No CPU oracle finishes these in
    seconds, so the check is through size-independent properties: the octree's own voxel count (root prefix sums) equals the
    number of unique keys, children precede parents, hit voxel ranks are in range, the silhouette matches the 4096^3 golden
    frame to within the voxel size, and a second build is bit-identical (determinism)."""
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    st = ctx.build_stats()
    D, U, N = st["dumpedVoxels"], st["uniqueVoxels"], st["numberOfNodes"]
    ref = golden["cases"]["4096"]
    scale = (grid / 4096.0) ** 2
    assert 0.8 * scale * ref["U"] < U <= D and D < 1.25 * scale * ref["D"]
    nodes = g.export(True, False, False)[0]
    assert nodes.size == N and N < (1 << 24)
    ch = nodes["children"].astype(np.uint32)
    internal = ch != 0xFFFFFFFF
    idx = (ch & 0xFFFFFF).astype(np.int64)
    own = np.arange(N, dtype=np.int64)[:, None]
    root = nodes[-1]
    # leaf-level nodes have only sentinels; elsewhere every child index is smaller than its parent (bottom-up order)
    assert (idx[internal] < np.broadcast_to(own, ch.shape)[internal]).all()
    # prefix sums are non-decreasing inside a node and the root's last prefix is < U
    ps = nodes["psum"].astype(np.int64)
    assert (np.diff(ps, axis=1) >= 0).all() and ps[-1, 7] < U and ps[:, 0].max() == 0
    W, H = 1920, 1080
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    _, t, nm, vi = g.render(cam, W, H, rgba=False)
    hit = t != FLT_MAX
    assert abs(int(hit.sum()) - ref["hits"]) < 0.01 * ref["hits"]
    assert int(vi[hit].max()) < U and (nm[hit] >= 0).all() and (nm[hit] <= 2).all() and (vi[~hit] == 0xFFFFFFFF).all()
    # the largest hit rank seen plus the root's bookkeeping: ranks cover the upper end of [0, U)
    assert int(vi[hit].max()) > 0.5 * U
    g2 = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    nodes2 = g2.export(True, False, False)[0]
    assert np.array_equal(nodes.view(np.uint8), nodes2.view(np.uint8))
    _, t2, nm2, vi2 = g2.render(cam, W, H, rgba=False)
    assert np.array_equal(t.view(np.uint32), t2.view(np.uint32)) and np.array_equal(vi, vi2)
    g.cleanUp()
    g2.cleanUp()


def test_build_beyond_the_24_bit_node_index(hk, ctx, bunny):
    """BASELINE config 3 at 2^14 on a dragon-like mesh: 18.9 M DAG nodes, more than the reference's child pointers can address.
    The tree is kept in the wide-index layout; invariants as in the sweep test, plus agreement of the silhouette with the same
    mesh at 4096^3 (reference layout)."""
    from scenes import dragon_standin
    tris, col, emi = dragon_standin(5, 0.15)
    W, H = 1920, 1080
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    hits = {}
    for grid in (4096, 16384):
        origin, dps = scene_params(tris, grid)
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
        info = g.info()
        st = ctx.build_stats()
        _, t, nm, vi = g.render(cam, W, H, rgba=False)
        hit = t != FLT_MAX
        hits[grid] = hit
        assert int(vi[hit].max()) < st["uniqueVoxels"] and (nm[hit] >= 0).all() and (nm[hit] <= 2).all()
        if grid == 16384:
            assert info.wideIndex == 1 and info.numberOfNodes >= (1 << 24) - 1
            assert st["uniqueVoxels"] > 500_000_000
            with pytest.raises(hk.HakoError):
                g.export(True, False, False)
        else:
            assert info.wideIndex == 0
        g.cleanUp()
    assert (hits[4096] != hits[16384]).mean() < 0.01
    assert abs(int(hits[4096].sum()) - int(hits[16384].sum())) < 0.01 * hits[4096].sum()


def test_build_subdivided_mesh_matches_oracle(hk, ctx, oracle, bunny):
    """many small triangles (the dragon-like regime): 82k triangles at 512^3"""
    tris = subdivide(bunny[0], 2, 0.15)
    col, emi = vertex_colors(tris)
    grid = 512
    origin, dps, keys, c, e, mk, mc, me, tree = _oracle_stages(oracle, tris, col, emi, grid)
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    nodes, gk, ga = g.export(True, True, True)
    np.testing.assert_array_equal(gk, mk)
    np.testing.assert_array_equal(ga["color"], mc)
    np.testing.assert_array_equal(ga["emission"], me)
    _cmp_nodes(nodes, tree.nodes())
    assert g.hasEmission() == bool(me[:, :3].any())


def test_build_rejects_bad_grid(hk, ctx, bunny):
    tris, col, emi = bunny
    with pytest.raises(hk.HakoError):
        hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, (0, 0, 0), 0.1, 100)      # not a power of two (reference abort()s)


def test_context_reserve(hk):
    """hako_ctx_reserve maps memory into the context's pool up front; an impossible size fails with an error and leaves the context usable"""
    c = hk.Context(0)
    c.reserve(0)
    c.reserve(64 << 20)
    with pytest.raises(hk.HakoError):
        c.reserve(1 << 50)
    keys = np.arange(1000, dtype=np.uint64)[::-1].copy()
    np.testing.assert_array_equal(c.sort_pairs(keys), np.arange(1000, dtype=np.uint64))
    c.close()
