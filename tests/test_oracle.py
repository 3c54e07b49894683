"""CPU tests of the oracle: against the committed golden vectors (produced by the reference's own CPU
code, tests/golden/make_golden.py), against the reference itself when oracle/_ref is built, and the
reference's own unit-test properties for this path (unittest.cpp:183-227 morton)."""
import numpy as np
import pytest

from scenes import scene_params

FLT_MAX = np.float32(3.402823466e38)


def _pipeline(o, bunny, grid, W, H):
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    keys, c, e = o.voxelize(tris, col, emi, origin, float(dps), grid)
    mk, mc, me = o.merge_voxels(keys, c, e)
    tree = o.build_octree(mk, origin, float(dps), grid)
    cam = o.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H)
    t, nm, vi = tree.render_primary(cam, W, H)
    return dict(keys=keys, c=c, e=e, mk=mk, mc=mc, me=me, tree=tree, cam=cam, t=t, nm=nm, vi=vi, origin=origin, dps=dps)


@pytest.mark.parametrize("grid", [16, 64, 256, 1024])
def test_oracle_matches_golden(oracle, bunny, golden, grid):
    g = golden["cases"][str(grid)]
    W, H = g["image"]
    r = _pipeline(oracle, bunny, grid, W, H)
    f = lambda a: "%016x" % oracle.fnv1a64(a)
    assert float(r["dps"]).hex() == g["dps_hex"]
    assert r["keys"].size == g["D"]
    assert r["mk"].size == g["U"]
    assert f(np.sort(r["keys"])) == g["fnv_sorted_dumped_keys"]
    assert f(r["mk"]) == g["fnv_merged_keys"]
    assert f(np.ascontiguousarray(r["mc"][:, :3])) == g["fnv_merged_rgb"]
    assert f(np.ascontiguousarray(r["me"][:, :3])) == g["fnv_merged_emission_rgb"]
    nodes = r["tree"].nodes()
    assert nodes.size == g["dag_nodes"]
    assert f(np.ascontiguousarray(nodes["mask"])) == g["fnv_node_mask"]
    assert f(np.ascontiguousarray(nodes["children"])) == g["fnv_node_children"]
    assert f(np.ascontiguousarray(nodes["psum"])) == g["fnv_node_psum"]
    assert bytes(r["cam"]).hex() == g["camera_bytes_hex"]
    hit = r["t"] != FLT_MAX
    assert int(hit.sum()) == g["hits"]
    assert f(r["vi"]) == g["fnv_vindex"]
    assert f(r["nm"]) == g["fnv_nmajor"]
    assert f(r["t"]) == g["fnv_t"]
    origin = np.asarray(g["origin"], np.float32)
    naive = oracle.build_octree(r["mk"], origin, float(r["dps"]), grid, dag=False)
    assert naive.num_nodes == g["naive_nodes"]


@pytest.mark.parametrize("grid", [8192, 16384])
def test_oracle_matches_golden_large_grid_subset(oracle, bunny, golden, grid):
    """BASELINE config 4 grids (39 / 42-bit Morton keys) on the first 64 bunny triangles, in the full mesh's frame"""
    g = golden["subset_cases"][str(grid)]
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    n = g["triangles"]
    keys, c, e = oracle.voxelize(tris[:n], col[:n], emi[:n], origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    tree = oracle.build_octree(mk, origin, float(dps), grid)
    f = lambda a: "%016x" % oracle.fnv1a64(a)
    assert (keys.size, mk.size, tree.num_nodes) == (g["D"], g["U"], g["dag_nodes"])
    assert f(np.sort(keys)) == g["fnv_sorted_dumped_keys"] and f(mk) == g["fnv_merged_keys"]
    assert f(np.ascontiguousarray(mc[:, :3])) == g["fnv_merged_rgb"] and f(np.ascontiguousarray(me[:, :3])) == g["fnv_merged_emission_rgb"]
    nodes = tree.nodes()
    assert f(np.ascontiguousarray(nodes["mask"])) == g["fnv_node_mask"] and f(np.ascontiguousarray(nodes["children"])) == g["fnv_node_children"]
    assert f(np.ascontiguousarray(nodes["psum"])) == g["fnv_node_psum"]
    W, H = g["image"]
    cam = oracle.camera_lookat((-2.2, 1.2, 1.0), tuple(float(x) for x in tris[:n].reshape(-1, 3).mean(0)), (0, 1, 0), None, W, H)
    assert bytes(cam).hex() == g["camera_bytes_hex"]
    t, nm, vi = tree.render_primary(cam, W, H)
    assert int((t != FLT_MAX).sum()) == g["hits"] and f(vi) == g["fnv_vindex"] and f(nm) == g["fnv_nmajor"] and f(t) == g["fnv_t"]


def test_survey_probe_values(golden):
    """values the surveyor measured by driving the reference headers."""
    c = golden["cases"]["256"]
    assert (c["D"], c["U"], c["naive_nodes"], c["dag_nodes"], c["hits"]) == (185985, 133016, 60917, 10935, 416497)
    assert abs(c["sum_t_hits"] - 2042694.863875) < 1e-5
    c = golden["cases"]["1024"]
    assert (c["D"], c["U"], c["naive_nodes"], c["dag_nodes"], c["hits"]) == (2320463, 2115628, 978199, 86489, 458220)
    c = golden["cases"]["4096"]     # the C2* row of SURVEY 8(a): bunny at the bench grid
    assert (c["D"], c["U"], c["naive_nodes"], c["dag_nodes"], c["hits"]) == (34653745, 33810276, 15658017, 734869, 456677)


@pytest.mark.parametrize("grid", [16, 64])
def test_oracle_small_cases_elementwise(oracle, bunny, small_cases, grid):
    r = _pipeline(oracle, bunny, grid, 64 if grid == 16 else 256, 64 if grid == 16 else 256)
    s = small_cases
    np.testing.assert_array_equal(r["keys"], s["keys_%d" % grid])       # same emission order as the reference loop
    np.testing.assert_array_equal(r["c"], s["colors_%d" % grid])
    np.testing.assert_array_equal(r["e"], s["emissions_%d" % grid])
    np.testing.assert_array_equal(r["mk"], s["mkeys_%d" % grid])
    np.testing.assert_array_equal(r["mc"], s["mcolors_%d" % grid])
    np.testing.assert_array_equal(r["me"], s["memissions_%d" % grid])
    np.testing.assert_array_equal(r["tree"].nodes().view(np.uint8).reshape(-1, 68), s["nodes_%d" % grid])
    np.testing.assert_array_equal(r["t"], s["t_%d" % grid])
    np.testing.assert_array_equal(r["nm"].astype(np.int8), s["nmajor_%d" % grid])
    np.testing.assert_array_equal(r["vi"], s["vindex_%d" % grid])


def test_oracle_degenerate_triangles(oracle, small_cases):
    s = small_cases
    deg = s["deg_tris"]
    for grid in (8, 32):
        for six in (1, 0):
            k, _, _ = oracle.voxelize(deg, np.ones_like(deg), np.zeros_like(deg), np.zeros(3, np.float32), 1.0 / grid, grid, six=bool(six))
            np.testing.assert_array_equal(np.sort(k), s["deg_keys_%d_%d" % (grid, six)])
            cnt = [oracle.voxelize(deg[i:i + 1], None, None, np.zeros(3, np.float32), 1.0 / grid, grid, six=bool(six), attributes=False)[0].size
                   for i in range(deg.shape[0])]
            np.testing.assert_array_equal(cnt, s["deg_counts_%d_%d" % (grid, six)])
            assert cnt[0] == 0 and cnt[1] == 0 and cnt[2] == 0     # collinear / point / off-grid => nothing


def test_oracle_random_rays(oracle, bunny, small_cases):
    s = small_cases
    tris, col, emi = bunny
    origin, dps = scene_params(tris, 64)
    tree = oracle.build_octree(s["mkeys_64"], origin, float(dps), 64)
    t, nm, vi = tree.intersect(s["rays_ro"], s["rays_rd"])
    np.testing.assert_array_equal(t, s["rays_t"])
    np.testing.assert_array_equal(nm.astype(np.int8), s["rays_nmajor"])
    np.testing.assert_array_equal(vi, s["rays_vindex"])
    ts, _, vis = tree.intersect(s["rays_ro"], s["rays_rd"], shadow=True)
    np.testing.assert_array_equal(ts, s["rays_shadow_t"])
    np.testing.assert_array_equal(vis, s["rays_shadow_vindex"])
    assert (vis[ts != FLT_MAX] == 0).all()        # shadow rays skip the nVoxelsPSum accumulation
    np.testing.assert_array_equal(ts, t)


def test_morton_variants_agree(oracle):
    """unittest.cpp:183-216 morton.encodedecode: Naive == PDEP, decode inverts."""
    rng = np.random.default_rng(0)
    for x, y, z in rng.integers(0, 1 << 21, (20000, 3)).tolist():
        m = oracle.morton_encode(x, y, z)
        assert m == oracle.morton_encode_naive(x, y, z)
        assert oracle.morton_decode(m) == (x, y, z)
    assert oracle.morton_encode(1, 0, 0) == 1 and oracle.morton_encode(0, 1, 0) == 2 and oracle.morton_encode(0, 0, 1) == 4


def test_morton_sort_bits(oracle):
    """unittest.cpp:218-227 morton.sortBit."""
    for i in range(21):
        m = oracle.morton_encode((1 << i) - 1, (1 << i) - 1, (1 << i) - 1)
        assert oracle.lib.orc_sort_bits_morton(1 << i) == bin(m).count("1")


def test_oracle_equals_reference_when_available(oracle, ref_oracle, bunny):
    """Direct A/B against the reference's own headers (build container only)."""
    for grid, W, H in ((32, 128, 96), (128, 320, 200)):
        a = _pipeline(oracle, bunny, grid, W, H)
        b = _pipeline(ref_oracle, bunny, grid, W, H)
        for k in ("keys", "c", "e", "mk", "mc", "me", "t", "nm", "vi"):
            np.testing.assert_array_equal(a[k], b[k], err_msg=k)
        np.testing.assert_array_equal(a["tree"].nodes(), b["tree"].nodes())
        assert bytes(a["cam"]) == bytes(b["cam"])
        for (x, y, ox, oy) in ((0, 0, 0.5, 0.5), (W - 1, H - 1, 0.25, 0.75), (17, 33, 0.0, 1.0)):
            for u, v in zip(oracle.camera_shoot(a["cam"], x, y, ox, oy, W, H), ref_oracle.camera_shoot(b["cam"], x, y, ox, oy, W, H)):
                np.testing.assert_array_equal(u, v)
    cam = oracle.camera_lookat((1, 2, 3), (0.5, 0, -1), (0, 1, 0), 0.6, 640, 480, focus=2.5, lens_r=0.1)
    cam2 = ref_oracle.camera_lookat((1, 2, 3), (0.5, 0, -1), (0, 1, 0), 0.6, 640, 480, focus=2.5, lens_r=0.1)
    assert bytes(cam) == bytes(cam2)
    for args in ((3, 5, 0.3, 0.6, 640, 480, 0.1, 0.9), (639, 479, 0.9, 0.1, 640, 480, 0.7, 0.2)):
        for u, v in zip(oracle.camera_shoot_thin_lens(cam, *args), ref_oracle.camera_shoot_thin_lens(cam2, *args)):
            np.testing.assert_array_equal(u, v)


def test_traversal_agrees_with_dense_grid_dda(oracle, bunny):
    from dda import check_against_dda
    tris, col, emi = bunny
    grid = 32
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk = oracle.merge_voxels(keys, c, e)[0]
    tree = oracle.build_octree(mk, origin, float(dps), grid)
    _, both = check_against_dda(lambda ro, rd: tree.intersect(ro, rd), mk, origin, dps, grid)
    assert both > 5000
