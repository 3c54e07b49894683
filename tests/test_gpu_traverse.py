"""GPU parity of the traversal path (hako_octree_import / hako_intersect_batch / hako_render_primary) against the
CPU oracle: bit-exact t / nMajor / vIndex (fp32 operation order is the reference's; the library is built -fmad=false)."""
import numpy as np
import pytest

from scenes import scene_params

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


def _oracle_tree(oracle, bunny, grid):
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    tree = oracle.build_octree(mk, origin, float(dps), grid)
    return tree, mk, mc, me, origin, dps


def _import(hk, ctx, tree, mc, me, dps):
    lo, up = tree.bounds()
    attrs = np.zeros(mc.shape[0], hk.VOXEL_ATTRIB)
    attrs["color"] = mc
    attrs["emission"] = me
    g = hk.IntersectorOctreeGPU(ctx)
    g.import_nodes(tree.nodes(), lo, up, float(dps), attrs)
    return g, attrs


@pytest.mark.parametrize("grid,W,H", [(16, 64, 64), (64, 256, 256), (256, 1024, 1024)])
def test_render_primary_matches_oracle(hk, ctx, oracle, bunny, golden, grid, W, H):
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, grid)
    g, attrs = _import(hk, ctx, tree, mc, me, dps)
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    ocam = oracle.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H)
    assert bytes(cam) == bytes(ocam)
    t, nm, vi = tree.render_primary(ocam, W, H, threads=8)
    rgba, gt, gnm, gvi = g.render(cam, W, H)
    np.testing.assert_array_equal(gvi.reshape(-1), vi)
    np.testing.assert_array_equal(gnm.reshape(-1), nm)
    np.testing.assert_array_equal(gt.reshape(-1).view(np.uint32), t.view(np.uint32))   # bit-exact, stricter than 1e-6 rel
    case = golden["cases"][str(grid)]
    if case["image"] == [W, H]:
        assert "%016x" % oracle.fnv1a64(gvi) == case["fnv_vindex"]
        assert "%016x" % oracle.fnv1a64(gnm) == case["fnv_nmajor"]
        assert "%016x" % oracle.fnv1a64(gt) == case["fnv_t"]
        assert int((gt != FLT_MAX).sum()) == case["hits"]
    # normal-colour image: miss -> (0,0,0,255); hit -> (hitN + 1) / 2 (voxKernel.cu:470-476)
    hit = (gt != FLT_MAX)
    assert (rgba[~hit] == np.array([0, 0, 0, 255], np.uint8)).all()
    assert (rgba[hit][:, 3] == 255).all() and set(np.unique(rgba[hit][:, :3])) <= {0, 128, 255}
    # vertex-colour image == attribute lookup by hit id (voxKernel.cu:466-468)
    rgba2, _, _, _ = g.render(cam, W, H, showVertexColor=True, ids=False)
    np.testing.assert_array_equal(rgba2[hit], attrs["color"][gvi[hit]])
    # row-range sharding (how image tiles are split over GPUs) gives the same pixels
    y0, y1 = H // 3, H // 3 + 37 if H > 64 else H // 3 + 5
    _, t2, nm2, vi2 = g.render(cam, W, H, y0=y0, y1=y1)
    np.testing.assert_array_equal(vi2, gvi[y0:y1])
    np.testing.assert_array_equal(t2.view(np.uint32), gt[y0:y1].view(np.uint32))
    g.cleanUp()


def test_render_views_equals_per_frame_calls(hk, ctx, oracle, bunny):
    """hako_render_primary_views (a ring of 4 device frames, frame i+1 traced while frame i is copied out) returns what one
    hako_render_primary call per camera returns -- more views than ring slots, odd image size, both output kinds"""
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, 64)
    g, attrs = _import(hk, ctx, tree, mc, me, dps)
    W, H = 203, 117
    cams = [hk.camera_lookat((4 * np.cos(a), 1.0 + 0.5 * k, 4 * np.sin(a)), (0, 0, 0), width=W, height=H) for k, a in enumerate(np.linspace(0, 5, 11))]
    rgba, t, nm, vi = g.render_views(cams, W, H, showVertexColor=True)
    assert rgba.shape == (11, H, W, 4) and t.shape == (11, H, W)
    for k, cam in enumerate(cams):
        r1, t1, nm1, vi1 = g.render(cam, W, H, showVertexColor=True)
        np.testing.assert_array_equal(rgba[k], r1)
        np.testing.assert_array_equal(t[k].view(np.uint32), t1.view(np.uint32))
        np.testing.assert_array_equal(nm[k], nm1)
        np.testing.assert_array_equal(vi[k], vi1)
    assert len({int((t[k] != FLT_MAX).sum()) for k in range(11)}) > 3     # the views differ
    # one output kind only, fewer views than ring slots, and the empty run
    only = g.render_views(cams[:2], W, H, rgba=True, ids=False)
    assert only[1] is None and np.array_equal(only[0], g.render_views(cams[:2], W, H, showVertexColor=False)[0])
    assert g.render_views([], W, H)[0].shape == (0, H, W, 4)
    g.cleanUp()


def test_export_roundtrip(hk, ctx, oracle, bunny):
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, 64)
    g, attrs = _import(hk, ctx, tree, mc, me, dps)
    nodes, _, a = g.export(nodes=True, attrs=True)
    ref = tree.nodes()
    for f in ("mask", "children", "psum"):
        np.testing.assert_array_equal(nodes[f], ref[f])
    np.testing.assert_array_equal(a, attrs)
    assert g.m_numberOfNodes == ref.size and g.m_numberOfVoxels == mk.size
    assert g.hasEmission() == bool(me[:, :3].any())


def test_import_validates_and_accepts_any_float_sequence(hk, ctx, oracle, bunny):
    """hako_octree_import rejects trees the traversal could not walk safely (child index not below its parent, embedded mask that
    is not the child's mask, a child without its mask bit), and lower / upper may be tuples or float64 arrays (the binding keeps
    the converted arrays alive across the call)"""
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, 32)
    lo, up = tree.bounds()
    nodes = tree.nodes()
    g = hk.IntersectorOctreeGPU(ctx)
    g.import_nodes(nodes, tuple(float(x) for x in lo), np.asarray(up, np.float64), float(dps))
    i = g.info()
    assert np.array_equal(np.array(i.lower, np.float32), lo) and np.array_equal(np.array(i.upper, np.float32), up)
    assert i.gridRes == 32
    ro = np.array([[4.0, 4.0, 4.0]], np.float32)
    rd = np.array([[-1.0, -1.0, -1.0]], np.float32)
    t, nm, vi = g.intersect(ro, rd)
    ot, onm, ovi = tree.intersect(ro, rd)
    assert t.view(np.uint32)[0] == ot.view(np.uint32)[0] and vi[0] == ovi[0]
    g.cleanUp()
    root = nodes.size - 1
    c = int(np.nonzero(nodes["children"][root] != 0xFFFFFFFF)[0][0])
    for kind in ("forward", "mask", "bit"):
        bad = nodes.copy()
        if kind == "forward":
            bad["children"][5][0] = np.uint32(root | (int(nodes["mask"][root]) << 24))       # a cycle / forward reference
            bad["mask"][5] |= 1
        elif kind == "mask":
            bad["children"][root][c] ^= np.uint32(1 << 24)                                   # embedded mask != the child's mask
        else:
            bad["mask"][root] &= np.uint8(~(1 << c) & 0xFF)                                  # child present, mask bit clear
        with pytest.raises(hk.HakoError):
            hk.IntersectorOctreeGPU(ctx).import_nodes(bad, lo, up, float(dps))


def test_random_rays_and_shadow(hk, ctx, oracle, bunny, small_cases):
    s = small_cases
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, 64)
    g, _ = _import(hk, ctx, tree, mc, me, dps)
    t, nm, vi = g.intersect(s["rays_ro"], s["rays_rd"])
    np.testing.assert_array_equal(t.view(np.uint32), s["rays_t"].view(np.uint32))
    np.testing.assert_array_equal(nm.astype(np.int8), s["rays_nmajor"])
    np.testing.assert_array_equal(vi, s["rays_vindex"])
    ts, _, vis = g.intersect(s["rays_ro"], s["rays_rd"], isShadowRay=True)
    np.testing.assert_array_equal(ts.view(np.uint32), s["rays_shadow_t"].view(np.uint32))
    np.testing.assert_array_equal(vis, s["rays_shadow_vindex"])
    # caller presets survive a miss (voxRT.cpp:313-325)
    n = s["rays_ro"].shape[0]
    t0 = np.full(n, 123.0, np.float32)
    nm0 = np.full(n, 77, np.int32)
    vi0 = np.full(n, 99, np.uint32)
    g.intersect(s["rays_ro"], s["rays_rd"], t0, nm0, vi0)
    miss = s["rays_t"] == FLT_MAX
    assert miss.any() and (t0[miss] == 123.0).all() and (nm0[miss] == 77).all() and (vi0[miss] == 99).all()
    np.testing.assert_array_equal(vi0[~miss], s["rays_vindex"][~miss])


def test_edge_rays(hk, ctx, oracle, bunny):
    """axis-parallel rays (rd component = +-0), rays starting inside, negative directions, zero-length batches"""
    tree, mk, mc, me, origin, dps = _oracle_tree(oracle, bunny, 64)
    g, _ = _import(hk, ctx, tree, mc, me, dps)
    lo, up = tree.bounds()
    c = (lo + up) * np.float32(0.5)
    rng = np.random.default_rng(3)
    ro, rd = [], []
    for ax in range(3):
        for sgn in (1.0, -1.0):
            for k in range(200):
                o = c + (rng.random(3).astype(np.float32) - np.float32(0.5)) * (up - lo)
                d = np.zeros(3, np.float32)
                d[ax] = sgn
                if k % 2:
                    d[(ax + 1) % 3] = -0.0
                o[ax] = (lo[ax] - 1.0) if sgn > 0 else (up[ax] + 1.0)
                ro.append(o)
                rd.append(d)
    inside = c + (rng.random((500, 3)).astype(np.float32) - np.float32(0.5)) * (up - lo) * np.float32(0.9)
    ro += list(inside)
    rd += list(rng.standard_normal((500, 3)).astype(np.float32) * np.float32(1e-3))   # un-normalised, tiny
    ro = np.array(ro, np.float32)
    rd = np.array(rd, np.float32)
    t, nm, vi = tree.intersect(ro, rd)
    gt, gnm, gvi = g.intersect(ro, rd)
    np.testing.assert_array_equal(gt.view(np.uint32), t.view(np.uint32))
    np.testing.assert_array_equal(gnm, nm)
    np.testing.assert_array_equal(gvi, vi)
    assert (t != FLT_MAX).sum() > 100
    e = np.zeros((0, 3), np.float32)
    assert g.intersect(e, e)[0].size == 0


def test_wide_index_layout_matches_reference_layout(hk, ctx, bunny, small_cases, monkeypatch):
    """Trees with >= 2^24 - 1 nodes cannot use the reference's child pointers (index | mask << 24, voxCommon.hpp:183-195); they are
    kept with plain 32-bit indices and the traversal reads child masks from the side array.  Forced on a small tree, that layout must
    give bit-identical hits (primary frame, random + exotic rays, shadow rays, a path-traced step) and refuse the 68-byte export."""
    from scenes import procedural_hdri
    tris, col, emi = bunny
    grid = 64
    origin, dps = scene_params(tris, grid)
    W, H = 320, 200
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    env = procedural_hdri(128, 64)
    res = []
    for force in ("0", "1"):
        monkeypatch.setenv("HAKO_FORCE_WIDE", force)
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
        assert g.info().wideIndex == int(force)
        rgba, t, nm, vi = g.render(cam, W, H)
        rt = g.intersect(small_cases["rays_ro"], small_cases["rays_rd"])
        rs = g.intersect(small_cases["rays_ro"], small_cases["rays_rd"], isShadowRay=True)
        pt = hk.PathTracer(ctx).setup()
        pt.m_intersectorOctreeGPU = g
        pt.maxDepth = 4
        pt.resizeFrameBufferIfNeeded(None, 96, 64)
        pt.loadHDRI(None, env)
        pcam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=96, height=64, focus=6.0, lensR=0.01)
        pt.step(None, pcam)
        acc = pt.accumulators().copy()
        # the trace waves keep their traversal stack in shared memory by default; the local-memory form must agree (wide trees carry
        # the node's mask in the stack entry)
        monkeypatch.setenv("HAKO_PT_SMEM_STACK", "0")
        pt2 = hk.PathTracer(ctx).setup()
        monkeypatch.delenv("HAKO_PT_SMEM_STACK")
        pt2.m_intersectorOctreeGPU = g
        pt2.maxDepth = 4
        pt2.resizeFrameBufferIfNeeded(None, 96, 64)
        pt2.loadHDRI(None, env)
        pt2.step(None, pcam)
        assert np.array_equal(pt2.accumulators().view(np.uint32), acc.view(np.uint32))
        pt2.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
        pt2.cleanUp()
        if force == "1":
            with pytest.raises(hk.HakoError):
                g.export(True, False, False)
            assert g.export(False, True, False)[1].size == g.info().numberOfVoxels      # keys / attributes still export
        pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
        pt.cleanUp()
        res.append((rgba, t, nm, vi, rt, rs, acc, g.info().numberOfNodes))
        g.cleanUp()
    a, b = res
    assert a[7] == b[7]
    for k in range(4):
        assert np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8))
    for k in (4, 5):
        for x, y in zip(a[k], b[k]):
            assert np.array_equal(np.asarray(x).view(np.uint8), np.asarray(y).view(np.uint8))
    assert np.array_equal(a[6].view(np.uint32), b[6].view(np.uint32))
    # the same rays against the golden values of the reference
    assert np.array_equal(b[4][0].view(np.uint32), small_cases["rays_t"].view(np.uint32))
    assert np.array_equal(b[4][2], small_cases["rays_vindex"])


@pytest.mark.parametrize("grid", [32, 64])
def test_cuda_traversal_agrees_with_dense_grid_dda(hk, ctx, bunny, grid):
    """the independent comparator (float64 Amanatides-Woo march over the dense occupancy grid, tests/dda.py) against the CUDA
    path itself: hako_octree_build + hako_intersect_batch, no oracle involved"""
    from dda import check_against_dda
    tris, col, emi = bunny
    origin, dps = scene_params(tris, grid)
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    _, mk, _ = g.export(False, True, False)
    compared, both = check_against_dda(lambda ro, rd: g.intersect(ro, rd), mk, origin, dps, grid, n=20000 if grid == 32 else 6000)
    assert both > 1500
    g.cleanUp()
