"""The reference's OWN GPU kernels on the B200 (oracle/_ref/voxKernel.sm_100a*.cubin = /root/reference/voxKernel.cu compiled ahead of
time by oracle/Makefile, launched by oracle/librefgpu.so) against libhako_b200.so and against the CPU restatement.

This is what pins the rows whose reference code exists only on the device (a16 renderPT, a18 SAT build +
HDRI::sampleNearest, a19 renderResolve).  What can and cannot be bit-equal:

  * the reference kernels are compiled as the reference compiles them: fmad on, __sinf/__cosf/__expf/__logf (vectorMath.hpp:86-91);
    ours follow the CPU operation order (no contraction, libdevice sinf/cosf).  So hit ids / SAT words / accumulators agree up to
    the STATED tolerances below, each bounded from underneath by an A/A run of the reference against itself where the reference is
    itself nondeterministic (renderPT sums a pixel's 16 samples with shared-memory atomicAdd, voxKernel.cu:763-765);
  * the `nofma` cubin (-fmad=false) isolates the fast-intrinsic residual from the contraction residual.

Scene: the bunny at 1024^3 (2.1 M voxels, 86 489 DAG nodes -- the golden fixture's counts), 1920x1080, depth 8 = the reference's
loop bound, 4 steps = 64 spp; tolerances are written next to each assert.
"""
import numpy as np
import pytest

from scenes import procedural_hdri, scene_params

pytestmark = pytest.mark.gpu

GRID, W, H = 1024, 1920, 1080
HW, HH = 1024, 512

# stated tolerances (measured values are printed by the tests and recorded in DESIGN.md section 2)
HIT_ID_MISMATCH_MAX = 5e-5      # fraction of primary rays whose hit voxel differs, fmad-contracted reference traversal vs ours (measured 6.8e-6);
                                # against the -fmad=false build of the reference kernel the frames must be IDENTICAL
SAT_REL_MAX = 1e-6              # max |SAT_ref - SAT_ours| / 2^32: the reference's importance weights use __sinf / __cosf (measured 3.1e-7)
PT_REL_RMSE_MAX = 5e-3          # relative RMSE of 64-spp accumulators vs the reference kernel (measured 1.3e-3; 8.7e-4 vs the -fmad=false build;
                                # the reference against itself: 0 -- its shared-memory atomicAdd order is stable on this GPU)
PT_MEAN_REL_MAX = 1e-4          # |mean radiance difference| / mean radiance, i.e. no bias (measured 7.9e-6)
PT_CLOSE_MIN = 0.99             # fraction of pixels within 1e-3 relative (measured 0.9973)


@pytest.fixture(scope="module")
def refgpu_mod():
    from oracle import refgpu
    if not refgpu.available():
        pytest.skip("oracle/_ref/voxKernel.sm_100a.cubin / oracle/librefgpu.so not built (needs /root/reference at build time)")
    return refgpu


@pytest.fixture(scope="module")
def hk():
    import massivevoxelraytracing_b200 as h
    return h


@pytest.fixture(scope="module")
def ctx(hk):
    c = hk.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def built(hk, ctx, bunny):
    tris, col, emi = bunny
    origin, dps = scene_params(tris, GRID)
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), GRID)
    nodes, mortons, attrs = g.export(True, True, True)
    info = g.info()
    yield dict(g=g, nodes=nodes, mortons=mortons, attrs=attrs, lower=np.array(info.lower, np.float32), upper=np.array(info.upper, np.float32),
               dps=float(dps), origin=origin, has_emission=bool(info.hasEmission))
    g.cleanUp()


@pytest.fixture(scope="module")
def env():
    return procedural_hdri(HW, HH)


@pytest.fixture(scope="module")
def env_primary():
    """the separate image primary rays see (PathTracer::loadHDRI's filePrimary).  Both reference apps pass one (voxPTGPU.cpp:91,
    RTCamp.cpp:123); without it HDRI::sampleNearest( rd, true ) runs with m_widthPrimary = m_heightPrimary = 0 and reads
    m_pixels[-1] (renderCommon.hpp:356-363) -- undefined in the reference, so every comparison loads one."""
    return procedural_hdri(2048, 1024, seed=7)


def _ref(refgpu_mod, built, variant="default"):
    r = refgpu_mod.RefGpu(variant)
    b = built
    r.set_octree(b["nodes"], b["attrs"].view(np.uint8).reshape(-1, 8), b["lower"], b["upper"], b["dps"], b["has_emission"])
    return r


def test_reference_render_kernel_vs_ours(hk, refgpu_mod, built):
    """`render` (voxKernel.cu:437-483) vs hako_render_primary: RGBA frames and, through the rank-in-colour trick, the hit voxel ids"""
    g = built["g"]
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    rgba_n, t, nm, vi = g.render(cam, W, H)
    rgba_c, _, _, _ = g.render(cam, W, H, showVertexColor=True, ids=False)
    res = {}
    for variant in ("default", "nofma"):
        r = _ref(refgpu_mod, built, variant)
        img_n, best, mean = r.render(cam, W, H, False, reps=5)
        img_c, _, _ = r.render(cam, W, H, True)
        ids = r.render_ids(cam, W, H)
        r.close()
        mism = float((ids != vi).mean())
        res[variant] = mism
        print("reference render [%s]: %.3f ms best / %.3f mean = %.2f Grays/s; hit-id mismatch %.3g (%d of %d), normal-frame pixels differing %.3g, "
              "colour-frame %.3g" % (variant, best, mean, W * H / best / 1e6, mism, int((ids != vi).sum()), W * H,
                                     float((img_n != rgba_n).any(-1).mean()), float((img_c != rgba_c).any(-1).mean())))
        assert (ids != 0xFFFFFFFF).sum() > 0.15 * W * H
        assert mism <= HIT_ID_MISMATCH_MAX
        assert (img_n != rgba_n).any(-1).mean() <= HIT_ID_MISMATCH_MAX
        assert (img_c != rgba_c).any(-1).mean() <= HIT_ID_MISMATCH_MAX
        if variant == "nofma":
            # without contraction the reference's traversal is ours operation for operation: every hit id and every pixel equal
            assert mism == 0.0 and np.array_equal(img_n, rgba_n) and np.array_equal(img_c, rgba_c)


@pytest.mark.parametrize("w,h", [(256, 128), (1024, 512), (2048, 1024), (600, 300)])
def test_reference_sats_vs_ours(hk, ctx, oracle, refgpu_mod, w, h):
    """HDRIstoreImportance / buildSATh / buildSATv / buildSAT2u32 (voxKernel.cu:485-608) vs hako_pt_load_hdri and the CPU restatement"""
    img = procedural_hdri(w, h, seed=w)
    pt = hk.PathTracer(ctx).setup()
    pt.loadHDRI(None, img)
    ours = pt.hdri_sats(w, h).astype(np.int64)
    pt.cleanUp()
    port = oracle.hdri_build_sats(img, w, h).astype(np.int64)
    assert np.array_equal(ours, port)
    for variant in ("default", "nofma"):
        r = refgpu_mod.RefGpu(variant)
        ms = r.load_hdri(img)
        ref = r.sats().astype(np.int64)
        r.close()
        d = np.abs(ref - ours)
        print("SAT %dx%d [%s]: reference build %.3f ms; words differing %.4f, max |diff| %d = %.3g of 2^32; table 0 (no cosine weights) differing %.4f"
              % (w, h, variant, ms, float((d != 0).mean()), int(d.max()), d.max() / 2.0 ** 32, float((d[: w * h] != 0).mean())))
        assert d.max() / 2.0 ** 32 <= SAT_REL_MAX
        # the last word of every table is the normalisation itself
        assert (ref.reshape(7, -1)[:, -1] == 0xFFFFFFFF).all() and (ours.reshape(7, -1)[:, -1] == 0xFFFFFFFF).all()


def _stats(a, r):
    a = a.reshape(-1, 4).astype(np.float64)
    r = r.reshape(-1, 4).astype(np.float64)
    mean = r[:, :3].mean()
    rel = np.sqrt(((a[:, :3] - r[:, :3]) ** 2).mean()) / mean
    close = (np.abs(a[:, :3] - r[:, :3]).max(1) <= 1e-3 * np.maximum(np.abs(r[:, :3]).max(1), 1e-3)).mean()
    bias = abs(a[:, :3].mean() - mean) / mean
    return rel, close, bias


def test_reference_renderpt_vs_ours_and_oracle(hk, ctx, oracle, refgpu_mod, built, env, env_primary):
    """renderPT (voxKernel.cu:610-777), 4 steps of 16 spp at 1920x1080, depth 8, thin lens on:
    reference A/A  <  reference vs ours  <=  stated tolerance; and the CPU restatement against the same reference image."""
    b = built
    steps = 4
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.01)
    pt = hk.PathTracer(ctx).setup()
    pt.m_intersectorOctreeGPU = b["g"]
    pt.maxDepth = 8
    pt.resizeFrameBufferIfNeeded(None, W, H)
    pt.loadHDRI(None, env, env_primary)
    pt.clearFrameBuffer()
    for it in range(steps):
        pt.step(None, cam)
    ours = pt.accumulators().copy()
    ours_img = pt.toImageAsync()
    pmj = pt.pmj_table()
    assert np.array_equal(pmj.view(np.uint32), oracle.pmj_table().view(np.uint32))

    runs = {}
    for variant in ("default", "default", "nofma"):
        r = _ref(refgpu_mod, built, variant)
        r.load_hdri(env, env_primary)
        r.set_pmj(pmj)
        r.pt_resize(W, H)
        ms = [r.pt_step(cam, it) for it in range(steps)]
        acc = r.pt_accum()
        key = variant if variant not in runs else variant + "_again"
        runs[key] = acc
        if key == "default":
            ref_img = r.pt_resolve()
            r.pt_write_accum(ours)
            ref_resolve_of_ours = r.pt_resolve()
        print("reference renderPT [%s]: %s ms per 16-spp step" % (key, ", ".join("%.2f" % m for m in ms)))
        r.close()
    ref = runs["default"]
    assert np.array_equal(ref[..., 3], ours[..., 3]) and (ref[..., 3] == 16 * steps).all()

    aa = _stats(runs["default_again"], ref)
    vs = _stats(ours, ref)
    nf = _stats(ours, runs["nofma"])
    print("renderPT 64 spp %dx%d depth 8, bunny %d^3: reference A/A rel RMSE %.3g (close %.5f); ours vs reference rel RMSE %.3g, pixels within 1e-3 %.4f, "
          "mean-radiance difference %.3g; ours vs reference -fmad=false rel RMSE %.3g, close %.4f, mean diff %.3g"
          % (W, H, GRID, aa[0], aa[1], vs[0], vs[1], vs[2], nf[0], nf[1], nf[2]))
    assert aa[0] <= vs[0] + 1e-9
    assert vs[0] <= PT_REL_RMSE_MAX and vs[1] >= PT_CLOSE_MIN and vs[2] <= PT_MEAN_REL_MAX
    assert nf[0] <= PT_REL_RMSE_MAX and nf[2] <= PT_MEAN_REL_MAX

    # renderResolve (voxKernel.cu:779-795, __expf/__logf) on OUR sums vs our resolve (powf): at most 1 LSB apart
    dr = np.abs(ref_resolve_of_ours.astype(int) - ours_img.astype(int))
    print("renderResolve on the same accumulators: max |diff| %d LSB, bytes differing %.4f; resolved frames (each its own sums) bytes differing %.4f"
          % (int(dr.max()), float((dr != 0).mean()), float((ref_img != ours_img).mean())))
    assert dr.max() <= 1 and (dr != 0).mean() <= 0.02

    # the CPU restatement (what the other PT tests compare against) vs the reference kernel: one step, rows 400..520 (where the model is)
    y0, y1 = 400, 520
    tree = oracle.build_octree(b["mortons"], b["origin"], b["dps"], GRID)
    sats = oracle.hdri_build_sats(env, HW, HH)
    ocam = oracle.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H, focus=6.0, lens_r=0.01)
    acc_o = np.zeros(W * H * 4, np.float32)
    a8 = b["attrs"].view(np.uint8).reshape(-1, 8)
    oracle.pt_render(tree, a8[:, :4], a8[:, 4:], int(b["has_emission"]), ocam, W, H, 0, pmj, acc_o, env, sats, (HW, HH), primary=env_primary,
                     primary_wh=(env_primary.shape[1], env_primary.shape[0]), max_depth=8, y0=y0, y1=y1)
    r = _ref(refgpu_mod, built, "default")
    r.load_hdri(env, env_primary)
    r.set_pmj(pmj)
    r.pt_resize(W, H)
    r.pt_step(cam, 0)
    ref1 = r.pt_accum()[y0:y1]
    r.close()
    pt.clearFrameBuffer()
    pt.step(None, cam, rows=(y0, y1), iteration=0)
    ours1 = pt.accumulators()[y0:y1]
    po = _stats(acc_o.reshape(H, W, 4)[y0:y1], ref1)
    oo = _stats(ours1, acc_o.reshape(H, W, 4)[y0:y1])
    print("16 spp, rows %d..%d: CPU restatement vs reference kernel rel RMSE %.3g, close %.4f, mean diff %.3g; ours vs CPU restatement rel RMSE %.3g"
          % (y0, y1, po[0], po[1], po[2], oo[0]))
    assert po[0] <= 2 * PT_REL_RMSE_MAX and po[2] <= 4 * PT_MEAN_REL_MAX     # 16 spp: twice the 64-spp noise
    assert oo[0] <= 2e-3
    pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
    pt.cleanUp()


def test_reference_renderpt_depth4_no_lens(hk, ctx, refgpu_mod, built, env, env_primary):
    """BASELINE config 3's bounce count: the d4 cubin (`depth < 8` -> `depth < 4`, one-token sed in oracle/Makefile) vs maxDepth = 4"""
    if not refgpu_mod.available("d4"):
        pytest.skip("d4 cubin not built")
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    pt = hk.PathTracer(ctx).setup()
    pt.m_intersectorOctreeGPU = built["g"]
    pt.maxDepth = 4
    pt.resizeFrameBufferIfNeeded(None, W, H)
    pt.loadHDRI(None, env, env_primary)
    for it in range(2):
        pt.step(None, cam)
    ours = pt.accumulators().copy()
    r = _ref(refgpu_mod, built, "d4")
    r.load_hdri(env, env_primary)
    r.set_pmj(pt.pmj_table())
    r.pt_resize(W, H)
    ms = [r.pt_step(cam, it) for it in range(2)]
    ref = r.pt_accum()
    r.close()
    vs = _stats(ours, ref)
    print("renderPT depth 4, 32 spp: reference %.2f ms per step; ours vs reference rel RMSE %.3g, close %.4f, mean diff %.3g" % (min(ms), vs[0], vs[1], vs[2]))
    assert vs[0] <= 1.5 * PT_REL_RMSE_MAX and vs[2] <= 2 * PT_MEAN_REL_MAX
    pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
    pt.cleanUp()
