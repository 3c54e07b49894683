# scratch: sweep the path-tracing kernel variants (HAKO_PT_SHADE_THRESHOLD x HAKO_PT_MIN_CTAS) on the bench PT workload
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
env = procedural_hdri(2048, 1024)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
ref = None
variants = [(0, 8, 8, 5)] + [(2, r, b, c) for c in (5, 6, 8) for (r, b) in ((8, 16), (16, 16), (24, 16), (16, 32), (28, 32), (32, 16))]
if len(sys.argv) > 2: variants = [(0, 8, 8, 5)] + [(2, r, b, c) for c in (5, 6) for (r, b) in ((8, 16), (16, 16), (24, 16), (28, 16), (28, 32), (32, 16), (32, 64))]
if os.environ.get('PT_SWEEP'):   # "refillMin:burst:ctas,..."
    variants = [(0, 8, 8, 5)] + [(2,) + tuple(int(x) for x in v.split(':')) for v in os.environ['PT_SWEEP'].split(',')]
for wf, rmin, burst, ctas in variants:
    os.environ['HAKO_PT_WAVEFRONT'] = str(wf); os.environ['HAKO_PT_REFILL_MIN'] = str(rmin); os.environ['HAKO_PT_BURST'] = str(burst); os.environ['HAKO_PT_TRACE_CTAS'] = str(ctas)
    pt = hk.PathTracer(ctx).setup()
    own = pt.m_intersectorOctreeGPU; pt.m_intersectorOctreeGPU = g
    pt.loadHDRI(None, env); pt.resizeFrameBufferIfNeeded(None, W, H)
    for depth in (4, 8):
        pt.maxDepth = depth
        pt.clearFrameBuffer(); pt.step(None, cam); ctx.synchronize(); pt.ray_count(reset=True)
        acc = pt.accumulators().copy()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(3): pt.step(None, cam)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 3; rays = pt.ray_count() / 3
        if ref is None: ref = {}
        same = None
        if (depth) in ref: same = bool(np.array_equal(ref[depth].view(np.uint32), acc.view(np.uint32)))
        else: ref[depth] = acc
        print('wavefront %d refillMin %2d burst %2d ctas %d depth %d: %.2f ms/step = %.3f ms/spp, %.1f Mrays/s, identical_to_variant0 %s' % (wf, rmin, burst, ctas, depth, ms, ms / 16, rays / ms / 1e3, same), flush=True)
    pt.m_intersectorOctreeGPU = own; pt.cleanUp()
