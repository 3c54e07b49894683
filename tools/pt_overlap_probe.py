# scratch: does a PT step gain from running two row bands concurrently on two streams (tails of the trace waves overlap)?
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
for nparts in (1, 2, 3, 4):
    pts = []
    for k in range(nparts):
        pt = hk.PathTracer(ctx).setup(); pt._own = pt.m_intersectorOctreeGPU; pt.m_intersectorOctreeGPU = g
        pt.loadHDRI(None, env); pt.resizeFrameBufferIfNeeded(None, W, H); pt.maxDepth = 4; pt.clearFrameBuffer()
        pts.append(pt)
    streams = [torch.cuda.Stream() for _ in range(nparts)]
    bounds = [(H * k // nparts) // 4 * 4 for k in range(nparts)] + [H]
    def step(it):
        for k in range(nparts):
            pts[k].step(streams[k], cam, rows=(bounds[k], bounds[k + 1]), iteration=it)
    step(0); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in range(1, 5): step(it)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / 4
    print('%d concurrent row bands: %.2f ms/step' % (nparts, ms), flush=True)
    for pt in pts:
        pt.m_intersectorOctreeGPU = pt._own; pt.cleanUp()
