# scratch: what the wide-index node layout costs (a second dependent load per descent: the child's mask from the side array): the same tree
# built narrow and with HAKO_FORCE_WIDE=1 -- primary frames alone and a path-tracing step
import sys, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
env = procedural_hdri(2048, 1024)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
pcam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
for force in ('0', '1'):
    os.environ['HAKO_FORCE_WIDE'] = force
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    vi = torch.empty(W * H, dtype=torch.int32, device='cuda'); t = torch.empty(W * H, dtype=torch.float32, device='cuda'); nm = torch.empty_like(vi)
    for _ in range(3): g.render_device(cam, W, H, t=t, nMajor=nm, vIndex=vi)
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20): g.render_device(cam, W, H, t=t, nMajor=nm, vIndex=vi)
    e.record(); torch.cuda.synchronize()
    frame_ms = s.elapsed_time(e) / 20
    pt = hk.PathTracer(ctx).setup()
    pt.m_intersectorOctreeGPU = g
    pt.loadHDRI(None, env); pt.resizeFrameBufferIfNeeded(None, W, H); pt.maxDepth = 4
    pt.clearFrameBuffer(); pt.step(None, pcam); ctx.synchronize()
    s.record()
    for it in range(3): pt.step(None, pcam, iteration=it)
    e.record(); torch.cuda.synchronize()
    step_ms = s.elapsed_time(e) / 3
    print('grid %d wide %d (%d nodes): primary frame alone %.3f ms, path-tracing step %.2f ms' % (grid, g.info().wideIndex, g.info().numberOfNodes, frame_ms, step_ms), flush=True)
    pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx); pt.cleanUp(); g.cleanUp()
