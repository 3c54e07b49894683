# scratch: PT throughput at BASELINE config[2] scale (stand-in dragon, 1080p, procedural 2k HDR)
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
W, H = 1920, 1080
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
pt = hk.PathTracer(ctx).setup()
t0 = time.time(); pt.updateScene(tris, col, emi, None, origin, float(dps), grid); print('build s', time.time() - t0, ctx.build_stats())
pt.loadHDRI(None, procedural_hdri(2048, 1024))
pt.resizeFrameBufferIfNeeded(None, W, H)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
for depth in (4, 8):
    pt.maxDepth = depth
    pt.clearFrameBuffer()
    pt.step(None, cam); ctx.synchronize()
    pt.ray_count(reset=True)
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps): pt.step(None, cam)
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / steps
    rays = pt.ray_count() / steps
    print('grid %d depth %d: %.2f ms/step (16 spp) = %.3f ms/spp, %.1f Mpaths/s, %.1f Mrays/s, %.2f rays/path' % (grid, depth, ms, ms / 16, W * H * 16 / ms / 1e3, rays / ms / 1e3, rays / (W * H * 16)))
acc = pt.accumulators()
print('mean radiance', acc[..., :3].mean() / acc[..., 3].mean(), 'voxels', pt.getNumberOfVoxels(), 'nodes', pt.m_intersectorOctreeGPU.m_numberOfNodes)
