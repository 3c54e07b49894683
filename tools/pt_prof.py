# scratch: two PT steps on the bench workload (for ncu)
import sys, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
pt = hk.PathTracer(ctx).setup()
pt.m_intersectorOctreeGPU = g
pt.loadHDRI(None, procedural_hdri(2048, 1024)); pt.resizeFrameBufferIfNeeded(None, W, H)
pt.maxDepth = 4
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
for _ in range(2): pt.step(None, cam)
ctx.synchronize()
print('rays', pt.ray_count())
