# scratch: ray binning in front of the trace waves (HAKO_PT_BIN = bits per axis of the origin grid, 0 = off) at BASELINE config[2] scale:
# one octree, one path tracer per setting (the knob is read at create time), device time per 16-spp step and checksum of the accumulators
import sys, os, hashlib, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
base = hk.PathTracer(ctx).setup()
base.updateScene(tris, col, emi, None, origin, float(dps), grid)
img = procedural_hdri(2048, 1024)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
settings = [dict(HAKO_PT_BIN='0')] + [dict(HAKO_PT_BIN=str(b), HAKO_PT_BIN_CELL_MAJOR=str(c)) for b in (2, 3, 4, 5, 6) for c in (0, 1)]
for depth in [int(d) for d in os.environ.get('PT_DEPTHS', '4,8').split(',')]:
    for env in settings:
        for k in ('HAKO_PT_BIN', 'HAKO_PT_BIN_CELL_MAJOR'):
            os.environ.pop(k, None)
        os.environ.update(env)
        pt = hk.PathTracer(ctx).setup()
        pt.m_intersectorOctreeGPU = base.m_intersectorOctreeGPU
        pt.loadHDRI(None, img)
        pt.maxDepth = depth
        pt.resizeFrameBufferIfNeeded(None, W, H)
        pt.clearFrameBuffer()
        pt.step(None, cam, iteration=0); ctx.synchronize()
        pt.clearFrameBuffer()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for it in range(steps): pt.step(None, cam, iteration=it)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / steps
        digest = hashlib.sha1(pt.accumulators().tobytes()).hexdigest()[:12]
        print('grid %d depth %d %-52s %.2f ms/step = %.3f ms/spp  accumulators sha1 %s' % (grid, depth, env, ms, ms / 16, digest), flush=True)
        pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
        pt.cleanUp()
