# scratch: the RTCamp-style frame loop (hako_seq_*: per frame updateScene -> clear -> steps -> resolve -> download) on a 256 -> 8192
# resolution ramp (RTCamp.cpp:138-196 animates the grid), with updateScene( f + 1 ) overlapped with the rendering of frame f
# (two streams, the default) against everything on one stream (HAKO_SEQ_OVERLAP=0).  Frames are identical either way.
import os, sys, time, hashlib, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from massivevoxelraytracing_b200.io import FrameSequence
from scenes import dragon_standin, scene_params, procedural_hdri
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
ramp = [256, 512, 1024, 2048, 4096, 8192, 4096, 2048, 1024, 512] * 2
ctx = hk.Context(0)
pt = hk.PathTracer(ctx).setup(); pt.resizeFrameBufferIfNeeded(None, W, H); env = procedural_hdri(1024, 512); pt.loadHDRI(None, env, env)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
params = {g: scene_params(tris, g) for g in set(ramp)}
out = os.environ.get('HAKO_SEQ_OUT')
for overlap in ('0', '1', '0', '1'):
    os.environ['HAKO_SEQ_OVERLAP'] = overlap
    seq = FrameSequence(pt, 4, 2)
    for g in ramp[:3]:                                  # warm the pools
        o, d = params[g]; seq.render_frame(tris, col, emi, o, float(d), g, cam, iterations=4, maxDepth=4)
    seq.wait()
    t0 = time.perf_counter()
    for i, g in enumerate(ramp):
        o, d = params[g]
        seq.render_frame(tris, col, emi, o, float(d), g, cam, iterations=4, maxDepth=4, png=(os.path.join(out, 'f%s_%02d.png' % (overlap, i)) if out else None))
    seq.wait()
    dt = time.perf_counter() - t0
    seq.close()
    print('HAKO_SEQ_OVERLAP=%s: %d frames (grid ramp 256..8192, 64 spp, depth 4, 1920x1080) in %.1f ms = %.2f frames/s' % (overlap, len(ramp), dt * 1e3, len(ramp) / dt), flush=True)
