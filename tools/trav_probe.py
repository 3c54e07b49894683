# scratch: A/B timing of the primary-ray kernel (bench workload: stand-in dragon, 16 orbit views at 1080p) and of a PT step
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
import bench
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
do_pt = len(sys.argv) > 2 and sys.argv[2] == 'pt'
W, H, V = 1920, 1080, 16
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
pt = hk.PathTracer(ctx).setup()
pt.updateScene(tris, col, emi, None, origin, float(dps), grid)
g = pt.m_intersectorOctreeGPU
cams = [hk.camera_lookat(bench.view_eye(k, V), (0, 0, 0), width=W, height=H) for k in range(V)]
t = torch.empty((V, W * H), dtype=torch.float32, device='cuda'); nm = torch.empty((V, W * H), dtype=torch.int32, device='cuda'); vi = torch.empty_like(nm)
def step():
    for i, c in enumerate(cams): g.render_device(c, W, H, t=t[i], nMajor=nm[i], vIndex=vi[i])
for _ in range(3): step()
torch.cuda.synchronize()
s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
N = 10
s.record()
for _ in range(N): step()
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / N
print('grid %d primary: %.3f ms/step, %.1f Mrays/s, checksum %d' % (grid, ms, V * W * H / ms / 1e3, int(vi.to(torch.int64).sum().item())))
if do_pt:
    pt.loadHDRI(None, procedural_hdri(2048, 1024))
    pt.resizeFrameBufferIfNeeded(None, W, H)
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
    for depth in (4, 8):
        pt.maxDepth = depth
        pt.clearFrameBuffer(); pt.step(None, cam); ctx.synchronize(); pt.ray_count(reset=True)
        s.record()
        for _ in range(4): pt.step(None, cam)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 4; rays = pt.ray_count() / 4
        print('grid %d PT depth %d: %.2f ms/step = %.3f ms/spp, %.1f Mpaths/s, %.1f Mrays/s' % (grid, depth, ms, ms / 16, W * H * 16 / ms / 1e3, rays / ms / 1e3))
    print('accum checksum', float(pt.accumulators().astype(np.float64).sum()))
