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
for _ in range(1 if os.environ.get('PROFILE') else 3): step()
torch.cuda.synchronize()
s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
N = 1 if os.environ.get('PROFILE') else 10
s.record()
for _ in range(N): step()
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / N
print('grid %d primary: %.3f ms/step, %.1f Mrays/s, checksum %d' % (grid, ms, V * W * H / ms / 1e3, int(vi.to(torch.int64).sum().item())))
import ctypes as C
t_h = torch.empty(W * H, dtype=torch.float32).pin_memory().numpy(); nm_h = torch.empty(W * H, dtype=torch.int32).pin_memory().numpy(); vi_h = torch.empty(W * H, dtype=torch.int32).pin_memory().numpy()
rgba_h = torch.empty(W * H * 4, dtype=torch.uint8).pin_memory().numpy()
for name, args in (('t+nMajor+vIndex (12 B/ray)', (None, t_h.ctypes.data, nm_h.ctypes.data, vi_h.ctypes.data)), ('rgba+vIndex (8 B/ray)', (rgba_h.ctypes.data, None, None, vi_h.ctypes.data)), ('rgba (4 B/ray)', (rgba_h.ctypes.data, None, None, None))):
    def hstep():
        for c in cams: hk._abi.check(ctx.lib.hako_render_primary(g.h, None, C.byref(c), W, H, 0, H, 0, *args))
    hstep(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3): hstep()
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3 / 3
    print('e2e host buffers %s: %.3f ms/step, %.1f Mrays/s' % (name, ms, V * W * H / ms / 1e3))
assert (vi_h.view(np.uint32) == vi[V - 1].cpu().numpy().view(np.uint32)).all()
if do_pt:
    pt.loadHDRI(None, procedural_hdri(2048, 1024))
    pt.resizeFrameBufferIfNeeded(None, W, H)
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
    for depth in ((8,) if os.environ.get('PROFILE') else (4, 8)):
        pt.maxDepth = depth
        pt.clearFrameBuffer(); pt.step(None, cam); ctx.synchronize(); pt.ray_count(reset=True)
        s.record()
        K = 1 if os.environ.get('PROFILE') else 4
        for _ in range(K): pt.step(None, cam)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / K; rays = pt.ray_count() / K
        print('grid %d PT depth %d: %.2f ms/step = %.3f ms/spp, %.1f Mpaths/s, %.1f Mrays/s' % (grid, depth, ms, ms / 16, W * H * 16 / ms / 1e3, rays / ms / 1e3))
    print('accum checksum', float(pt.accumulators().astype(np.float64).sum()))
