# scratch: D2H bandwidth, call overhead, and overlap check for the host-buffer primary-ray call
import sys, time, os, numpy as np, torch, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
import bench
for mb in (1, 3, 8, 25, 100):
    d = torch.empty(mb << 20, dtype=torch.uint8, device='cuda'); h = torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    h.copy_(d); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    print('D2H %3d MB pinned: %.3f ms  %.1f GB/s' % (mb, dt * 1e3, (mb << 20) / dt / 1e9))
grid = 4096; W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
vi_h = torch.empty(W * H, dtype=torch.int32).pin_memory().numpy(); t_h = torch.empty(W * H, dtype=torch.float32).pin_memory().numpy(); nm_h = torch.empty(W * H, dtype=torch.int32).pin_memory().numpy()
vi_d = torch.empty(W * H, dtype=torch.int32, device='cuda')
def timeit(f, n=20):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print('device call (vIndex only)    %.3f ms' % timeit(lambda: g.render_device(cam, W, H, vIndex=vi_d)))
scam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=16, height=16)
print('host call 16x16 (overhead)   %.3f ms' % timeit(lambda: hk._abi.check(ctx.lib.hako_render_primary(g.h, None, C.byref(scam), 16, 16, 0, 16, 0, None, None, None, vi_h.ctypes.data))))
print('host call 1080p vIndex       %.3f ms' % timeit(lambda: hk._abi.check(ctx.lib.hako_render_primary(g.h, None, C.byref(cam), W, H, 0, H, 0, None, None, None, vi_h.ctypes.data))))
print('host call 1080p t+nm+vIndex  %.3f ms' % timeit(lambda: hk._abi.check(ctx.lib.hako_render_primary(g.h, None, C.byref(cam), W, H, 0, H, 0, None, t_h.ctypes.data, nm_h.ctypes.data, vi_h.ctypes.data))))
pg = np.empty(W * H, np.uint32)
print('host call 1080p vIndex PAGEABLE %.3f ms' % timeit(lambda: hk._abi.check(ctx.lib.hako_render_primary(g.h, None, C.byref(cam), W, H, 0, H, 0, None, None, None, pg.ctypes.data))))
