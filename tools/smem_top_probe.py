# scratch: A/B of HAKO_SMEM_TOP (0 = off, 64 / 512 = that many top nodes of the octree staged in shared memory per CTA) on the
# bench scene: primary frames overlapped on 4 streams, one frame alone, and a path-tracing step at 8192^3 depth 4.
# The variable is read once per process: run this script once per value.
import os, sys, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params, procedural_hdri
W, H = 1920, 1080
tris, col, emi = dragon_standin(5, 0.15)
ctx = hk.Context(0)
dev = [torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda() for a in (tris, col, emi)]
stream = torch.cuda.current_stream()
def ev_ms(fn, reps):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps): fn()
    b.record(stream); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
origin, dps = scene_params(tris, 4096)
g = hk.IntersectorOctreeGPU(ctx).build_device(dev[0], dev[1], dev[2], None, origin, float(dps), 4096)
cams = [hk.camera_lookat((float(4 * np.cos(a) + 4 * np.sin(a)), 4.0, float(-4 * np.sin(a) + 4 * np.cos(a))), (0, 0, 0), width=W, height=H) for a in 2 * np.pi * np.arange(64) / 64]
t_d = torch.empty((16, W * H), dtype=torch.float32, device='cuda'); nm_d = torch.empty((16, W * H), dtype=torch.int32, device='cuda'); vi_d = torch.empty((16, W * H), dtype=torch.int32, device='cuda')
sides = [torch.cuda.Stream() for _ in range(3)]; fork = torch.cuda.Event(); joins = [torch.cuda.Event() for _ in sides]
def views():
    fork.record(stream)
    for sd in sides: sd.wait_event(fork)
    for i, cam in enumerate(cams):
        k = i % 4
        g.render_device(cam, W, H, t=t_d[i % 16], nMajor=nm_d[i % 16], vIndex=vi_d[i % 16], stream=sides[k - 1] if k else stream)
    for sd, j in zip(sides, joins):
        j.record(sd); stream.wait_event(j)
ms64 = ev_ms(views, 5)
alone = ev_ms(lambda: g.render_device(cams[0], W, H, t=t_d[0], nMajor=nm_d[0], vIndex=vi_d[0], stream=stream), 20)
chk = int(vi_d[0].to(torch.int64).sum().item())
g.cleanUp()
origin, dps = scene_params(tris, 8192)
g = hk.IntersectorOctreeGPU(ctx).build_device(dev[0], dev[1], dev[2], None, origin, float(dps), 8192)
pt = hk.PathTracer(ctx).setup(); pt.m_intersectorOctreeGPU = g; pt.maxDepth = 4
pt.resizeFrameBufferIfNeeded(None, W, H); env = procedural_hdri(2048, 1024); pt.loadHDRI(None, env, env)
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
it = [0]
def step():
    pt.step(None, cam, iteration=it[0] % 4); it[0] += 1
pt_ms = ev_ms(step, 8)
pt.clearFrameBuffer(); pt.step(None, cam, iteration=0)
mean = float(pt.accumulators()[..., :3].mean())
print('HAKO_SMEM_TOP=%s: 64 views on 4 streams %.3f ms = %.2f Grays/s; one frame alone %.1f us = %.2f Grays/s; PT 8192^3 depth 4: %.2f ms per 16-spp step | checks: id sum %d, mean accum %.6f' % (
    os.environ.get('HAKO_SMEM_TOP', '0'), ms64, 64 * W * H / ms64 / 1e6, alone * 1e3, W * H / alone / 1e6, pt_ms, chk, mean), flush=True)
