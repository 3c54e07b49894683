# scratch: what the FIRST build of a context pays on top of a warm one (bench scene).  Fresh process per line (run with an argument):
#   host    : hako_octree_build from pageable host arrays (pinned staging chunks + memory-pool growth + first launches)
#   device  : hako_octree_build_device from resident vertex arrays (no staging)
#   reserve : like host, after hako_ctx_reserve( 4 GiB )
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
mode = sys.argv[1] if len(sys.argv) > 1 else 'host'
grid = 4096
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
torch.cuda.init(); torch.zeros(1, device='cuda'); torch.cuda.synchronize()
t0 = time.perf_counter(); ctx = hk.Context(0); t_ctx = (time.perf_counter() - t0) * 1e3
t_res = 0.0
if mode == 'reserve':
    t0 = time.perf_counter(); ctx.reserve(4 << 30); t_res = (time.perf_counter() - t0) * 1e3
dev = [torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda() for a in (tris, col, emi)] if mode == 'device' else None
walls = []
for i in range(3):
    t0 = time.perf_counter()
    if mode == 'device':
        g = hk.IntersectorOctreeGPU(ctx).build_device(dev[0], dev[1], dev[2], None, origin, float(dps), grid)
    else:
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    ctx.synchronize()
    walls.append((time.perf_counter() - t0) * 1e3)
    g.cleanUp()
print('%-8s context %.1f ms, reserve %.1f ms, builds %s ms' % (mode, t_ctx, t_res, ' / '.join('%.1f' % w for w in walls)), flush=True)
