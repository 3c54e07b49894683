# scratch: host-call build time (pageable vertex arrays in, octree resident) against the number of staging threads (HAKO_H2D_WORKERS)
import sys, time, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
print('host threads', os.cpu_count(), flush=True)
for workers in [int(w) for w in os.environ.get('PROBE_WORKERS', '4,6,8,10,12,16').split(',')]:
    os.environ['HAKO_H2D_WORKERS'] = str(workers)
    ctx = hk.Context(0)
    walls = []
    for i in range(6):
        t0 = time.perf_counter()
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
        walls.append((time.perf_counter() - t0) * 1e3)
        st = dict(ctx.build_stats()); g.cleanUp()
    print('workers %2d: host call best %.2f median %.2f ms | last build: h2d %.2f vox %.2f total %.2f' % (workers, min(walls[1:]), sorted(walls[1:])[2], st['ms_h2d'], st['ms_voxelize'], st['ms_total']), flush=True)
    ctx.close()
