# scratch: per-stage device times of hako_octree_build_device (vertex data resident: no upload in the way) on the bench scene
import sys, time, os, numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
dev = [torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda() for a in (tris, col, emi)]
best = None
for i in range(6):
    t0 = time.perf_counter()
    g = hk.IntersectorOctreeGPU(ctx).build_device(dev[0], dev[1], dev[2], None, origin, float(dps), grid)
    wall = (time.perf_counter() - t0) * 1e3
    st = dict(ctx.build_stats()); g.cleanUp()
    if best is None or st['ms_total'] < best[1]['ms_total']: best = (wall, st)
    if os.environ.get('PROBE_ALL'): print('   build %d: wall %.2f vox %.2f sort %.2f uniq %.2f oct %.2f' % (i, wall, st['ms_voxelize'], st['ms_sort'], st['ms_unique'], st['ms_octree']), flush=True)
wall, st = best
print('grid %d resident input: wall %.2f ms | vox %.2f sort %.2f uniq %.2f oct %.2f total %.2f | D %d U %d nodes %d' % (
    grid, wall, st['ms_voxelize'], st['ms_sort'], st['ms_unique'], st['ms_octree'], st['ms_total'], st['dumpedVoxels'], st['uniqueVoxels'], st['numberOfNodes']), flush=True)
