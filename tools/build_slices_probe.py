# scratch: host-call time of hako_octree_build on the bench scene; HAKO_BUILD_SLICES is read once per process
import sys, time, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
walls = []
for i in range(8):
    t0 = time.perf_counter()
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    walls.append((time.perf_counter() - t0) * 1e3)
    st = dict(ctx.build_stats()); info = g.info(); g.cleanUp()
print('slices %s: wall min %.1f median %.1f ms | h2d %.2f vox %.2f sort %.2f uniq %.2f oct %.2f total %.2f | D %d U %d nodes %d' % (
    os.environ.get('HAKO_BUILD_SLICES', 'default'), min(walls), float(np.median(walls)), st['ms_h2d'], st['ms_voxelize'], st['ms_sort'], st['ms_unique'], st['ms_octree'], st['ms_total'],
    st['dumpedVoxels'], st['uniqueVoxels'], st['numberOfNodes']), flush=True)
