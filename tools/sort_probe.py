# scratch: sort stage timing for digit width x occupancy on the bench scene
import sys, time, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
tris, col, emi = dragon_standin(5, 0.15)
ctx = hk.Context(0)
for grid in (4096, 8192):
    origin, dps = scene_params(tris, grid)
    for rb in (8, 9):
        for occ in (2, 3, 4):
            os.environ['HAKO_SORT_RBITS'] = str(rb); os.environ['HAKO_SORT_MIN_CTAS'] = str(occ)
            best = None
            for i in range(3):
                g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
                st = ctx.build_stats(); g.cleanUp()
                if best is None or st['ms_sort'] < best['ms_sort']: best = dict(st)
            print('grid %d rbits %d minCtas %d: sort %.2f ms (%d passes) vox %.2f uniq %.2f oct %.2f | U %d nodes %d' % (grid, rb, occ, best['ms_sort'], best['sortPasses'], best['ms_voxelize'], best['ms_unique'], best['ms_octree'], best['uniqueVoxels'], best['numberOfNodes']), flush=True)
