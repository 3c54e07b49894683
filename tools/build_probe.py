# scratch: build-stage timings on the bench scene for the voxelizer's columns-per-item knob and the H2D worker count
import sys, time, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, bunny_scene, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
scenes = {'standin': dragon_standin(5, 0.15), 'bunny': bunny_scene()}
for name, (tris, col, emi) in scenes.items():
    origin, dps = scene_params(tris, grid)
    for workers in ((1, 2, 4, 8) if name == 'standin' else (4,)):
        for k in ((1, 2, 4, 8, 16) if workers == 4 else (4,)):
            os.environ['HAKO_VOX_COLS_PER_ITEM'] = str(k); os.environ['HAKO_H2D_WORKERS'] = str(workers)
            ctx = hk.Context(0)
            best = None
            for i in range(3):
                t0 = time.perf_counter()
                g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
                wall = (time.perf_counter() - t0) * 1e3
                st = ctx.build_stats()
                if best is None or wall < best[0]: best = (wall, dict(st))
                info = g.info(); g.cleanUp()
            st = best[1]
            print('%s grid %d workers %d K %2d: wall %.1f ms | h2d %.2f vox %.2f sort %.2f uniq %.2f oct %.2f total %.2f | D %d U %d nodes %d' % (
                name, grid, workers, k, best[0], st['ms_h2d'], st['ms_voxelize'], st['ms_sort'], st['ms_unique'], st['ms_octree'], st['ms_total'],
                st['dumpedVoxels'], st['uniqueVoxels'], st['numberOfNodes']), flush=True)
            ctx.close()
