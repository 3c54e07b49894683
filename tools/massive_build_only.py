# scratch: one build of the stand-in dragon at a large grid (for ncu launch lists of the build kernels)
import sys, time, os, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from scenes import dragon_standin, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
tris, col, emi = dragon_standin(5, 0.15)
origin, dps = scene_params(tris, grid)
ctx = hk.Context(0)
for i in range(reps):
    t0 = time.perf_counter()
    g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    print('build %d: %.1f ms' % (i, (time.perf_counter() - t0) * 1e3), {k: (round(v, 1) if isinstance(v, float) else int(v)) for k, v in ctx.build_stats().items()}, flush=True)
    g.cleanUp()
