# scratch (CPU only): how long are the longest primary rays of a bench frame?  Loop visits per ray from the oracle port,
# per-warp (8x4 tile) maximum = the serial chain a warp of renderPrimaryKernel executes.
import sys, os, time, ctypes as C, numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
from oracle import orc
from scenes import dragon_standin, scene_params
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
levels = int(sys.argv[2]) if len(sys.argv) > 2 else 5
W, H = 1920, 1080
orc.build(quiet=True)
o = orc.Oracle("port")
tris, col, emi = dragon_standin(levels, 0.15)
origin, dps = scene_params(tris, grid)
t0 = time.time()
keys, c, e = o.voxelize(tris, col, emi, origin, float(dps), grid, threads=8) if 'threads' in o.voxelize.__code__.co_varnames else o.voxelize(tris, col, emi, origin, float(dps), grid)
mk, mc, me = o.merge_voxels(keys, c, e)
tree = o.build_octree(mk, origin, float(dps), grid)
print('built', grid, tree.num_nodes, 'nodes in %.1f s' % (time.time() - t0), flush=True)
cam = o.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H)
v = np.zeros(W * H, np.uint32)
o.lib.orc_render_primary_visits.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
o.lib.orc_render_primary_visits(tree.h, C.byref(cam), W, H, 8, v.ctypes.data)
v = v.reshape(H, W)
print('visits per ray: mean %.2f  p50 %d  p90 %d  p99 %d  p99.9 %d  max %d' % ((v.mean(),) + tuple(np.percentile(v, [50, 90, 99, 99.9, 100]).astype(int))))
Hc, Wc = H // 4 * 4, W // 8 * 8
tile = v[:Hc, :Wc].reshape(Hc // 4, 4, Wc // 8, 8).transpose(0, 2, 1, 3).reshape(Hc // 4, Wc // 8, 32)
wmax = tile.max(-1); wsum = tile.sum(-1)
print('per-warp max visits: mean %.1f  p50 %d  p90 %d  p99 %d  p99.9 %d  max %d' % ((wmax.mean(),) + tuple(np.percentile(wmax, [50, 90, 99, 99.9, 100]).astype(int))))
print('lane utilisation if a warp runs max(visits) iterations: %.3f' % (wsum.sum() / (32.0 * wmax.sum())))
print('sum over warps of max visits = %d; 64 warps/SM x 148 SMs -> %.0f iterations per warp slot; longest warp = %d iterations' % (wmax.sum(), wmax.sum() / (64 * 148), wmax.max()))
