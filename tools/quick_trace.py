# scratch: first timing of the traversal kernel on an oracle-built tree (not bench.py)
import sys, time, numpy as np, torch
import os; R=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R,'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from oracle import orc
from scenes import bunny_scene, scene_params
o = orc.Oracle('port')
tris, col, emi = bunny_scene()
ctx = hk.Context(0)
for grid, W, H in ((256, 1024, 1024), (1024, 1920, 1080), (2048, 1920, 1080)):
    origin, dps = scene_params(tris, grid)
    t0=time.time(); keys, c, e = o.voxelize(tris, col, emi, origin, float(dps), grid); mk, mc, me = o.merge_voxels(keys, c, e)
    tree = o.build_octree(mk, origin, float(dps), grid); print('oracle build s', time.time()-t0)
    lo, up = tree.bounds()
    g = hk.IntersectorOctreeGPU(ctx).import_nodes(tree.nodes(), lo, up, float(dps))
    cam = hk.camera_lookat((4,4,4),(0,0,0),width=W,height=H)
    vi = torch.empty(W*H, dtype=torch.int32, device='cuda'); t = torch.empty(W*H, dtype=torch.float32, device='cuda'); nm = torch.empty_like(vi)
    for _ in range(3): g.render_device(cam, W, H, t=t, nMajor=nm, vIndex=vi)
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20): g.render_device(cam, W, H, t=t, nMajor=nm, vIndex=vi)
    e2.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e2)/20
    print(grid, W, H, 'nodes', g.m_numberOfNodes, 'ms', ms, 'Mrays/s', W*H/ms/1e3, 'hits', int((t != 3.402823466e38).sum()))
