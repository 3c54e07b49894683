# sharded build versus build-on-rank-0 + broadcast, under torchrun:  torchrun --nproc-per-node N tools/sharded_build_probe.py [grid ...]
import json, os, sys, time
import numpy as np, torch, torch.distributed as dist
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from massivevoxelraytracing_b200 import dist as hdist
from scenes import dragon_standin, scene_params
world = int(os.environ.get('WORLD_SIZE', '1')); rank = int(os.environ.get('RANK', '0')); local = int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local); dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
ctx = hk.Context(local)
tris, col, emi = dragon_standin(5, 0.15)
warm = torch.zeros(1, device=dev); dist.all_reduce(warm); torch.cuda.synchronize()
for grid in [int(a) for a in (sys.argv[1:] or ['4096'])]:
    origin, dps = scene_params(tris, grid)
    row = {'grid': grid, 'n_gpus': world}
    best = None
    for rep in range(3):
        times = {}
        dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        g = hdist.build_sharded(hdist.LibStages(ctx, hk.IntersectorOctreeGPU), tris, col, emi, origin, dps, grid, times=times)
        torch.cuda.synchronize(); dist.barrier(); wall = (time.perf_counter() - t0) * 1e3
        info = g.info()
        if best is None or wall < best: best = wall; row['phases_ms'] = {k: round(v, 2) for k, v in times.items() if k.endswith('_ms')}; row['range_voxels_this_rank'] = times['range_voxels_this_rank']
        if rep < 2: g.cleanUp()
    row['sharded_build_ms'] = round(best, 2); row['voxels'] = int(info.numberOfVoxels); row['nodes'] = int(info.numberOfNodes)
    if os.environ.get('HAKO_PROBE_NO_REF') == '1':
        g.cleanUp()
        if rank == 0: print(json.dumps(row), flush=True)
        continue
    # the replicated alternative: rank 0 builds, everybody receives
    best = None
    for rep in range(2):
        dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        ref = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid) if rank == 0 else None
        tb = (time.perf_counter() - t0) * 1e3
        ref = hdist.broadcast_octree(ctx, hk.IntersectorOctreeGPU, ref, 0)
        torch.cuda.synchronize(); dist.barrier(); wall = (time.perf_counter() - t0) * 1e3
        if best is None or wall < best[0]: best = (wall, tb)
        if rep == 0: ref.cleanUp()
    row['rank0_build_plus_broadcast_ms'] = round(best[0], 2); row['rank0_build_ms'] = round(best[1], 2)
    same = torch.tensor([int(ref.info().numberOfNodes == info.numberOfNodes and ref.info().numberOfVoxels == info.numberOfVoxels)], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN); row['same_counts_on_every_rank'] = bool(same.item())
    ref.cleanUp(); g.cleanUp()
    if rank == 0: print(json.dumps(row), flush=True)
dist.destroy_process_group()
