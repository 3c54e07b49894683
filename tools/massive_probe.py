# BASELINE config 4 (scaled to the 32-bit voxel counters): the stand-in dragon at 32768^3 -- billions of surface voxels in a wide-index
# octree -- built on rank 0, replicated with NCCL, 3840x2160 primary frame + path tracing with spp sharded over the ranks.
# usage: python tools/massive_probe.py [grid]      or under torchrun for N GPUs
import json, os, sys, time
import numpy as np, torch, torch.distributed as dist
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(R, 'tests')); sys.path.insert(0, R)
import massivevoxelraytracing_b200 as hk
from massivevoxelraytracing_b200 import dist as hdist
from scenes import dragon_standin, scene_params, procedural_hdri
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
world = int(os.environ.get('WORLD_SIZE', '1')); rank = int(os.environ.get('RANK', '0')); local = int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local); dev = torch.device('cuda', local)
if world > 1: dist.init_process_group('nccl', device_id=dev)
ctx = hk.Context(local)
W, H = 3840, 2160
out = {'grid': grid, 'n_gpus': world, 'image': [W, H]}
g = None
if rank == 0:
    tris, col, emi = dragon_standin(5, 0.15)
    origin, dps = scene_params(tris, grid)
    t0 = time.perf_counter()
    try:
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
    except hk.HakoError as e:
        print('build failed:', e, ctx.build_stats(), 'free/total GB', [x / 1e9 for x in torch.cuda.mem_get_info()], flush=True)
        raise
    out['build_host_call_ms'] = (time.perf_counter() - t0) * 1e3
    st = ctx.build_stats(); info = g.info()
    out['build'] = {k: (round(v, 2) if isinstance(v, float) else int(v)) for k, v in st.items()}
    out['octree'] = {'voxels': int(info.numberOfVoxels), 'nodes': int(info.numberOfNodes), 'wide_index': int(info.wideIndex), 'device_GB': round(info.deviceBytes / 1e9, 2)}
    print('built', out['octree'], out['build'], flush=True)
if world > 1:
    warm = torch.zeros(1, device=dev); dist.broadcast(warm, 0); torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = hdist.broadcast_octree(ctx, hk.IntersectorOctreeGPU, g, 0)
    dist.barrier(); out['nccl_broadcast_ms'] = (time.perf_counter() - t0) * 1e3
cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
# primary frame (rank 0 checks it against the 4096^3 silhouette)
t_d = torch.empty(W * H, dtype=torch.float32, device=dev); nm_d = torch.empty(W * H, dtype=torch.int32, device=dev); vi_d = torch.empty(W * H, dtype=torch.int32, device=dev)
stream = torch.cuda.current_stream()
for _ in range(3): g.render_device(cam, W, H, t=t_d, nMajor=nm_d, vIndex=vi_d, stream=stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(10): g.render_device(cam, W, H, t=t_d, nMajor=nm_d, vIndex=vi_d, stream=stream)
e1.record(stream); torch.cuda.synchronize()
out['primary_ms_per_4k_frame'] = e0.elapsed_time(e1) / 10
out['primary_mrays_per_s_per_gpu'] = W * H / (out['primary_ms_per_4k_frame'] * 1e-3) / 1e6
hit = (t_d != 3.402823466e38)
out['hits'] = int(hit.sum().item()); out['max_voxel_rank_hit'] = int(vi_d[hit].to(torch.int64).remainder(1 << 32).max().item())
# path tracing: 4 steps x 16 spp per rank, one ncclReduce
pt = hk.PathTracer(ctx).setup(); own = pt.m_intersectorOctreeGPU; pt.m_intersectorOctreeGPU = g
pt.maxDepth = 4; pt.resizeFrameBufferIfNeeded(None, W, H); pt.loadHDRI(None, procedural_hdri(2048, 1024))
ptr, nbytes = pt.accum_device(); acc = hdist.as_tensor(ptr, nbytes, dev).view(torch.float32)
iters = hdist.spp_shard(4 * world, rank, world)
def frame():
    pt.clearFrameBuffer()
    for it in iters: pt.step(None, cam, iteration=it)
    if world > 1: hdist.reduce_accum(acc, 0)
frame(); torch.cuda.synchronize()
if world > 1: dist.barrier()
e0.record(stream); frame(); e1.record(stream); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
spp = 64 * world
out['pt'] = {'spp_per_frame': spp, 'ms_per_frame': float(ms.item()), 'ms_per_spp_per_gpu': float(ms.item()) / 64, 'gpaths_per_s': W * H * spp / (float(ms.item()) * 1e-3) / 1e9}
if rank == 0:
    a = pt.accumulators(); out['pt']['spp_count_ok'] = bool((a[..., 3] == spp).all()); out['pt']['mean_radiance'] = float(a[..., :3].mean() / spp)
    print(json.dumps(out), flush=True)
pt.m_intersectorOctreeGPU = own; pt.cleanUp()
if world > 1: dist.destroy_process_group()
