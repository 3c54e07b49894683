"""GPU tests of the multi-GPU layer over NCCL: the sharded build (triangles sharded, pairs exchanged by Morton range with one
all-to-all, unique voxels gathered, DAG built on every rank) produces the bytes of the single-GPU hako_octree_build.
world_size 1 runs on any GPU box; world_size 2 / 4 need that many devices and are skipped otherwise."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, grid, levels, q, n_tris=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import massivevoxelraytracing_b200 as hk
    from massivevoxelraytracing_b200 import dist as hdist
    from scenes import bunny_scene, dragon_standin, scene_params
    tris, col, emi = bunny_scene() if levels == 0 else dragon_standin(levels, 0.15)
    origin, dps = scene_params(tris, grid)
    if n_tris is not None:                       # fewer triangles than ranks: empty shards, possibly empty Morton ranges
        tris, col, emi = tris[:n_tris], col[:n_tris], emi[:n_tris]
    ctx = hk.Context(rank)
    times = {}
    g = hdist.build_sharded(hdist.LibStages(ctx, hk.IntersectorOctreeGPU), tris, col, emi, origin, dps, grid, times=times)
    nodes, mk, at = g.export(True, True, True)
    info = g.info()
    W, H = 320, 200
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    _, t, nm, vi = g.render(cam, W, H, rgba=False)
    out = dict(rank=rank, fnv=[int(hash(a.tobytes())) for a in (nodes, mk, at, t, vi)], n=(int(info.numberOfNodes), int(info.numberOfVoxels), int(info.hasEmission)), times=times)
    if rank == 0:
        ref = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
        rn, rm, ra = ref.export(True, True, True)
        _, rt, _, rvi = ref.render(cam, W, H, rgba=False)
        out["equal"] = bool(np.array_equal(nodes.view(np.uint8), rn.view(np.uint8)) and np.array_equal(mk, rm) and np.array_equal(at.view(np.uint8), ra.view(np.uint8))
                            and np.array_equal(t.view(np.uint32), rt.view(np.uint32)) and np.array_equal(vi, rvi)
                            and int(ref.info().hasEmission) == int(info.hasEmission))
        out["voxels"] = int(mk.size)
        ref.cleanUp()
    g.cleanUp()
    q.put(out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,grid,levels", [(1, 256, 0), (2, 256, 0), (2, 1024, 2), (4, 512, 1)])
def test_sharded_build_equals_single_gpu_build(world, grid, levels):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29700 + (os.getpid() % 1000) + world
    procs = [ctxm.Process(target=_worker, args=(r, world, port, grid, levels, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    r0 = [r for r in res if r["rank"] == 0][0]
    assert r0["equal"] and r0["voxels"] > 100_000
    for r in res:
        assert r["n"] == r0["n"]                       # every rank built the same tree ...
        assert r["times"]["unique_voxels"] == r0["voxels"]
    assert sum(r["times"]["range_voxels_this_rank"] for r in res) == r0["voxels"]
    if world > 1:
        # ... from a balanced split: no Morton range holds more than twice its share
        assert max(r["times"]["range_voxels_this_rank"] for r in res) < 2.0 * r0["voxels"] / world


def test_sharded_build_with_an_empty_shard():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29760 + (os.getpid() % 1000)
    procs = [ctxm.Process(target=_worker, args=(r, 2, port, 256, 0, q, 1)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    r0 = [r for r in res if r["rank"] == 0][0]
    assert r0["equal"] and r0["voxels"] > 0
    assert res[0]["n"] == res[1]["n"]


# ---------------------------------------------------------------------------------------------------------------------------
# The north-star partition through the C ABI (include/hako.h: hako_dist_*; NCCL called from C++, csrc/dist.cu):
# build on rank 0 -> hako_dist_broadcast_octree -> primary rays sharded by rows (no collective) and path tracing sharded by spp
# with ONE hako_dist_reduce_accum -> equal to the 1-GPU results.  torch.distributed (gloo) only carries the 128-byte NCCL id.
# ---------------------------------------------------------------------------------------------------------------------------
def _digest(a):
    import hashlib
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()   # hash() is salted per process


def _partition_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import massivevoxelraytracing_b200 as hk
    from massivevoxelraytracing_b200 import dist as hdist
    from scenes import bunny_scene, procedural_hdri, scene_params
    grid, W, H, iters, depth = 512, 640, 360, 4 * world, 4
    tris, col, emi = bunny_scene()
    origin, dps = scene_params(tris, grid)
    ctx = hk.Context(rank)
    nd = hdist.NativeDist(ctx)
    src = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid) if rank == 0 else None
    g = nd.broadcast_octree(hk.IntersectorOctreeGPU, src, 0)
    ctx.synchronize()
    out = dict(rank=rank)
    # replica == source: canonical nodes and attributes
    nodes, _, attrs = g.export(True, False, True)
    out["tree"] = (_digest(nodes), _digest(attrs), int(g.info().numberOfNodes), int(g.info().numberOfVoxels), int(g.info().hasEmission))
    # primary rays: this rank's row tile
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H)
    y0, y1 = hdist.row_shard(H, rank, world)
    rgba, t, nm, vi = g.render(cam, W, H, y0, y1)
    out["rows"] = (y0, y1, rgba, t, nm, vi)
    # path tracing: spp shard + one reduce
    env = procedural_hdri(256, 128)
    pt = hk.PathTracer(ctx).setup()
    pt.m_intersectorOctreeGPU = g
    pt.maxDepth = depth
    pt.resizeFrameBufferIfNeeded(None, W, H)
    pt.loadHDRI(None, env, env)
    pcam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.01)
    for it in hdist.spp_shard(iters, rank, world):
        pt.step(None, pcam, iteration=it)
    nd.reduce_accum(pt, 0)
    ctx.synchronize()
    if rank == 0:
        out["acc"] = pt.accumulators().copy()
        # the 1-GPU answers on the same device
        pt.clearFrameBuffer()
        for it in range(iters):
            pt.step(None, pcam, iteration=it)
        out["acc1"] = pt.accumulators().copy()
        out["full"] = g.render(cam, W, H)
    # the native sharded build: byte-identical to the single-GPU build
    times = {}
    gs = nd.build_sharded(hk.IntersectorOctreeGPU, tris, col, emi, origin, dps, grid, times=times)
    n2, m2, a2 = gs.export(True, True, True)
    out["sharded"] = (_digest(n2), _digest(a2), int(gs.info().numberOfNodes), int(gs.info().numberOfVoxels), int(gs.info().hasEmission))
    out["sharded_mortons_sorted"] = bool((np.diff(m2.astype(np.int64)) > 0).all())
    out["times"] = times
    gs.cleanUp()
    pt.m_intersectorOctreeGPU = hk.IntersectorOctreeGPU(ctx)
    pt.cleanUp()
    nd.barrier()
    if rank != 0:
        g.cleanUp()
    q.put(out)
    dist.barrier()
    if rank == 0:
        g.cleanUp()
    nd.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4])
def test_north_star_partition_equals_single_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29800 + (os.getpid() % 1000) + world
    procs = [ctxm.Process(target=_partition_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=600) for _ in range(world)], key=lambda r: r["rank"])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    r0 = res[0]
    for r in res:
        assert r["tree"] == r0["tree"]            # replica export() == source
        assert r["sharded"] == r0["tree"]         # the sharded build is the single-GPU build, byte for byte, on every rank
        assert r["sharded_mortons_sorted"]
    # row tiles concatenate to the full frame, bit for bit
    full = r0["full"]
    for k in range(4):
        cat = np.concatenate([r["rows"][2 + k] for r in res], 0)
        assert np.array_equal(cat.view(np.uint8), full[k].view(np.uint8))
    assert [r["rows"][0] for r in res][1:] == [r["rows"][1] for r in res][:-1]
    # spp-sharded accumulators: identical sample set, float summation order differs (per-rank partial sums, then the reduce)
    a, b = r0["acc"].astype(np.float64), r0["acc1"].astype(np.float64)
    assert np.array_equal(a[..., 3], b[..., 3]) and (a[..., 3] == 16 * 4 * world).all()
    err = np.abs(a[..., :3] - b[..., :3]) / np.maximum(np.abs(b[..., :3]), 1e-3)
    assert err.max() <= 1e-5, err.max()
    if world == 1:
        assert np.array_equal(r0["acc"].view(np.uint32), r0["acc1"].view(np.uint32))
