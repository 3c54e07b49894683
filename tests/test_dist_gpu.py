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
