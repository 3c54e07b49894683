"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard maps and the accumulator reduction."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_maps_cover_exactly():
    from massivevoxelraytracing_b200.dist import row_shard, spp_shard
    for h in (1, 7, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            rows = []
            for r in range(world):
                y0, y1 = row_shard(h, r, world)
                rows += list(range(y0, y1))
            assert rows == list(range(h))
    for it in (0, 1, 8, 13):
        for world in (1, 2, 4, 8):
            allit = sorted(sum((spp_shard(it, r, world) for r in range(world)), []))
            assert allit == list(range(it))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from massivevoxelraytracing_b200 import dist as hdist
    from massivevoxelraytracing_b200._abi import OctreeInfo
    # accumulator reduce: rank r holds the samples of its spp shard; the sum must equal the 1-rank accumulation
    W, H, iters = 8, 4, 6
    rng = np.random.default_rng(0)
    per_iter = rng.random((iters, H, W, 4)).astype(np.float32)
    per_iter[..., 3] = 16.0
    mine = torch.zeros(H, W, 4)
    for i in hdist.spp_shard(iters, rank, world):
        mine += torch.from_numpy(per_iter[i])
    hdist.reduce_accum(mine, 0)
    # header broadcast round trip (what broadcast_octree sends first)
    info = OctreeInfo()
    if rank == 0:
        info.numberOfNodes, info.numberOfVoxels, info.dps, info.gridRes, info.hasEmission = 1234, 56789, 0.125, 4096, 1
    hdr = hdist._info_to_tensor(info)
    dist.broadcast(hdr, 0)
    got = OctreeInfo.from_buffer_copy(hdr.numpy().tobytes())
    ok_hdr = (got.numberOfNodes, got.numberOfVoxels, got.dps, got.gridRes, got.hasEmission) == (1234, 56789, 0.125, 4096, 1)
    if rank == 0:
        q.put((mine.numpy(), per_iter.sum(0), ok_hdr))
    else:
        q.put(ok_hdr)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_reduce_and_header():
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29500 + (os.getpid() % 1000)
    procs = [ctxm.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r0 = [r for r in res if isinstance(r, tuple)][0]
    other = [r for r in res if not isinstance(r, tuple)][0]
    np.testing.assert_allclose(r0[0], r0[1], rtol=1e-6)
    assert (r0[0][..., 3] == 96.0).all()
    assert r0[2] and other


# ---- sharded build: the host logic of hdist.build_sharded on gloo, with the CPU oracle standing in for the device stages ----
class _OracleStages:
    """test double of hdist.LibStages: the same four methods on host tensors, computed by the oracle"""

    def __init__(self, oracle):
        self.o = oracle

    def voxelize_sorted(self, tris, col, emi, origin, dps, grid):
        if tris.shape[0] == 0:
            return torch.zeros(0, dtype=torch.int64), torch.zeros(0, dtype=torch.int64)
        keys, c, e = self.o.voxelize(tris.reshape(-1, 3, 3), col.reshape(-1, 3, 3), emi.reshape(-1, 3, 3), origin, float(dps), grid)
        vals = np.ascontiguousarray(np.concatenate([c, e], axis=1)).view(np.uint64).reshape(-1)   # colour | emission << 32
        order = np.argsort(keys, kind="stable")
        return torch.from_numpy(keys[order].view(np.int64).copy()), torch.from_numpy(vals[order].view(np.int64).copy())

    def sort_unique(self, keys, vals, grid):
        if keys.numel() == 0:
            return 0, False
        k = keys.numpy().view(np.uint64)
        order = np.argsort(k, kind="stable")
        a = np.ascontiguousarray(vals.numpy()[order]).view(np.uint8).reshape(-1, 8)
        mk, mc, me = self.o.merge_voxels(k[order], a[:, :4].copy(), a[:, 4:].copy())
        U = mk.size
        keys[:U] = torch.from_numpy(mk.view(np.int64).copy())
        vals[:U] = torch.from_numpy(np.ascontiguousarray(np.concatenate([mc, me], axis=1)).view(np.int64).reshape(-1))
        return U, bool(me[:, :3].any())

    def build_from_unique(self, mortons, attrs, origin, dps, grid, has_emission):
        return mortons.numpy().view(np.uint64).copy(), attrs.numpy().copy().view(np.uint8).reshape(-1, 8), has_emission

    def sync(self):
        pass


def _shard_worker(rank, world, port, q, n_tris=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from massivevoxelraytracing_b200 import dist as hdist
    from oracle import orc
    from scenes import bunny_scene, scene_params
    o = orc.Oracle("port")
    tris, col, emi = bunny_scene()
    grid = 64
    origin, dps = scene_params(tris, grid)
    if n_tris is not None:                       # a mesh with fewer triangles than ranks: some shards are empty
        tris, col, emi = tris[:n_tris], col[:n_tris], emi[:n_tris]
    times = {}
    mk, at, he = hdist.build_sharded(_OracleStages(o), tris, col, emi, origin, dps, grid, times=times)
    q.put((rank, mk, at, he, times))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_sharded_build_equals_single_rank(world, oracle, bunny):
    """every rank ends with the sorted unique voxel list (keys + integer-mean attributes) of the whole mesh: the shards are
    voxelized separately, exchanged by Morton range (splitters from samples), merged per range and gathered"""
    from scenes import scene_params
    tris, col, emi = bunny
    grid = 64
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29600 + (os.getpid() % 1000) + world
    procs = [ctxm.Process(target=_shard_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ranges = []
    for rank, gk, ga, he, times in res:
        assert np.array_equal(gk, mk)
        assert np.array_equal(ga[:, :4], mc) and np.array_equal(ga[:, 4:], me)
        assert he == bool(me[:, :3].any())
        ranges.append(times["range_voxels_this_rank"])
        assert times["unique_voxels"] == mk.size
    # the splitters balance the ranges (bunny at 64^3: within a factor of two of the mean)
    assert sum(ranges) == mk.size and max(ranges) < 2.0 * mk.size / world


def test_choose_splitters_and_shards():
    from massivevoxelraytracing_b200.dist import choose_splitters, triangle_shard
    for T in (0, 1, 7, 5110):
        for world in (1, 2, 3, 8):
            cover = []
            for r in range(world):
                t0, t1 = triangle_shard(T, r, world)
                cover += list(range(t0, t1))
            assert cover == list(range(T))
    rng = np.random.default_rng(0)
    a = np.sort(rng.integers(0, 1 << 40, (4, 256)), axis=1)
    s = choose_splitters(a, [1000, 1000, 1000, 1000], 4)
    assert s.shape == (3,) and (np.diff(s) >= 0).all()
    frac = [(a < x).mean() for x in s]
    assert np.allclose(frac, [0.25, 0.5, 0.75], atol=0.03)
    # a rank without pairs carries no weight; all-empty input is legal
    s2 = choose_splitters(np.stack([a[0], np.zeros(256, np.int64)]), [1000, 0], 2)
    assert abs((a[0] < s2[0]).mean() - 0.5) < 0.03
    assert choose_splitters(np.zeros((2, 16), np.int64), [0, 0], 2).tolist() == [0]


def test_gloo_sharded_build_with_empty_shards(oracle, bunny):
    """two triangles over three ranks: one shard is empty, and so may be a Morton range"""
    from scenes import scene_params
    tris, col, emi = bunny
    grid = 64
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris[:2], col[:2], emi[:2], origin, float(dps), grid)
    mk, mc, me = oracle.merge_voxels(keys, c, e)
    world = 3
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29650 + (os.getpid() % 1000)
    procs = [ctxm.Process(target=_shard_worker, args=(r, world, port, q, 2)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, gk, ga, he, times in res:
        assert np.array_equal(gk, mk) and np.array_equal(ga[:, :4], mc) and np.array_equal(ga[:, 4:], me)
