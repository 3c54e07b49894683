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
