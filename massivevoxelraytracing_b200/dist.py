"""Multi-GPU layer: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch) as plumbing.

The reference has no multi-GPU path at all (its only distribution is usecase2_submission/run.py:13-27: two processes on
two hosts split by frame range over SSH).  The path shards naturally: the octree is read-only after the
build and rays are independent, so

  * the octree is built on ONE rank and replicated with a broadcast of its resident arrays (nodes, masks, voxel attributes);
  * primary rays are sharded by image rows / by view, with no data-path collective (ids are disjoint);
  * path tracing shards samples-per-pixel (rank r renders iterations r, r+G, ...) and/or row tiles, and the per-rank
    float4 radiance accumulators are summed with ONE reduce (the w channel carries the spp count, so resolve is unchanged).

Device memory owned by libhako_b200.so is wrapped zero-copy as torch tensors (cuda array interface) so that
torch.distributed can move it; on CPU test runs (gloo) the same functions take host tensors.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from ._abi import OctreeInfo, check


class _DevBuf:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def as_tensor(ptr, nbytes, device):
    """zero-copy uint8 view of library-owned device memory"""
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DevBuf(ptr, nbytes), device=device)


def row_shard(height, rank, world):
    """contiguous row tile [y0, y1) of rank `rank` (image-tile sharding of primary rays / PT tiles)"""
    base, rem = divmod(height, world)
    y0 = rank * base + min(rank, rem)
    return y0, y0 + base + (1 if rank < rem else 0)


def spp_shard(iterations, rank, world):
    """PathTracer iterations (16 spp each) rendered by this rank: r, r+G, r+2G, ... -- the union over ranks is exactly
    the 1-GPU sample set (spp = iteration * 16 + localSpp, voxKernel.cu:642), so images agree up to float summation order"""
    return list(range(rank, iterations, world))


def _info_to_tensor(info):
    return torch.frombuffer(bytearray(bytes(info)), dtype=torch.uint8).clone()


def broadcast_octree(ctx, octree_cls, octree, src=0, group=None):
    """Replicate the octree resident on rank `src` to every rank.  `octree` is the built object on src and may be None
    elsewhere; returns an octree object on every rank.  Three NCCL broadcasts (nodes, masks, attributes) after a small
    header broadcast."""
    rank = dist.get_rank(group)
    device = torch.device("cuda", ctx.device)
    hdr = _info_to_tensor(octree.info()) if rank == src else torch.zeros(C.sizeof(OctreeInfo), dtype=torch.uint8)
    backend = dist.get_backend(group)
    hdr_dev = hdr.to(device) if backend == "nccl" else hdr
    dist.broadcast(hdr_dev, src, group=group)
    info = OctreeInfo.from_buffer_copy(bytes(hdr_dev.cpu().numpy().tobytes()))
    if rank != src:
        h = C.c_void_p()
        check(ctx.lib.hako_octree_create_replica(ctx.h, C.byref(info), C.byref(h)))
        octree = octree_cls(ctx)
        octree._adopt(h)
    for ptr, nbytes in octree.device_arrays():
        if nbytes == 0:
            continue
        t = as_tensor(ptr, nbytes, device)
        dist.broadcast(t, src, group=group)
    torch.cuda.synchronize(device)
    return octree


def reduce_accum(tensor, dst=0, group=None):
    """sum the per-rank float4 accumulators onto rank dst (in place)"""
    dist.reduce(tensor, dst, op=dist.ReduceOp.SUM, group=group)
    return tensor
