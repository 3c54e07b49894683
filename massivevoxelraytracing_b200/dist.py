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


class NativeDist:
    """The multi-GPU layer of the C ABI (include/hako.h: hako_dist_*, csrc/dist.cu) -- NCCL called from C++, no torch tensor on
    the data path.  torch.distributed (any backend, gloo is enough) only ships the 128-byte NCCL id from rank 0 to the others,
    as any launcher of a C++ host would."""

    def __init__(self, ctx, rank=None, world=None, group=None, unique_id=None):
        self.ctx = ctx
        self.lib = ctx.lib
        self.rank = dist.get_rank(group) if rank is None else int(rank)
        self.world = dist.get_world_size(group) if world is None else int(world)
        if unique_id is None:
            box = [None]
            if self.rank == 0:
                buf = (C.c_uint8 * 128)()
                check(self.lib.hako_dist_unique_id(buf))
                box[0] = bytes(buf)
            dist.broadcast_object_list(box, src=0, group=group)
            unique_id = box[0]
        assert len(unique_id) == 128
        self._id = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        h = C.c_void_p()
        check(self.lib.hako_dist_init(ctx.h, self._id, self.rank, self.world, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.hako_dist_destroy(self.h)
            self.h = None

    def nccl_version(self):
        v = C.c_int()
        check(self.lib.hako_dist_info(self.h, None, None, C.byref(v)))
        return v.value

    def barrier(self, stream=None):
        check(self.lib.hako_dist_barrier(self.h, getattr(stream, "cuda_stream", stream)))

    def max_f32(self, value, stream=None):
        v = C.c_float(float(value))
        check(self.lib.hako_dist_max_f32(self.h, getattr(stream, "cuda_stream", stream), C.byref(v)))
        return v.value

    def broadcast_octree(self, octree_cls, octree=None, src=0, stream=None):
        """replicate the octree of rank `src`; returns an octree object on every rank (the source object itself on src).
        The broadcast is enqueued on `stream`; synchronise before timing or destroying the source."""
        out = C.c_void_p()
        check(self.lib.hako_dist_broadcast_octree(self.h, getattr(stream, "cuda_stream", stream), int(src),
                                                  octree.h if (self.rank == src and octree is not None) else None, C.byref(out)))
        if self.rank == src:
            return octree
        o = octree_cls(self.ctx)
        o._adopt(out)
        return o

    def reduce_accum(self, pt, dst=0, stream=None):
        """sum the float4 accumulators of PathTracer `pt` onto rank dst, in place (one ncclReduce)"""
        check(self.lib.hako_dist_reduce_accum(self.h, getattr(stream, "cuda_stream", stream), pt.h, int(dst)))

    def build_sharded(self, octree_cls, tris, col, emi, origin, dps, grid, times=None, stream=None):
        """every rank passes the SAME float32 [T, 3, 3] host arrays; rank r voxelizes triangles [T r / G, T (r+1) / G)"""
        T = int(np.asarray(tris).reshape(-1, 9).shape[0])
        t0, t1 = triangle_shard(T, self.rank, self.world)
        f = lambda a: np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1, 9)[t0:t1])
        p, c, e = f(tris), f(col), f(emi)
        org = np.ascontiguousarray(origin, np.float32)
        ms = (C.c_float * 6)()
        out = C.c_void_p()
        check(self.lib.hako_dist_build_sharded(self.h, getattr(stream, "cuda_stream", stream), p.ctypes.data, c.ctypes.data, e.ctypes.data, t1 - t0,
                                               org.ctypes.data, float(dps), int(grid), C.byref(out), ms))
        if times is not None:
            for k, v in zip(("voxelize_sort_ms", "splitters_ms", "exchange_ms", "sort_unique_ms", "gather_unique_ms", "octree_ms"), ms):
                times[k] = times.get(k, 0.0) + float(v)
        o = octree_cls(self.ctx)
        o._adopt(out)
        return o


def reduce_accum(tensor, dst=0, group=None):
    """sum the per-rank float4 accumulators onto rank dst (in place)"""
    dist.reduce(tensor, dst, op=dist.ReduceOp.SUM, group=group)
    return tensor


# ---------------------------------------------------------------------------------------------------------------------------
# Build sharded over the ranks ("sharding the build by Morton range (one all-to-all of keys by top digits)").
#
#   rank r:  triangles [T r / G, T (r+1) / G)  --voxelize + sort-->  sorted (key, packed attribute) pairs          (own GPU)
#            splitters = weighted quantiles of evenly spaced key samples of every rank                    (all_gather, tiny)
#            pairs with splitter[j-1] <= key < splitter[j] go to rank j                          (ONE all-to-all over NVLink)
#            received runs --sort + unique / integer-mean merge-->  the unique voxels of the rank's Morton range   (own GPU)
#            unique voxels of every range, in range order = the sorted unique list of the scene        (gather, broadcasts)
#            DAG octree from that list on every rank (its node numbering is global, so this stage is replicated)   (own GPU)
#
# Every dumped voxel of a key lands on the rank that owns the key, so the merged attributes are the single-GPU integer means
# and the octree is the single-GPU octree, byte for byte (tests/test_dist_gpu.py; host logic: tests/test_dist_cpu.py on gloo).
# No rank ever holds the whole triangle soup or the whole dumped-voxel list, and no octree broadcast is needed afterwards.
# ---------------------------------------------------------------------------------------------------------------------------
SPLITTER_SAMPLES = 2048
import os as _os
GATHER_MODE = _os.environ.get("HAKO_SHARD_GATHER", "padded")   # "padded": native all-gather of padded ranges; "uneven": grouped broadcasts


def triangle_shard(n_triangles, rank, world):
    """contiguous triangle range [t0, t1) of rank `rank`"""
    return (n_triangles * rank) // world, (n_triangles * (rank + 1)) // world


def choose_splitters(samples, counts, world):
    """samples: [world, S] evenly spaced keys of every rank's sorted pairs (int64); counts: [world] pairs per rank.
    Returns world-1 keys that cut the union into ranges of about equal size (weighted quantiles: a sample of rank r stands for
    counts[r] / S pairs).  Pure host arithmetic -- every rank computes the same splitters from the same gathered input."""
    samples = np.asarray(samples, np.int64)
    counts = np.asarray(counts, np.float64)
    S = samples.shape[1]
    w = np.repeat(counts / max(S, 1), S)
    flat = samples.reshape(-1)
    keep = np.repeat(counts > 0, S)
    flat, w = flat[keep], w[keep]
    if flat.size == 0:
        return np.zeros(world - 1, np.int64)
    order = np.argsort(flat, kind="stable")
    flat, cw = flat[order], np.cumsum(w[order])
    total = cw[-1]
    cuts = [flat[min(int(np.searchsorted(cw, total * j / world, side="left")), flat.size - 1)] for j in range(1, world)]
    return np.maximum.accumulate(np.asarray(cuts, np.int64))


def _all_to_all_v(send, send_counts, recv_counts, group=None):
    """variable all-to-all of a 1-D tensor: NCCL all_to_all_single; on gloo (CPU tests) a loop of broadcasts"""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    recv = torch.empty(int(sum(recv_counts)), dtype=send.dtype, device=send.device)
    if dist.get_backend(group) == "nccl":
        dist.all_to_all_single(recv, send, [int(c) for c in recv_counts], [int(c) for c in send_counts], group=group)
        return recv
    sizes = [None] * world
    dist.all_gather_object(sizes, [int(c) for c in send_counts], group=group)
    roff = np.concatenate([[0], np.cumsum(recv_counts)]).astype(np.int64)
    for src in range(world):
        buf = send if src == rank else torch.empty(int(sum(sizes[src])), dtype=send.dtype, device=send.device)
        if buf.numel():
            dist.broadcast(buf, src, group=group)
        o = int(sum(sizes[src][:rank]))
        recv[roff[src]:roff[src + 1]] = buf[o:o + sizes[src][rank]]
    return recv


class LibStages:
    """the three device stages of a sharded build, on libhako_b200.so (include/hako.h: hako_voxelize_sorted,
    hako_sort_unique_device, hako_octree_build_from_morton_device)"""

    def __init__(self, ctx, octree_cls):
        self.ctx, self.octree_cls = ctx, octree_cls
        self.device = torch.device("cuda", ctx.device)
        self._pairs = None

    def voxelize_sorted(self, tris, col, emi, origin, dps, grid):
        f = lambda a: np.ascontiguousarray(a, np.float32)
        tris, col, emi = f(tris), f(col), f(emi)
        h = C.c_void_p()
        org = np.ascontiguousarray(origin, np.float32)
        check(self.ctx.lib.hako_voxelize_sorted(self.ctx.h, None, tris.ctypes.data, col.ctypes.data, emi.ctypes.data, tris.size // 9,
                                                org.ctypes.data, float(dps), int(grid), C.byref(h)))
        n, k, v = C.c_uint64(), C.c_void_p(), C.c_void_p()
        check(self.ctx.lib.hako_pairs_info(h, C.byref(n), C.byref(k), C.byref(v)))
        self._pairs = h
        keys = as_tensor(k.value or 0, n.value * 8, self.device).view(torch.int64)
        vals = as_tensor(v.value or 0, n.value * 8, self.device).view(torch.int64)
        return keys, vals

    def release_pairs(self):
        if self._pairs is not None:
            self.ctx.lib.hako_pairs_destroy(self._pairs)
            self._pairs = None

    def sort_unique(self, keys, vals, grid):
        U, he = C.c_uint64(), C.c_uint32()
        check(self.ctx.lib.hako_sort_unique_device(self.ctx.h, None, keys.data_ptr() if keys.numel() else None, vals.data_ptr() if vals.numel() else None,
                                                   keys.numel(), int(grid), C.byref(U), C.byref(he)))
        return int(U.value), bool(he.value)

    def build_from_unique(self, mortons, attrs, origin, dps, grid, has_emission):
        h = C.c_void_p()
        org = np.ascontiguousarray(origin, np.float32)
        check(self.ctx.lib.hako_octree_build_from_morton_device(self.ctx.h, None, mortons.data_ptr() if mortons.numel() else None,
                                                                attrs.data_ptr() if attrs.numel() else None, mortons.numel(), org.ctypes.data,
                                                                float(dps), int(grid), 1, int(bool(has_emission)), C.byref(h)))
        o = self.octree_cls(self.ctx)
        o._adopt(h)
        return o

    def alloc_unique(self, total, origin, dps, grid):
        """the octree object with its resident key / attribute arrays (library memory) as int64 tensors to gather into"""
        h, m, a = C.c_void_p(), C.c_void_p(), C.c_void_p()
        org = np.ascontiguousarray(origin, np.float32)
        check(self.ctx.lib.hako_octree_create_for_voxels(self.ctx.h, None, int(total), org.ctypes.data, float(dps), int(grid), C.byref(h), C.byref(m), C.byref(a)))
        self._shell = h
        return (as_tensor(m.value or 0, total * 8, self.device).view(torch.int64), as_tensor(a.value or 0, total * 8, self.device).view(torch.int64))

    def build_nodes(self, has_emission):
        h, self._shell = self._shell, None
        o = self.octree_cls(self.ctx)
        o._adopt(h)
        check(self.ctx.lib.hako_octree_build_nodes(h, None, 1, int(bool(has_emission))))
        return o

    def sync(self):
        torch.cuda.synchronize(self.device)


def build_sharded(stages, tris, col, emi, origin, dps, grid, group=None, times=None):
    """Build the octree of the whole triangle soup with every rank doing 1/G of the voxelization, sort and merge.
    `tris / col / emi`: float32 [T, 3, 3] host arrays -- every rank passes the SAME arrays (or at least its own shard's rows
    valid); `stages`: LibStages (GPU) or a test double with the same four methods.  Returns the octree object of this rank
    (identical on every rank).  `times`, if a dict, receives wall-clock milliseconds per phase (synchronised)."""
    import time
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    T = int(np.asarray(tris).reshape(-1, 9).shape[0])

    def mark(name, t0):
        if times is not None:
            stages.sync()
            times[name] = times.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
        return time.perf_counter()

    t = time.perf_counter()
    t0, t1 = triangle_shard(T, rank, world)
    sl = lambda a: np.asarray(a, np.float32).reshape(-1, 9)[t0:t1]
    keys, vals = stages.voxelize_sorted(sl(tris), sl(col), sl(emi), origin, dps, grid)
    dev = keys.device
    t = mark("voxelize_sort_ms", t)

    # splitters from evenly spaced samples of every rank's sorted keys
    n = keys.numel()
    S = SPLITTER_SAMPLES
    if n:
        idx = (torch.arange(S, dtype=torch.int64, device=dev) * (n - 1)) // (S - 1)   # integer arithmetic: float32 cannot index 10^8 pairs
        mine = torch.cat([keys[idx], torch.tensor([n], dtype=torch.int64, device=dev)])
    else:
        mine = torch.zeros(S + 1, dtype=torch.int64, device=dev)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    g = torch.stack(gathered).cpu().numpy()
    splitters = choose_splitters(g[:, :S], g[:, S], world)
    cuts = torch.searchsorted(keys, torch.from_numpy(splitters).to(dev), right=False) if n else torch.zeros(world - 1, dtype=torch.int64, device=dev)
    bounds = np.concatenate([[0], cuts.cpu().numpy(), [n]]).astype(np.int64)
    send_counts = np.diff(bounds)
    cm = [torch.empty(world, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(cm, torch.from_numpy(send_counts).to(dev), group=group)
    recv_counts = torch.stack(cm).cpu().numpy()[:, rank]
    t = mark("splitters_ms", t)

    rk = _all_to_all_v(keys, send_counts, recv_counts, group)
    rv = _all_to_all_v(vals, send_counts, recv_counts, group)
    del keys, vals
    if hasattr(stages, "release_pairs"):
        stages.sync()
        stages.release_pairs()
    t = mark("exchange_ms", t)

    U, he = stages.sort_unique(rk, rv, grid)
    t = mark("sort_unique_ms", t)

    meta = [torch.empty(2, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(meta, torch.tensor([U, int(he)], dtype=torch.int64, device=dev), group=group)
    meta = torch.stack(meta).cpu().numpy()
    offs = np.concatenate([[0], np.cumsum(meta[:, 0])]).astype(np.int64)
    total = int(offs[-1])
    in_place = hasattr(stages, "alloc_unique")   # LibStages: gather straight into the octree's own arrays
    if in_place:
        mortons, attrs = stages.alloc_unique(total, origin, dps, grid)
    else:
        mortons = torch.empty(total, dtype=torch.int64, device=dev)
        attrs = torch.empty(total, dtype=torch.int64, device=dev)
    mortons[offs[rank]:offs[rank + 1]] = rk[:U]
    attrs[offs[rank]:offs[rank + 1]] = rv[:U]
    del rk, rv
    parts_m = [mortons[offs[r]:offs[r + 1]] for r in range(world)]
    parts_a = [attrs[offs[r]:offs[r + 1]] for r in range(world)]
    if dist.get_backend(group) == "nccl" and world > 1 and GATHER_MODE == "padded":
        # native NCCL all-gather needs equal pieces: pad every range to the longest (the splitters keep them within a few per cent of
        # each other) and copy the pieces to their places -- measured faster than the grouped broadcasts of an uneven all-gather
        pad = int(meta[:, 0].max())
        for dst, parts in ((mortons, parts_m), (attrs, parts_a)):
            mine = torch.empty(pad, dtype=torch.int64, device=dev)
            mine[:U] = parts[rank]
            wide = torch.empty(world * pad, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(wide, mine, group=group)
            for r in range(world):
                if r != rank:
                    parts[r].copy_(wide[r * pad:r * pad + int(meta[r, 0])])
            del wide, mine
    elif dist.get_backend(group) == "nccl" and world > 1:
        # uneven all-gather: torch issues the per-rank broadcasts as one NCCL group, so they share the links
        dist.all_gather(parts_m, parts_m[rank], group=group)
        dist.all_gather(parts_a, parts_a[rank], group=group)
    else:
        for r in range(world):
            if offs[r + 1] > offs[r]:
                dist.broadcast(parts_m[r], r, group=group)
                dist.broadcast(parts_a[r], r, group=group)
    t = mark("gather_unique_ms", t)

    if in_place:
        stages.sync()
        del mortons, attrs, parts_m, parts_a
        octree = stages.build_nodes(bool(meta[:, 1].any()))
    else:
        octree = stages.build_from_unique(mortons, attrs, origin, dps, grid, bool(meta[:, 1].any()))
    mark("octree_ms", t)
    if times is not None:
        times["dumped_voxels_this_rank"] = int(n)
        times["range_voxels_this_rank"] = int(U)
        times["unique_voxels"] = total
    return octree
