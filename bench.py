#!/usr/bin/env python
"""bench.py -- headline benchmark of the massive-voxel hot path on B200.

Metric (BASELINE.json): Mrays/s of primary-ray octree traversal; the voxelize + sort + unique + octree build time in ms is
reported beside it.  Workload at N=1 = BASELINE config[1]: the (synthetic stand-in for the) xyzrgb dragon voxelized at
4096^3, 1920x1080 pixel-centre primary rays.  A "step" casts VIEWS full frames from cameras orbiting the model (view 0 is
voxRT's default camera, voxRT.cpp:69-72), so one step is ~33 M rays and is not launch-bound.

  value      rays / s with the octree resident in HBM (device outputs), CUDA events on the launching stream
  e2e        same rays through the host-buffer C-ABI call (hako_render_primary_views, the frames of a step in one call): the
             cameras go in by value and the RGBA8 frames (4 B per ray, what voxRTGPU downloads per frame) come back to pinned
             host memory inside the timed region; e2e.hit_records is the same call returning t / nMajor / vIndex (12 B per
             ray), e2e.per_frame_call the one-synchronous-call-per-frame form (hako_render_primary)
  roofline   dominant kernel = renderPrimaryKernel; algorithmic bytes = rays * (12 + 8 * descents/ray)
  cpu_baseline / --impl reference   the reference's own CPU traversal (oracle/_ref when present, else the oracle port)
             on the box's host cores, on a bounded sample of the same views

Multi-GPU (torchrun, one rank per GPU): the octree is built on rank 0 and broadcast over NCCL; every rank then casts its
own VIEWS views (weak scaling: per-GPU work fixed, no data-path collective -- rays are independent).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GRID = 4096
W, H = 1920, 1080
VIEWS = 384              # full frames per step per GPU: 796 M rays = ~50 ms of GPU work, so 20 timed steps are ~1 s
RING = 16                # device output sets kept per GPU (16 x 24.9 MB = 398 MB > the 126 MB L2); view i writes set i % RING
CHUNK = 64               # views per hako_render_primary_views call of the e2e leg (64 pinned RGBA frames = 531 MB)
REF_BANDS = 16           # the CPU arm casts row band (k % 16) of view k: 1/16 of the rays of a step, a bounded sample of every view
STRONG_VIEWS = 128       # strong-scaling block: this many views in TOTAL, split over the ranks
STRONG_SPP_STEPS = 32    # strong-scaling block: one 512-spp 1080p frame (32 PathTracer steps) in TOTAL, split over the ranks
MESH_LEVELS = 5          # bunny x 4^5 = 5.23 M triangles (the dragon blob is missing: SURVEY.md 8d)


def make_scene():
    from scenes import dragon_standin, scene_params
    t0 = time.time()
    tris, col, emi = dragon_standin(MESH_LEVELS, 0.15)
    origin, dps = scene_params(tris, GRID)
    return tris, col, emi, origin, float(dps), time.time() - t0


def view_eye(k, n_views_total):
    """orbit of voxRT's default eye (4,4,4) about the y axis; view 0 is the reference camera"""
    a = 2.0 * np.pi * k / n_views_total
    c, s = np.cos(a), np.sin(a)
    return (float(4 * c + 4 * s), 4.0, float(-4 * s + 4 * c))


def rank_views(rank, world):
    """the views of a step are dealt round-robin: every rank's 16 views span the whole orbit, so the ranks do the same work"""
    return [rank + i * world for i in range(VIEWS)]


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.rows)}


def load_oracle():
    """the CPU reference: oracle/_ref (the reference's own headers) when it travelled with the snapshot, else the port"""
    from oracle import orc
    port = orc.Oracle("port")
    ref = orc.Oracle("reference") if orc.ref_available() else None
    return port, ref


def cpu_tree(port, ref, tris, col, emi, origin, dps):
    """reference CPU pipeline (voxRT.cpp:198-270): voxelize -> mergeVoxels -> buildDAGReference, timed per stage.  Every stage runs
    the reference's own code (oracle/_ref) when it travelled with the snapshot, else the port; `kind` says which."""
    o = ref if ref is not None else port
    t = {"kind": "reference" if ref is not None else "port"}
    t0 = time.time()
    keys, c, e = o.voxelize(tris, col, emi, origin, dps, GRID)
    t["voxelize_ms"] = (time.time() - t0) * 1e3
    t0 = time.time()
    mk, mc, me = o.merge_voxels(keys, c, e)
    t["sort_unique_ms"] = (time.time() - t0) * 1e3
    del keys, c, e
    t0 = time.time()
    tree = o.build_octree(mk, origin, dps, GRID)
    t["octree_ms"] = (time.time() - t0) * 1e3
    t["total_ms"] = t["voxelize_ms"] + t["sort_unique_ms"] + t["octree_ms"]
    return o, tree, mk, t


def cpu_cast(o, tree, views_total, view_ids, threads, bands, stats=None):
    """cast row band (k % REF_BANDS) of view k (bands=True: 1/16 of each view) or whole views; returns rays, seconds.
    Output buffers are allocated once, outside the timed calls."""
    import ctypes as C
    from oracle.orc import FLT_MAX
    t = np.full(W * H, FLT_MAX, np.float32)
    nm = np.full(W * H, -1, np.int32)
    vi = np.full(W * H, 0xFFFFFFFF, np.uint32)
    st = np.zeros(4, np.uint64)
    rays = 0
    secs = 0.0
    for i, k in enumerate(view_ids):
        cam = o.camera_lookat(view_eye(k, views_total), (0, 0, 0), (0, 1, 0), None, W, H)
        if bands:
            j = i % REF_BANDS
            y0, y1 = (H * j) // REF_BANDS, (H * (j + 1)) // REF_BANDS
        else:
            y0, y1 = 0, H
        t0 = time.time()
        o.lib.orc_render_primary(tree.h, C.byref(cam), W, H, y0, y1, threads, t, nm, vi, st.ctypes.data if stats is not None else None)
        secs += time.time() - t0
        rays += (y1 - y0) * W
        if stats is not None:
            stats[:3] += st[:3]
            stats[3] = max(stats[3], st[3])
    return rays, secs


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def base_config(n_gpus):
    return {"workload": "config[1]: xyzrgb-dragon stand-in (bunny.obj midpoint-subdivided x%d + hash noise, %d triangles; the dragon blob is "
                        "missing from the reference checkout) voxelized at %d^3, %dx%d pixel-centre primary rays, %d orbit views per step per GPU"
                        % (MESH_LEVELS, 5110 * 4 ** MESH_LEVELS, GRID, W, H, VIEWS),
            "grid": GRID, "image": [W, H], "views_per_step_per_gpu": VIEWS,
            "reference_arm_sample": "--impl reference casts row band (k mod %d) of view k for the same %d views: 1/%d of the rays of a step (a rate, so the ratio stands)" % (REF_BANDS, VIEWS, REF_BANDS),
            "l2": "the outputs of a step cycle through %d sets of 24.9 MB (%.0f MB > the 126 MB L2); the octree is L2-resident by design; no explicit flush" % (RING, RING * 24.9),
            "parallelism": "octree built on rank 0 + NCCL broadcast; views dealt round-robin over %d GPU(s) (every rank spans the whole orbit), no data-path collective" % n_gpus,
            "streams": "the frames of a step alternate between CUDA streams (4 per GPU unless HAKO_BENCH_STREAMS says otherwise; tail of frame i overlaps the head of frame i+1); "
                       "timed with events on the stream that forks and joins them"}


def run_reference(args):
    """--impl reference: the reference's own CPU traversal on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    tris, col, emi, origin, dps, _ = make_scene()
    port, ref = load_oracle()
    o, tree, mk, build = cpu_tree(port, ref, tris, col, emi, origin, dps)
    views_total = VIEWS * args.gpus
    # one step = a bounded sample of the workload's step: row band k of each of the VIEWS views (= one frame's worth of rays)
    for _ in range(args.warmup):
        cpu_cast(o, tree, views_total, rank_views(0, views_total // VIEWS), threads, True)
    rays = 0
    secs = 0.0
    for _ in range(args.steps):
        r, s = cpu_cast(o, tree, views_total, rank_views(0, views_total // VIEWS), threads, True)
        rays += r
        secs += s
    value = rays / secs / 1e6
    sample = "row band (k mod %d) of view k for the %d views of a step (%d rays per step = 1/%d of the step), %d host threads over image rows" % (
        REF_BANDS, VIEWS, rays // max(args.steps, 1), REF_BANDS, threads)
    line = {"impl": "reference", "metric": "Mrays/s (octree traversal, primary rays)", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / max(args.steps, 1) * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": base_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "reference" if ref is not None else "port", "sample": sample,
                             "cpu": cpu_model(), "build_ms": build},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "unique_voxels": int(mk.size), "octree_nodes": tree.num_nodes}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


class _TorchDist:
    """fallback with the interface of dist.NativeDist over torch.distributed (used only if the C-ABI NCCL layer fails to start)"""

    def __init__(self, hdist, ctx, dev):
        self.hdist, self.ctx, self.dev, self.kind = hdist, ctx, dev, "torch.distributed fallback"

    def broadcast_octree(self, octree_cls, octree=None, src=0, stream=None):
        return self.hdist.broadcast_octree(self.ctx, octree_cls, octree, src)

    def reduce_accum(self, pt, dst=0, stream=None):
        import torch
        ptr, nbytes = pt.accum_device()
        self.hdist.reduce_accum(self.hdist.as_tensor(ptr, nbytes, self.dev).view(torch.float32), dst)

    def build_sharded(self, octree_cls, tris, col, emi, origin, dps, grid, times=None, stream=None):
        return self.hdist.build_sharded(self.hdist.LibStages(self.ctx, octree_cls), tris, col, emi, origin, dps, grid, times=times)

    def close(self):
        pass


def run_ours(args):
    import torch
    import torch.distributed as dist
    import massivevoxelraytracing_b200 as hk
    from massivevoxelraytracing_b200 import dist as hdist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = hk.Context(local)
    # the multi-GPU layer of the C ABI (hako_dist_*: NCCL called from C++); torch.distributed only ships the 128-byte id and
    # synchronises the timed regions
    nd = None
    if world > 1:
        try:
            nd = hdist.NativeDist(ctx)
        except Exception as exc:             # the bench line must survive: fall back to torch.distributed for the collectives
            print("NativeDist unavailable, using torch.distributed: %r" % (exc,), file=sys.stderr)
            nd = _TorchDist(hdist, ctx, dev)
    views_total = VIEWS * world
    nrays_view = W * H
    stream = torch.cuda.current_stream()

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def device_ms(fn, reps=1):
        """device time of fn() on the current stream, events on that stream, max over ranks"""
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        sync_all()
        return max_over_ranks(a.elapsed_time(b) / reps)

    # ---- build (rank 0), host-buffer call = the e2e build; then replicate ----
    build = None
    octree = None
    scene = None
    multi = {}
    if rank == 0:
        scene = make_scene()
        tris, col, emi, origin, dps, gen_s = scene
        wall = []
        for i in range(3):
            if octree is not None:
                octree.cleanUp()
            t0 = time.perf_counter()
            octree = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, dps, GRID)
            wall.append((time.perf_counter() - t0) * 1e3)
        st = ctx.build_stats()
        build = {"e2e_host_call_ms": min(wall), "e2e_host_call_ms_first": wall[0],
                 "device_ms": {k: round(st[k], 3) for k in ("ms_h2d", "ms_voxelize", "ms_sort", "ms_unique", "ms_octree", "ms_total")},
                 "triangles": int(st["nTriangles"]), "dumped_voxels": int(st["dumpedVoxels"]), "unique_voxels": int(st["uniqueVoxels"]),
                 "octree_nodes": int(st["numberOfNodes"]), "sort_passes": int(st["sortPasses"]), "host_syncs": int(st["hostSyncs"]),
                 "kernel_launches": int(st["kernelLaunches"]), "mesh_generation_s": round(gen_s, 2)}
    if world > 1:
        # replicate: hako_dist_broadcast_octree (64-byte header, then nodes + masks + attributes as one NCCL group).  A first call
        # sizes the receivers' memory pools; the timed one is the second (device time on the stream, max over ranks)
        src = octree
        rep = nd.broadcast_octree(hk.IntersectorOctreeGPU, src, 0)
        sync_all()
        if rank != 0:
            rep.cleanUp()
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(stream)
        octree = nd.broadcast_octree(hk.IntersectorOctreeGPU, src, 0, stream=stream)
        b.record(stream)
        sync_all()
        wall_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        bc_ms = max_over_ranks(a.elapsed_time(b))
        bytes_ = sum(nb for _, nb in octree.device_arrays())
        multi["octree_broadcast"] = {"device_ms": round(bc_ms, 3), "wall_ms": round(wall_ms, 3), "MB": round(bytes_ / 1e6, 1), "GBps": round(bytes_ / bc_ms / 1e6, 1)}
        # the alternative to build-on-rank-0 + broadcast: hako_dist_build_sharded -- every rank voxelizes / sorts / merges 1/N of the
        # scene, ONE all-to-all of the voxel pairs by Morton range, unique voxels gathered, DAG built on every rank (same bytes)
        try:
            if os.environ.get("HAKO_BENCH_SHARDED", "1") == "0":
                raise RuntimeError("skipped (HAKO_BENCH_SHARDED=0)")
            if scene is None:
                scene = make_scene()            # deterministic generator: every rank derives the same triangles
            tris, col, emi, origin, dps, _ = scene
            best = None
            for _ in range(3):
                times = {}
                sync_all()
                t0 = time.perf_counter()
                sh = nd.build_sharded(hk.IntersectorOctreeGPU, tris, col, emi, origin, dps, GRID, times=times)
                sync_all()
                wall_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
                same = int(sh.info().numberOfNodes == octree.info().numberOfNodes and sh.info().numberOfVoxels == octree.info().numberOfVoxels)
                sh.cleanUp()
                if best is None or wall_ms < best[0]:
                    best = (wall_ms, {k: round(v, 2) for k, v in times.items() if k.endswith("_ms")}, same)
            ok = torch.tensor([best[2]], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            multi["sharded_build"] = {"e2e_ms": round(best[0], 2), "phases_rank0_ms": best[1], "same_tree_on_every_rank": bool(ok.item())}
        except Exception as exc:                # the bench line must survive a failure of this optional leg
            multi["sharded_build"] = {"error": repr(exc)[:200]}
    info = octree.info()

    # ---- the timed step: VIEWS full frames per GPU, device-resident ----
    my_views = rank_views(rank, world)
    cams = [hk.camera_lookat(view_eye(k, views_total), (0, 0, 0), width=W, height=H) for k in my_views]
    t_d = torch.empty((RING, nrays_view), dtype=torch.float32, device=dev)
    nm_d = torch.empty((RING, nrays_view), dtype=torch.int32, device=dev)
    vi_d = torch.empty((RING, nrays_view), dtype=torch.int32, device=dev)
    # frames alternate between streams (what hako_render_primary_views does inside): a frame is one wave of CTAs whose
    # tail -- the few long rays -- would otherwise idle most SMs before the next frame starts
    n_streams = max(1, int(os.environ.get("HAKO_BENCH_STREAMS", "4")))
    sides = [torch.cuda.Stream(device=dev) for _ in range(n_streams - 1)]
    fork, joins = torch.cuda.Event(), [torch.cuda.Event() for _ in sides]

    def cast_views(cam_list):
        fork.record(stream)
        for sd in sides:
            sd.wait_event(fork)
        for i, cam in enumerate(cam_list):
            k = i % n_streams
            octree.render_device(cam, W, H, t=t_d[i % RING], nMajor=nm_d[i % RING], vIndex=vi_d[i % RING], stream=sides[k - 1] if k else stream)
        for sd, j in zip(sides, joins):
            j.record(sd)
            stream.wait_event(j)

    def step_device():
        cast_views(cams)

    for _ in range(max(args.warmup, 3)):
        step_device()
    sampler = ClockSampler(local)
    sync_all()
    sampler.start()
    l0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    sync_all()
    launches = ctx.launch_count - l0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    # one frame launched ALONE (no overlap with its neighbours): the isolated-launch figure beside the overlapped one
    alone_ms = device_ms(lambda: octree.render_device(cams[0], W, H, t=t_d[0], nMajor=nm_d[0], vIndex=vi_d[0], stream=stream), reps=10) if rank == 0 or world > 1 else None

    # ---- e2e: host-buffer C-ABI calls, results into pinned host memory ----
    # headline e2e = the voxRTGPU frame loop (voxRTGPU.cpp:150-208): `render` -> RGBA8 image -> D2H, 4 B per ray, the views of a
    # step in CHUNK-view calls of hako_render_primary_views (frame i+1 traced while frame i is copied out, one host wait per call);
    # the hit-record variant (t / nMajor / vIndex, 12 B per ray) and the one-synchronous-call-per-frame form are timed beside it
    import ctypes as C
    n_chunks = (VIEWS + CHUNK - 1) // CHUNK
    cam_chunks = [(hk.Camera * len(cams[c * CHUNK:(c + 1) * CHUNK]))(*cams[c * CHUNK:(c + 1) * CHUNK]) for c in range(n_chunks)]
    rgba_all = torch.empty(CHUNK * nrays_view * 4, dtype=torch.uint8).pin_memory().numpy()
    t_all = torch.empty(CHUNK * nrays_view, dtype=torch.float32).pin_memory().numpy()
    nm_all = torch.empty(CHUNK * nrays_view, dtype=torch.int32).pin_memory().numpy()
    vi_all = torch.empty(CHUNK * nrays_view, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    rgba_h = rgba_all[:nrays_view * 4]

    def step_host_rgba():
        for cc in cam_chunks:
            hk._abi.check(ctx.lib.hako_render_primary_views(octree.h, None, cc, len(cc), W, H, 0, rgba_all.ctypes.data, None, None, None))

    def step_host_records():
        for cc in cam_chunks:
            hk._abi.check(ctx.lib.hako_render_primary_views(octree.h, None, cc, len(cc), W, H, 0, None, t_all.ctypes.data, nm_all.ctypes.data, vi_all.ctypes.data))

    def step_host_rgba_per_frame():
        for cam in cams:
            hk._abi.check(ctx.lib.hako_render_primary(octree.h, None, C.byref(cam), W, H, 0, H, 0, rgba_h.ctypes.data, None, None, None))

    def time_host(fn, steps):
        fn()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        sync_all()
        return max_over_ranks((time.perf_counter() - t0) * 1e3) / steps

    e2e_steps = max(2, min(args.steps, 10))
    e2e_ms = time_host(step_host_rgba, e2e_steps)
    e2e_rec_ms = time_host(step_host_records, max(2, e2e_steps // 3))
    e2e_frame_ms = time_host(step_host_rgba_per_frame, max(2, e2e_steps // 3))
    # the frames that came back are the frames the device-resident leg produces: the last RING views of a step
    step_device()
    torch.cuda.synchronize(dev)
    last = len(cam_chunks[-1])
    host_t = t_all[(last - RING) * nrays_view:last * nrays_view].view(np.uint32).reshape(RING, nrays_view)
    host_vi = vi_all[(last - RING) * nrays_view:last * nrays_view].reshape(RING, nrays_view)
    slots = [(VIEWS - RING + j) % RING for j in range(RING)]
    e2e_parity = bool(all(np.array_equal(host_t[j], t_d[slots[j]].cpu().numpy().view(np.uint32)) and
                          np.array_equal(host_vi[j], vi_d[slots[j]].cpu().numpy().view(np.uint32)) for j in range(RING)))
    hits_last = int((t_d[slots[-1]] != 3.402823466e38).sum().item())
    # the host link's own ceiling in the same run: every rank copies CHUNK frames' worth of bytes device -> pinned host with plain
    # cudaMemcpyAsync, all ranks at once (what the e2e leg cannot exceed: the ranks of one box share the host's PCIe / memory)
    src_d = torch.empty(CHUNK * nrays_view * 4, dtype=torch.uint8, device=dev)
    dst_h = torch.from_numpy(rgba_all)

    def d2h():
        dst_h.copy_(src_d, non_blocking=True)
    d2h()
    link_ms = device_ms(d2h, reps=4)
    host_link_gbs = CHUNK * nrays_view * 4 * world / link_ms / 1e6
    del src_d
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- strong scaling (fixed total work split over the ranks): STRONG_VIEWS frames in total ----
    sviews = [k for k in range(STRONG_VIEWS) if k % world == rank]
    scams = [hk.camera_lookat(view_eye(k, STRONG_VIEWS), (0, 0, 0), width=W, height=H) for k in sviews]
    cast_views(scams)
    strong_ms = device_ms(lambda: cast_views(scams), reps=3)
    multi["strong_primary"] = {"views_total": STRONG_VIEWS, "ms": round(strong_ms, 3), "mrays_per_s": round(STRONG_VIEWS * nrays_view / strong_ms / 1e3, 1),
                               "collective": "none: frames stay on the rank that traced them"}

    # ---- path tracing leg (BASELINE config[2] shape): its own octree at 8192^3, 1920x1080, 4 bounces, procedural 2k environment;
    # samples-per-pixel sharded over ranks (rank r renders iterations r, r+G, ...), ONE ncclReduce of the accumulators ----
    pt_octree = None
    pt_build_ms = None
    if rank == 0:
        tris, col, emi, _, _, _ = scene
        from scenes import scene_params
        pt_origin, pt_dps = scene_params(tris, PT_GRID)
        pt_build_all = []
        for _ in range(2):          # the first build at this grid grows the memory pool (hundreds of ms on a run without the build sweep)
            if pt_octree is not None:
                pt_octree.cleanUp()
            t0 = time.perf_counter()
            pt_octree = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, pt_origin, float(pt_dps), PT_GRID)
            pt_build_all.append((time.perf_counter() - t0) * 1e3)
        pt_build_ms = min(pt_build_all)
    if world > 1:
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        pt_octree = nd.broadcast_octree(hk.IntersectorOctreeGPU, pt_octree, 0, stream=stream)
        b.record(stream)
        sync_all()
        bc_ms = max_over_ranks(a.elapsed_time(b))
        bytes_ = sum(nb for _, nb in pt_octree.device_arrays())
        multi["octree_broadcast_8192"] = {"device_ms": round(bc_ms, 3), "MB": round(bytes_ / 1e6, 1), "GBps": round(bytes_ / bc_ms / 1e6, 1)}
    pt_res = run_pt_leg(hk, hdist, nd, ctx, pt_octree, dev, rank, world, args, multi)
    if rank == 0 and pt_res is not None:
        pt_info = pt_octree.info()
        pt_res["octree"] = {"grid": PT_GRID, "voxels": int(pt_info.numberOfVoxels), "nodes": int(pt_info.numberOfNodes), "build_e2e_ms": pt_build_ms, "build_e2e_ms_first": pt_build_all[0],
                            "device_bytes": int(pt_info.deviceBytes)}
    # the reference's own GPU kernels on this GPU, same scenes (the bar the new kernels must beat)
    ref_gpu = run_reference_gpu(hk, ctx, octree, pt_octree, cams[0], args) if (rank == 0 and not args.no_cpu) else None
    pt_octree.cleanUp()
    # BASELINE config[4] (the massive scene, 8 GPUs): on by default when the job has 8 ranks, HAKO_BENCH_MASSIVE=1 / 0 overrides
    want_massive = os.environ.get("HAKO_BENCH_MASSIVE", "1" if world == 8 else "0") == "1"
    if want_massive:
        octree_keep = octree
        try:
            multi["massive_32768"] = run_massive_leg(hk, hdist, nd, ctx, dev, rank, world, scene)
        except Exception as exc:
            multi["massive_32768"] = {"error": repr(exc)[:200]}
        octree = octree_keep
    sweep = run_build_sweep(hk, ctx, scene) if (rank == 0 and world == 1 and not args.no_sweep) else None
    l2_bytes, l2_gbs = ctx.l2_probe(0, 20) if rank == 0 else (0, 0.0)

    if rank != 0:
        if nd is not None:
            nd.close()
        if world > 1:
            dist.destroy_process_group()
        return
    rays_step_all = nrays_view * VIEWS * world
    value = rays_step_all * args.steps / (ms_total * 1e-3) / 1e6
    e2e_value = rays_step_all / (e2e_ms * 1e-3) / 1e6

    # ---- CPU baseline + descents/ray + parity spot check (rank 0, N=1 only: bounded CPU work) ----
    cpu = None
    dbar = None
    parity = None
    if world == 1 and not args.no_cpu:
        port, ref = load_oracle()
        tris, col, emi, origin, dps, _ = scene
        o, tree, mk, cpu_build = cpu_tree(port, ref, tris, col, emi, origin, dps)
        threads = host_threads()
        sample_views = rank_views(0, 1)[::VIEWS // 64]   # 64 of the step's views spread over the orbit, row band (k mod 16) of each: 8.3 M rays
        cpu_cast(o, tree, views_total, sample_views[:2], threads, True)
        rays_c, secs_c = 0, 0.0
        reps = 0
        while secs_c < 10.0 and reps < 40:
            r, s = cpu_cast(o, tree, views_total, sample_views, threads, True)
            rays_c += r
            secs_c += s
            reps += 1
        r1, s1 = cpu_cast(o, tree, views_total, sample_views[:16], 1, True)
        ptree = tree if o is port else port.build_octree(mk, origin, dps, GRID)
        stats = np.zeros(4, np.uint64)
        rs, _ = cpu_cast(port, ptree, views_total, sample_views, threads, True, stats)
        dbar = float(stats[1]) / rs
        cpu = {"value": rays_c / secs_c / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference" if ref is not None else "port",
               "sample": "row band (k mod %d) of view k for 64 of the step's %d views (%d rays), repeated %d times (%.1f s); %d host threads over image rows" % (
                   REF_BANDS, VIEWS, rs, reps, secs_c, threads),
               "single_thread_mrays": r1 / s1 / 1e6, "cpu": cpu_model(), "build_ms": {k: (round(v, 1) if isinstance(v, float) else v) for k, v in cpu_build.items()},
               "traversal_stats_per_ray": {"loop_visits": float(stats[0]) / rs, "descents": dbar, "pushes": float(stats[2]) / rs, "max_stack": int(stats[3])}}
        # parity spot check: view 0 complete, GPU vs the CPU reference
        octree.render_device(cams[0], W, H, t=t_d[0], nMajor=nm_d[0], vIndex=vi_d[0], stream=stream)
        torch.cuda.synchronize(dev)
        cam0 = o.camera_lookat(view_eye(my_views[0], views_total), (0, 0, 0), (0, 1, 0), None, W, H)
        ct, cnm, cvi = tree.render_primary(cam0, W, H, 0, H, threads)
        parity = bool((t_d[0].cpu().numpy().view(np.uint32) == ct.view(np.uint32)).all() and (vi_d[0].cpu().numpy().view(np.uint32) == cvi).all()
                      and (nm_d[0].cpu().numpy() == cnm).all() and mk.size == info.numberOfVoxels and tree.num_nodes == info.numberOfNodes)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    if dbar is None:
        dbar = load_profile_value("descents_per_ray", 7.84)
    bytes_per_ray = 12.0 + 8.0 * dbar
    launch_ms = ms_total / max(launches, 1)
    achieved = nrays_view * bytes_per_ray / (launch_ms * 1e-3) / 1e9
    roofline = {"kernel": "renderPrimaryKernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
                "traffic": load_profile_value("render_dram_bytes_per_launch", None),
                "algorithmic_bytes_per_ray": bytes_per_ray, "descents_per_ray": dbar, "rays_per_launch": nrays_view, "avg_launch_ms": launch_ms,
                "avg_launch_ms_is": "timed region / launches with the frames of a step overlapped on %d streams; a frame launched ALONE is `isolated`" % n_streams,
                "isolated": {"launch_ms": alone_ms, "achieved": nrays_view * bytes_per_ray / (alone_ms * 1e-3) / 1e9,
                             "frac": nrays_view * bytes_per_ray / (alone_ms * 1e-3) / 1e9 / peak, "mrays_per_s": nrays_view / alone_ms / 1e3},
                # the roof that matters while the node array fits L2: measured read bandwidth over a resident set (hako_ctx_l2_probe),
                # against the sector-granular node traffic of the traversal (one 32-byte sector per descent + the 12-byte record)
                "l2": {"l2_bytes": l2_bytes, "peak_gbs": l2_gbs, "peak_source": "hako_ctx_l2_probe: ld.global.cg sweep of a %d MB resident set, best of 3" % (l2_bytes // 4 // 1000000),
                       "achieved_gbs": nrays_view * (12.0 + 32.0 * dbar) / (launch_ms * 1e-3) / 1e9,
                       "frac": (nrays_view * (12.0 + 32.0 * dbar) / (launch_ms * 1e-3) / 1e9 / l2_gbs) if l2_gbs else None,
                       "ncu_lts_throughput_pct": load_profile_value("render_l2_throughput_pct_of_peak", None)},
                "ncu": {k: load_profile_value(k, None) for k in ("render_issue_active_pct", "render_lanes_active_per_instruction", "render_l2_throughput_pct_of_peak",
                                                                  "render_l1_hit_pct", "render_l2_hit_pct", "render_warp_instructions_per_launch", "source")},
                "note": "pointer-chasing traversal with the %.0f MB octree L2-resident: the binding limit is the warp-instruction issue rate, not HBM (issue_bound below); a launch "
                        "timed ALONE is bound by its longest rays (a ray is a serial chain of dependent node fetches), which is why the frames of a step are overlapped on several streams"
                        % (info.numberOfNodes * 64 / 1e6)}
    winst = load_profile_value("render_warp_instructions_per_launch", None)
    clocks = sampler.summary()
    if winst:
        # issue roofline: every SM has 4 warp schedulers issuing at most 1 warp instruction per clock
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
        floor_ms = float(winst) / (sm_count * 4 * mhz * 1e6) * 1e3
        roofline["issue_bound"] = {"warp_instructions_per_launch": winst, "schedulers": sm_count * 4, "sm_mhz": mhz, "floor_ms_per_launch": floor_ms,
                                   "frac": floor_ms / launch_ms, "frac_isolated": floor_ms / alone_ms,
                                   "source": "instruction count of view 0 from the ncu capture named in roofline.ncu.source (this round)"}
    cfg = base_config(world)
    e2e_bytes = 4 * nrays_view * VIEWS
    line = {"metric": "Mrays/s (octree traversal, primary rays)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 60 * VIEWS, "d2h_bytes_per_step": e2e_bytes,
                    "ms_per_step": e2e_ms, "steps_timed": e2e_steps,
                    "host_link_gbs": host_link_gbs, "d2h_gbs": e2e_bytes * world / e2e_ms / 1e6, "frac_of_host_link": e2e_bytes * world / e2e_ms / 1e6 / host_link_gbs,
                    "host_link_is": "all %d rank(s) copying %d MB device -> pinned host with plain cudaMemcpyAsync at the same time: the ceiling of this leg" % (world, CHUNK * nrays_view * 4 // 1000000),
                    "call": "hako_render_primary_views (pinned host frames): the %d views of a step in %d-view calls, RGBA8 frames out = the voxRTGPU frame loop (voxRTGPU.cpp:150-208), frame i+1 traced while frame i is copied" % (VIEWS, CHUNK),
                    "per_frame_call": {"value": rays_step_all / (e2e_frame_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_frame_ms, "call": "hako_render_primary, one synchronous call per frame"},
                    "frames_equal_device_leg": e2e_parity,
                    "hit_records": {"value": rays_step_all / (e2e_rec_ms * 1e-3) / 1e6, "unit": "Mrays/s", "d2h_bytes_per_step": 12 * nrays_view * VIEWS,
                                    "ms_per_step": e2e_rec_ms, "call": "same call returning t / nMajor / vIndex (12 B per ray) instead of the images"}},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "build": build, "build_sweep": sweep,
            "hits_last_view": hits_last, "parity_vs_cpu_reference_view0": parity,
            "published_reference": "1.04 Grays/s @ 8192^3 dragon on RTX 3090 Ti / RX 7900 XTX (seminar slide 66), not this metric/hardware",
            # short blocks LAST, so that they survive a log tail: the reference's own GPU kernels on this box, path tracing, multi-GPU
            "reference_gpu": ref_gpu, "path_tracing": pt_res, "multi_gpu": dict(multi, collectives=getattr(nd, "kind", "hako_dist_* (NCCL from C++, include/hako.h)") if nd is not None else "none (1 GPU)")}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if nd is not None:
        nd.close()
    if world > 1:
        dist.destroy_process_group()


PT_STEPS = 4            # PathTracer::step calls per rank inside the timed region (16 spp each)
PT_DEPTH = 4            # BASELINE config[2]: 4 bounces (the reference's literal is 8, voxKernel.cu:691)
PT_GRID = 8192          # BASELINE config[2]: the dragon at 8192^3
SWEEP_GRIDS = (1024, 2048, 4096, 8192, 16384)   # BASELINE config[3]: voxelize + sort + octree build, 2^10 .. 2^14


def run_pt_leg(hk, hdist, nd, ctx, octree, dev, rank, world, args, multi):
    import torch
    import torch.distributed as dist
    from scenes import procedural_hdri
    pt = hk.PathTracer(ctx).setup()
    own = pt.m_intersectorOctreeGPU
    pt.m_intersectorOctreeGPU = octree          # share the resident octree (no rebuild)
    pt.maxDepth = PT_DEPTH
    pt.resizeFrameBufferIfNeeded(None, W, H)
    env = procedural_hdri(2048, 1024)
    pt.loadHDRI(None, env, env)
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
    stream = torch.cuda.current_stream()

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def mx(x, op=None):
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op or dist.ReduceOp.MAX)
        return float(t.item())

    def frame(iters):
        pt.clearFrameBuffer()
        for it in iters:
            pt.step(None, cam, iteration=it)
        if world > 1:
            nd.reduce_accum(pt, 0)              # hako_dist_reduce_accum: the ONE collective of the path

    def timed(iters, reps):
        frame(iters)
        sync_all()
        pt.ray_count(reset=True)
        pt.trace_steps(reset=True)
        l0 = ctx.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            frame(iters)
        e1.record(stream)
        sync_all()
        timed.steps = mx(pt.trace_steps() / reps, dist.ReduceOp.SUM if world > 1 else None)   # node visits of the secondary rays, all ranks
        return mx(e0.elapsed_time(e1) / reps), mx(pt.ray_count() / reps, dist.ReduceOp.SUM if world > 1 else None), (ctx.launch_count - l0) // reps

    # weak scaling: PT_STEPS steps per rank, whatever N
    iters = hdist.spp_shard(PT_STEPS * world, rank, world)
    reps = 3
    ms, rays, launches = timed(iters, reps)
    trace_steps = timed.steps
    spp = 16 * PT_STEPS * world
    a = pt.accumulators() if rank == 0 else None
    # e2e: the PathTracer calls a voxPTGPU frame makes (clear, steps, reduce, toImageAsync = resolve + D2H of the RGBA8 frame)
    t0 = time.perf_counter()
    for _ in range(reps):
        frame(iters)
        if rank == 0:
            pt.toImageAsync()
    sync_all()
    e2e_ms = mx((time.perf_counter() - t0) * 1e3 / reps)
    # strong scaling: ONE 512-spp frame (STRONG_SPP_STEPS steps in total) split over the ranks, reduce included
    s_iters = hdist.spp_shard(STRONG_SPP_STEPS, rank, world)
    s_ms, _, _ = timed(s_iters, 2)
    s_ok = bool((pt.accumulators()[..., 3] == 16 * STRONG_SPP_STEPS).all()) if rank == 0 else None
    multi["strong_path_tracing"] = {"spp": 16 * STRONG_SPP_STEPS, "ms_per_frame": round(s_ms, 2), "mpaths_per_s": round(W * H * 16 * STRONG_SPP_STEPS / s_ms / 1e3, 1),
                                    "spp_count_ok": s_ok, "collective": "ncclReduce(sum) of the 33 MB float4 accumulators (hako_dist_reduce_accum)" if world > 1 else "none (1 GPU)"}
    res = None
    if rank == 0:
        ok = bool((a[..., 3] == spp).all())
        paths = W * H * spp
        res = {"workload": "config[2]: stand-in dragon %d^3, %dx%d, maxDepth %d, procedural 2048x1024 environment (monks_forest_2k.hdr is missing), %d spp per frame = %d steps x 16 spp per GPU"
                           % (PT_GRID, W, H, PT_DEPTH, spp, PT_STEPS),
               "ms_per_frame": ms, "ms_per_spp_per_gpu": ms / (16 * PT_STEPS), "mpaths_per_s": paths / (ms * 1e-3) / 1e6,
               "mrays_per_s": rays / (ms * 1e-3) / 1e6, "rays_per_path": rays / paths,
               # how long the secondary rays are: node visits per ray of the trace waves (hako_pt_trace_steps) -- the primary rays of the headline
               # need ~12; visits/s of the whole frame time is a LOWER bound of the trace kernel's rate (primary wave and shading are inside it)
               "secondary_rays": {"per_path": (rays - paths) / paths, "node_visits_per_ray": trace_steps / max(rays - paths, 1.0),
                                  "gvisits_per_s_over_frame_time": trace_steps / (ms * 1e-3) / 1e9},
               "e2e_ms_per_frame": e2e_ms, "e2e_mpaths_per_s": paths / (e2e_ms * 1e-3) / 1e6, "d2h_bytes_per_frame": W * H * 4,
               "launches_per_frame_per_gpu": int(launches), "spp_count_ok": ok, "mean_radiance": float(a[..., :3].mean() / spp),
               "collective": "ncclReduce(sum) of %d MB float4 accumulators" % (W * H * 16 // 1000000) if world > 1 else "none (1 GPU)",
               "published_reference": "12.5 ms/spp at 1920x1080, 8 bounces, 4096^3 scene, RX 7900 XTX (seminar slide 67); different scene and hardware"}
    pt.m_intersectorOctreeGPU = own
    pt.cleanUp()
    return res


def run_massive_leg(hk, hdist, nd, ctx, dev, rank, world, scene):
    """BASELINE config[4], scaled to the 32-bit voxel counters of the reference format: the stand-in dragon at 32768^3 -- billions of
    surface voxels, a wide-index octree of ~47 GB -- built on rank 0, replicated to every GPU over NVLink (hako_dist_broadcast_octree),
    then a 3840x2160 primary frame per GPU and 4K path tracing with the spp sharded over the ranks and one ncclReduce."""
    import torch
    import torch.distributed as dist
    from scenes import procedural_hdri, scene_params
    MW, MH, MG = 3840, 2160, 32768
    stream = torch.cuda.current_stream()
    out = {"grid": MG, "image": [MW, MH]}

    def mx(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    g = None
    if rank == 0:
        tris, col, emi, _, _, _ = scene
        origin, dps = scene_params(tris, MG)
        t0 = time.perf_counter()
        g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), MG)
        out["build_host_call_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
        st = ctx.build_stats()
        info = g.info()
        out["voxels"], out["nodes"], out["wide_index"], out["device_GB"] = int(info.numberOfVoxels), int(info.numberOfNodes), int(info.wideIndex), round(info.deviceBytes / 1e9, 2)
        out["build_device_ms"] = {k: round(st[k], 1) for k in ("ms_voxelize", "ms_sort", "ms_unique", "ms_octree", "ms_total")}
    if world > 1:
        # replicate twice: the first call makes the receivers' memory pools map 47 GB (host time in front of the collective),
        # the second is the steady-state replication a frame loop would see
        src = g
        times = []
        for rep in range(2):
            if rep and rank != 0:
                g.cleanUp()
            torch.cuda.synchronize(dev)
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            g = nd.broadcast_octree(hk.IntersectorOctreeGPU, src, 0, stream=stream)
            b.record(stream)
            torch.cuda.synchronize(dev)
            dist.barrier()
            times.append(mx(a.elapsed_time(b)))
        nbytes = sum(nb for _, nb in g.device_arrays())
        out["broadcast"] = {"first_ms": round(times[0], 1), "device_ms": round(times[1], 1), "GB": round(nbytes / 1e9, 2), "GBps": round(nbytes / times[1] / 1e6, 1)}
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=MW, height=MH, focus=6.0, lensR=0.0)
    t_d = torch.empty(MW * MH, dtype=torch.float32, device=dev)
    nm_d = torch.empty(MW * MH, dtype=torch.int32, device=dev)
    vi_d = torch.empty(MW * MH, dtype=torch.int32, device=dev)
    for _ in range(3):
        g.render_device(cam, MW, MH, t=t_d, nMajor=nm_d, vIndex=vi_d, stream=stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        g.render_device(cam, MW, MH, t=t_d, nMajor=nm_d, vIndex=vi_d, stream=stream)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = mx(e0.elapsed_time(e1) / 10)
    out["primary_4k"] = {"ms_per_frame_per_gpu": round(ms, 3), "mrays_per_s_all_gpus": round(MW * MH * world / ms / 1e3, 1), "hits": int((t_d != 3.402823466e38).sum().item())}
    pt = hk.PathTracer(ctx).setup()
    own = pt.m_intersectorOctreeGPU
    pt.m_intersectorOctreeGPU = g
    pt.maxDepth = 4
    pt.resizeFrameBufferIfNeeded(None, MW, MH)
    env = procedural_hdri(2048, 1024)
    pt.loadHDRI(None, env, env)
    iters = hdist.spp_shard(4 * world, rank, world)

    def frame():
        pt.clearFrameBuffer()
        for it in iters:
            pt.step(None, cam, iteration=it)
        if world > 1:
            nd.reduce_accum(pt, 0)
    frame()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0.record(stream)
    frame()
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = mx(e0.elapsed_time(e1))
    spp = 64 * world
    out["pt_4k"] = {"spp": spp, "ms_per_frame": round(ms, 1), "mpaths_per_s": round(MW * MH * spp / ms / 1e3, 1)}
    if rank == 0:
        acc = pt.accumulators()
        out["pt_4k"]["spp_count_ok"] = bool((acc[..., 3] == spp).all())
        out["pt_4k"]["mean_radiance"] = round(float(acc[..., :3].mean() / spp), 4)
    pt.m_intersectorOctreeGPU = own
    pt.cleanUp()
    g.cleanUp()
    del t_d, nm_d, vi_d
    return out


def run_reference_gpu(hk, ctx, octree, pt_octree, cam0, args):
    """The reference's OWN kernels (voxKernel.cu compiled ahead of time for sm_100a by oracle/Makefile, launched by
    oracle/librefgpu.so) on this GPU and these scenes, beside ours: `render` on the bench octree (view 0) and `renderPT` at depth 4
    on the 8192^3 octree.  Test infrastructure used as a reported baseline, after the timed regions -- like cpu_baseline."""
    try:
        from oracle import refgpu
        from scenes import procedural_hdri
        if not refgpu.available():
            return {"unavailable": "oracle/_ref/voxKernel.sm_100a.cubin not built (needs /root/reference at build time)"}
        out = {"what": "reference voxKernel.cu kernels, nvcc -arch=sm_100a (fmad on, as the reference's nvrtc build), same octree bytes (68-byte nodes) and cameras"}
        info = octree.info()
        if info.wideIndex:
            return {"unavailable": "the bench octree exceeds the reference's 24-bit node index"}
        nodes, _, attrs = octree.export(True, False, True)
        r = refgpu.RefGpu("default")
        r.set_octree(nodes, attrs.view(np.uint8).reshape(-1, 8), np.array(info.lower, np.float32), np.array(info.upper, np.float32), float(info.dps), bool(info.hasEmission))
        _, best, mean = r.render(cam0, W, H, False, reps=10)
        ours_rgba, _, _, _ = octree.render(cam0, W, H, ids=False)
        img, _, _ = r.render(cam0, W, H, False, reps=1)
        out["render"] = {"ms_best": best, "ms_mean": mean, "mrays_per_s": W * H / best / 1e3, "pixels_differing_from_ours": float((img != ours_rgba).any(-1).mean())}
        r.close()
        del nodes, attrs
        pinfo = pt_octree.info()
        if not pinfo.wideIndex and refgpu.available("d4"):
            nodes, _, attrs = pt_octree.export(True, False, True)
            r = refgpu.RefGpu("d4")
            r.set_octree(nodes, attrs.view(np.uint8).reshape(-1, 8), np.array(pinfo.lower, np.float32), np.array(pinfo.upper, np.float32), float(pinfo.dps), bool(pinfo.hasEmission))
            env = procedural_hdri(2048, 1024)
            r.load_hdri(env, env)
            pt = hk.PathTracer(ctx).setup()
            r.set_pmj(pt.pmj_table())
            pt.cleanUp()
            r.pt_resize(W, H)
            cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
            ms = [r.pt_step(cam, it) for it in range(3)]
            acc = r.pt_accum()
            out["renderPT"] = {"depth": PT_DEPTH, "grid": PT_GRID, "ms_per_16spp_step": min(ms), "ms_per_spp": min(ms) / 16.0, "mean_radiance": float(acc[..., :3].mean() / 48.0)}
            r.close()
        return out
    except Exception as exc:
        return {"error": repr(exc)[:300]}


def run_build_sweep(hk, ctx, scene):
    """BASELINE config[3]: voxelization + Morton radix sort + octree build at 2^10 .. 2^14 on the bench mesh (host-buffer call,
    best of 3 after a warm-up; per-stage device times from hako_ctx_build_stats).  A grid whose octree exceeds the reference
    format's 24-bit node index is reported as such (HAKO_E_LIMIT), not skipped silently."""
    from scenes import scene_params
    tris, col, emi, _, _, _ = scene
    rows = []
    for grid in SWEEP_GRIDS:
        origin, dps = scene_params(tris, grid)
        row = {"grid": grid}
        try:
            best = None
            for i in range(4):
                t0 = time.perf_counter()
                g = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, origin, float(dps), grid)
                wall = (time.perf_counter() - t0) * 1e3
                st = ctx.build_stats()
                g.cleanUp()
                if i > 0 and (best is None or wall < best[0]):
                    best = (wall, dict(st))
            wall, st = best
            row.update({"e2e_host_call_ms": round(wall, 3), "dumped_voxels": int(st["dumpedVoxels"]), "unique_voxels": int(st["uniqueVoxels"]),
                        "octree_nodes": int(st["numberOfNodes"]), "sort_passes": int(st["sortPasses"]),
                        "device_ms": {k: round(st[k], 3) for k in ("ms_h2d", "ms_voxelize", "ms_sort", "ms_unique", "ms_octree", "ms_total")}})
        except hk.HakoError as e:
            row["error"] = str(e)[:200]
        rows.append(row)
    return rows


_REAL_STDOUT = sys.stdout


def load_profile_value(key, default):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "latest.json"))).get(key, default)
    except Exception:
        return default


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the build sweep (profiling runs)")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: libraries that write to fd 1 (NCCL prints its version banner there) go to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
