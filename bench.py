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
    views_total = VIEWS * world
    nrays_view = W * H

    # ---- build (rank 0), host-buffer call = the e2e build; then replicate ----
    build = None
    octree = None
    scene = None
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
        warm = torch.zeros(1, device=dev)
        dist.broadcast(warm, 0)                 # communicator set-up is not part of the replication time
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        octree = hdist.broadcast_octree(ctx, hk.IntersectorOctreeGPU, octree, 0)
        dist.barrier()
        if rank == 0:
            build["nccl_broadcast_ms"] = (time.perf_counter() - t0) * 1e3
        # the alternative to build-on-rank-0 + broadcast: every rank voxelizes / sorts / merges 1/N of the scene, one all-to-all by
        # Morton range, unique voxels gathered, DAG built on every rank (massivevoxelraytracing_b200/dist.py: build_sharded).
        # Reported beside the replicated build; the octree used below is the replicated one (they are the same bytes).
        try:
            if os.environ.get("HAKO_BENCH_SHARDED", "1") == "0":
                raise RuntimeError("skipped (HAKO_BENCH_SHARDED=0)")
            if scene is None:
                scene = make_scene()            # deterministic generator: every rank derives the same triangles
            tris, col, emi, origin, dps, _ = scene
            stages = hdist.LibStages(ctx, hk.IntersectorOctreeGPU)
            best = None
            for _ in range(3):
                times = {}
                dist.barrier()
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                sh = hdist.build_sharded(stages, tris, col, emi, origin, dps, GRID, times=times)
                torch.cuda.synchronize(dev)
                dist.barrier()
                wall_ms = (time.perf_counter() - t0) * 1e3
                same = int(sh.info().numberOfNodes == octree.info().numberOfNodes and sh.info().numberOfVoxels == octree.info().numberOfVoxels)
                sh.cleanUp()
                if best is None or wall_ms < best[0]:
                    best = (wall_ms, {k: round(v, 2) for k, v in times.items() if k.endswith("_ms")}, same)
            ok = torch.tensor([best[2]], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if rank == 0:
                build["sharded_build"] = {"e2e_ms": best[0], "phases_rank0_ms": best[1], "same_counts_as_replicated_on_every_rank": bool(ok.item()),
                                          "what": "host triangles in -> identical octree resident on every rank; compare with e2e_host_call_ms + nccl_broadcast_ms"}
            torch.cuda.empty_cache()
        except Exception as exc:                # the bench line must survive a failure of this optional leg
            if rank == 0:
                build["sharded_build"] = {"error": repr(exc)[:300]}
    info = octree.info()

    my_views = rank_views(rank, world)
    cams = [hk.camera_lookat(view_eye(k, views_total), (0, 0, 0), width=W, height=H) for k in my_views]
    # one output set per view: 16 x 24.9 MB = 398 MB per step, larger than the 126 MB L2
    t_d = torch.empty((VIEWS, nrays_view), dtype=torch.float32, device=dev)
    nm_d = torch.empty((VIEWS, nrays_view), dtype=torch.int32, device=dev)
    vi_d = torch.empty((VIEWS, nrays_view), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream()
    # frames alternate between two streams (what hako_render_primary_views does inside): a frame is one wave of CTAs whose
    # tail -- the few long rays -- would otherwise idle most SMs before the next frame starts
    n_streams = max(1, int(os.environ.get("HAKO_BENCH_STREAMS", "4")))
    sides = [torch.cuda.Stream(device=dev) for _ in range(n_streams - 1)]
    fork, joins = torch.cuda.Event(), [torch.cuda.Event() for _ in sides]

    def step_device():
        fork.record(stream)
        for sd in sides:
            sd.wait_event(fork)
        for i, cam in enumerate(cams):
            k = i % n_streams
            octree.render_device(cam, W, H, t=t_d[i], nMajor=nm_d[i], vIndex=vi_d[i], stream=sides[k - 1] if k else stream)
        for sd, j in zip(sides, joins):
            j.record(sd)
            stream.wait_event(j)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

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
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())

    # ---- e2e: host-buffer C-ABI call, results into pinned host memory ----
    # headline e2e = the voxRTGPU frame (voxRTGPU.cpp:191-208): `render` -> RGBA8 image -> D2H, 4 B per ray;
    # the hit-record variant (t / nMajor / vIndex, 12 B per ray: what the parity checks read) is timed beside it.
    t_h = torch.empty(nrays_view, dtype=torch.float32).pin_memory().numpy()
    nm_h = torch.empty(nrays_view, dtype=torch.int32).pin_memory().numpy()
    vi_h = torch.empty(nrays_view, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    rgba_h = torch.empty(nrays_view * 4, dtype=torch.uint8).pin_memory().numpy()
    import ctypes as C

    # the frames of a step land in VIEWS distinct pinned host frames (hako_render_primary_views: frame i+1 is traced while
    # frame i is copied out); the one-call-per-frame form (hako_render_primary, trace and copy of a frame in sequence) is
    # timed beside it
    cam_arr = (hk.Camera * VIEWS)(*cams)
    rgba_all = torch.empty(VIEWS * nrays_view * 4, dtype=torch.uint8).pin_memory().numpy()
    t_all = torch.empty(VIEWS * nrays_view, dtype=torch.float32).pin_memory().numpy()
    nm_all = torch.empty(VIEWS * nrays_view, dtype=torch.int32).pin_memory().numpy()
    vi_all = torch.empty(VIEWS * nrays_view, dtype=torch.int32).pin_memory().numpy().view(np.uint32)

    def step_host_rgba():
        hk._abi.check(ctx.lib.hako_render_primary_views(octree.h, None, cam_arr, VIEWS, W, H, 0, rgba_all.ctypes.data, None, None, None))

    def step_host_records():
        hk._abi.check(ctx.lib.hako_render_primary_views(octree.h, None, cam_arr, VIEWS, W, H, 0, None, t_all.ctypes.data, nm_all.ctypes.data, vi_all.ctypes.data))

    def step_host_rgba_per_frame():
        for cam in cams:
            hk._abi.check(ctx.lib.hako_render_primary(octree.h, None, C.byref(cam), W, H, 0, H, 0, rgba_h.ctypes.data, None, None, None))

    def time_host(fn):
        fn()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        sync_all()
        ms_ = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms_, op=dist.ReduceOp.MAX)
        return float(ms_.item())

    e2e_total = time_host(step_host_rgba)
    e2e_rec_total = time_host(step_host_records)
    e2e_frame_total = time_host(step_host_rgba_per_frame)
    # the frames that came back are the frames the device-resident leg produced
    e2e_parity = bool(np.array_equal(t_all.view(np.uint32), t_d.cpu().numpy().reshape(-1).view(np.uint32))
                      and np.array_equal(vi_all, vi_d.cpu().numpy().reshape(-1).view(np.uint32)))
    sampler.stop_flag = True
    sampler.join(timeout=2)
    hits_view0 = int((t_d[0] != 3.402823466e38).sum().item())

    # ---- path tracing leg (BASELINE config[2] shape): same octree, 1920x1080, 4 bounces, procedural 2k environment;
    # samples-per-pixel sharded over ranks (rank r renders iterations r, r+G, ...), ONE ncclReduce of the accumulators ----
    pt_octree = None
    pt_build_ms = None
    if rank == 0:
        tris, col, emi, _, _, _ = scene
        from scenes import scene_params
        pt_origin, pt_dps = scene_params(tris, PT_GRID)
        t0 = time.perf_counter()
        pt_octree = hk.IntersectorOctreeGPU(ctx).build(tris, col, emi, None, pt_origin, float(pt_dps), PT_GRID)
        pt_build_ms = (time.perf_counter() - t0) * 1e3
    pt_bcast_ms = None
    if world > 1:
        dist.barrier()
        t0 = time.perf_counter()
        pt_octree = hdist.broadcast_octree(ctx, hk.IntersectorOctreeGPU, pt_octree, 0)
        dist.barrier()
        pt_bcast_ms = (time.perf_counter() - t0) * 1e3
    pt_res = run_pt_leg(hk, hdist, ctx, pt_octree, dev, rank, world, args)
    if rank == 0 and pt_res is not None:
        pt_info = pt_octree.info()
        pt_res["octree"] = {"grid": PT_GRID, "voxels": int(pt_info.numberOfVoxels), "nodes": int(pt_info.numberOfNodes), "build_e2e_ms": pt_build_ms,
                            "device_bytes": int(pt_info.deviceBytes), "nccl_broadcast_ms": pt_bcast_ms}
    pt_octree.cleanUp()
    sweep = run_build_sweep(hk, ctx, scene) if (rank == 0 and not args.no_sweep) else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    rays_step_all = nrays_view * VIEWS * world
    value = rays_step_all * args.steps / (ms_total * 1e-3) / 1e6
    e2e_value = rays_step_all * args.steps / (e2e_total * 1e-3) / 1e6

    # ---- CPU baseline + descents/ray + parity spot check (rank 0, N=1 only: bounded CPU work) ----
    cpu = None
    dbar = None
    parity = None
    if world == 1 and not args.no_cpu:
        port, ref = load_oracle()
        tris, col, emi, origin, dps, _ = scene
        o, tree, mk, cpu_build = cpu_tree(port, ref, tris, col, emi, origin, dps)
        threads = host_threads()
        cpu_cast(o, tree, views_total, range(2), threads, True)
        rays_c, secs_c = 0, 0.0
        reps = 0
        while secs_c < 3.0 and reps < 8:
            r, s = cpu_cast(o, tree, views_total, rank_views(0, views_total // VIEWS), threads, True)
            rays_c += r
            secs_c += s
            reps += 1
        r1, s1 = cpu_cast(o, tree, views_total, rank_views(0, views_total // VIEWS), 1, True)
        ptree = tree if o is port else port.build_octree(mk, origin, dps, GRID)
        stats = np.zeros(4, np.uint64)
        rs, _ = cpu_cast(port, ptree, views_total, rank_views(0, views_total // VIEWS), threads, True, stats)
        dbar = float(stats[1]) / rs
        cpu = {"value": rays_c / secs_c / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference" if ref is not None else "port",
               "sample": "row band k of view k for the %d views of a step (%d rays), repeated %d times; %d host threads over image rows" % (VIEWS, rs, reps, threads),
               "single_thread_mrays": r1 / s1 / 1e6, "cpu": cpu_model(), "build_ms": {k: (round(v, 1) if isinstance(v, float) else v) for k, v in cpu_build.items()},
               "traversal_stats_per_ray": {"loop_visits": float(stats[0]) / rs, "descents": dbar, "pushes": float(stats[2]) / rs, "max_stack": int(stats[3])}}
        # parity spot check: view 0 complete, GPU (device buffers of the last step) vs the CPU reference
        cam0 = o.camera_lookat(view_eye(0, views_total), (0, 0, 0), (0, 1, 0), None, W, H)
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
                "ncu": {k: load_profile_value(k, None) for k in ("render_issue_active_pct", "render_lanes_active_per_instruction", "render_l2_throughput_pct_of_peak",
                                                                  "render_l1_hit_pct", "render_l2_hit_pct", "render_warp_instructions_per_launch", "source")},
                "note": "pointer-chasing traversal with the %.0f MB octree L2-resident: the binding limit is the warp-instruction issue rate, not HBM (issue_bound below); a launch "
                        "timed ALONE is bound by its longest rays (a ray is a serial chain of dependent node fetches; ~29 %% of such a launch the average SM idles), which is why "
                        "the frames of a step are overlapped on several streams" % (info.numberOfNodes * 64 / 1e6)}
    winst = load_profile_value("render_warp_instructions_per_launch", None)
    if winst:
        # issue roofline: every SM has 4 warp schedulers issuing at most 1 warp instruction per clock
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        clocks = sampler.summary()
        mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
        floor_ms = float(winst) / (sm_count * 4 * mhz * 1e6) * 1e3
        roofline["issue_bound"] = {"warp_instructions_per_launch": winst, "schedulers": sm_count * 4, "sm_mhz": mhz, "floor_ms_per_launch": floor_ms,
                                   "frac": floor_ms / launch_ms, "source": "instruction count from the ncu capture named in roofline.ncu.source"}
    cfg = base_config(world)
    line = {"metric": "Mrays/s (octree traversal, primary rays)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "clocks": sampler.summary(),
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 60 * VIEWS, "d2h_bytes_per_step": 4 * nrays_view * VIEWS,
                    "ms_per_step": e2e_total / args.steps,
                    "call": "hako_render_primary_views (pinned host frames): the 16 views of a step in one call, RGBA8 frames out = the voxRTGPU frame loop (render kernel + D2H of the image, voxRTGPU.cpp:150-208), frame i+1 traced while frame i is copied",
                    "per_frame_call": {"value": rays_step_all * args.steps / (e2e_frame_total * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_frame_total / args.steps,
                                       "call": "hako_render_primary, one synchronous call per frame"},
                    "frames_equal_device_leg": e2e_parity,
                    "hit_records": {"value": rays_step_all * args.steps / (e2e_rec_total * 1e-3) / 1e6, "unit": "Mrays/s", "d2h_bytes_per_step": 12 * nrays_view * VIEWS,
                                    "ms_per_step": e2e_rec_total / args.steps, "call": "same call returning t / nMajor / vIndex (12 B per ray) instead of the images"}},
            "gpu_launches": int(launches), "roofline": roofline, "path_tracing": pt_res, "cpu_baseline": cpu, "build": build, "build_sweep": sweep,
            "hits_view0": hits_view0,
            "parity_vs_cpu_reference_view0": parity, "published_reference": "1.04 Grays/s @ 8192^3 dragon on RTX 3090 Ti / RX 7900 XTX (seminar slide 66), not this metric/hardware"}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


PT_STEPS = 4            # PathTracer::step calls per rank inside the timed region (16 spp each)
PT_DEPTH = 4            # BASELINE config[2]: 4 bounces (the reference's literal is 8, voxKernel.cu:691)
PT_GRID = 8192          # BASELINE config[2]: the dragon at 8192^3
SWEEP_GRIDS = (1024, 2048, 4096, 8192, 16384)   # BASELINE config[3]: voxelize + sort + octree build, 2^10 .. 2^14


def run_pt_leg(hk, hdist, ctx, octree, dev, rank, world, args):
    import torch
    import torch.distributed as dist
    from scenes import procedural_hdri
    pt = hk.PathTracer(ctx).setup()
    own = pt.m_intersectorOctreeGPU
    pt.m_intersectorOctreeGPU = octree          # share the resident octree (no rebuild)
    pt.maxDepth = PT_DEPTH
    pt.resizeFrameBufferIfNeeded(None, W, H)
    pt.loadHDRI(None, procedural_hdri(2048, 1024))
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.0)
    ptr, nbytes = pt.accum_device()
    acc = hdist.as_tensor(ptr, nbytes, dev).view(torch.float32)
    stream = torch.cuda.current_stream()
    iters = hdist.spp_shard(PT_STEPS * world, rank, world)

    def frame():
        pt.clearFrameBuffer()
        for it in iters:
            pt.step(None, cam, iteration=it)
        if world > 1:
            hdist.reduce_accum(acc, 0)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    frame()
    sync_all()
    pt.ray_count(reset=True)
    l0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record(stream)
    for _ in range(reps):
        frame()
    e1.record(stream)
    sync_all()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    rays = torch.tensor([pt.ray_count() / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    launches = (ctx.launch_count - l0) // reps
    # e2e: the PathTracer calls a voxPTGPU frame makes (clear, steps, reduce, toImageAsync = resolve + D2H of the RGBA8 frame)
    t0 = time.perf_counter()
    for _ in range(reps):
        frame()
        if rank == 0:
            img = pt.toImageAsync()
    sync_all()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    res = None
    if rank == 0:
        spp = 16 * PT_STEPS * world
        a = pt.accumulators()
        ok = bool((a[..., 3] == spp).all())
        paths = W * H * spp
        res = {"workload": "config[2]: stand-in dragon %d^3, %dx%d, maxDepth %d, procedural 2048x1024 environment (monks_forest_2k.hdr is missing), %d spp per frame = %d steps x 16 spp per GPU"
                           % (PT_GRID, W, H, PT_DEPTH, spp, PT_STEPS),
               "ms_per_frame": float(ms.item()), "ms_per_spp_per_gpu": float(ms.item()) / (16 * PT_STEPS), "mpaths_per_s": paths / (float(ms.item()) * 1e-3) / 1e6,
               "mrays_per_s": float(rays.item()) / (float(ms.item()) * 1e-3) / 1e6, "rays_per_path": float(rays.item()) / paths,
               "e2e_ms_per_frame": float(e2e_ms.item()), "e2e_mpaths_per_s": paths / (float(e2e_ms.item()) * 1e-3) / 1e6, "d2h_bytes_per_frame": W * H * 4,
               "launches_per_frame_per_gpu": int(launches), "spp_count_ok": ok, "mean_radiance": float(a[..., :3].mean() / spp),
               "collective": "ncclReduce(sum) of %d MB float4 accumulators" % (W * H * 16 // 1000000) if world > 1 else "none (1 GPU)",
               "published_reference": "12.5 ms/spp at 1920x1080, 8 bounces, 4096^3 scene, RX 7900 XTX (seminar slide 67); different scene and hardware"}
    pt.m_intersectorOctreeGPU = own
    pt.cleanUp()
    return res


def run_build_sweep(hk, ctx, scene):
    """BASELINE config[3]: voxelization + Morton radix sort + octree build at 2^10 .. 2^14 on the bench mesh (host-buffer call,
    best of 2 after a warm-up; per-stage device times from hako_ctx_build_stats).  A grid whose octree exceeds the reference
    format's 24-bit node index is reported as such (HAKO_E_LIMIT), not skipped silently."""
    from scenes import scene_params
    tris, col, emi, _, _, _ = scene
    rows = []
    for grid in SWEEP_GRIDS:
        origin, dps = scene_params(tris, grid)
        row = {"grid": grid}
        try:
            best = None
            for i in range(3):
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
