#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE'S OWN CPU CODE
(oracle/_ref/libhako_ref.so, built from /root/reference by oracle/Makefile).

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py

Outputs
  bunny_tris.npz       the test input: scenes/bunny.obj flattened to float32 [T,3,3] triangle soup in face
                       order (what trianglesFlattened, voxUtil.hpp:8-65, hands to the voxelizer).  This is
                       a derived input vector, not reference source.
  golden.json          counts, FNV-1a-64 checksums and sums produced by the reference for every case in
                       SURVEY.md section 8(c) plus a few small-grid cases; per-level node profile.
  small_cases.npz      full arrays (keys, merged keys/colours, node arrays, hit images) for tiny grids so
                       that oracle and CUDA tests can diff element-wise without the reference present.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402

REF_OBJ = "/root/reference/scenes/bunny.obj"


def load_obj_triangles(path):
    """Minimal OBJ reader (v / f with triangles only), vertex order = face order like tiny_obj_loader."""
    vs = []
    tris = []
    with open(path) as f:
        for line in f:
            if line.startswith("v "):
                vs.append([float(x) for x in line.split()[1:4]])
            elif line.startswith("f "):
                idx = [int(tok.split("/")[0]) for tok in line.split()[1:]]
                assert len(idx) == 3
                tris.append([i - 1 if i > 0 else len(vs) + i for i in idx])
    v = np.asarray(vs, np.float64).astype(np.float32)
    return v[np.asarray(tris, np.int64)]  # [T,3,3]


def fnv(o, arr):
    return "%016x" % o.fnv1a64(arr)


def vertex_colors(tris):
    """Deterministic per-vertex colours/emissions in [0,1] so the attribute path is exercised
    (the bunny has no Cd / Emission attributes; the reference then uses colour 1 / emission 0)."""
    p = tris.reshape(-1, 3)
    lo, hi = p.min(0), p.max(0)
    col = ((p - lo) / (hi - lo)).astype(np.float32)
    emi = np.zeros_like(col)
    emi[p[:, 1] > (lo[1] + 0.8 * (hi[1] - lo[1]))] = np.float32(0.75)
    return col.reshape(-1, 3, 3), emi.reshape(-1, 3, 3)


def scene_params(tris, grid):
    p = tris.reshape(-1, 3)
    lo = p.min(0)
    hi = p.max(0)
    size = hi - lo
    dps = np.float32(np.float32(max(size[0], size[1], size[2])) / np.float32(grid))
    return lo.astype(np.float32), dps


def main():
    orc.build(quiet=False)
    ref = orc.Oracle("reference")
    assert ref.lib.orc_impl_name() == b"reference"

    tris = load_obj_triangles(REF_OBJ)
    np.savez_compressed(os.path.join(HERE, "bunny_tris.npz"), tris=tris)
    col, emi = vertex_colors(tris)

    golden = {"source": "Ushio/MassiveVoxelRayTracing@1ebdd4f CPU headers via oracle/_ref/libhako_ref.so",
              "mesh": "scenes/bunny.obj", "triangles": int(tris.shape[0]), "cases": {}}
    small = {}

    cases = [(16, 64, 64), (64, 256, 256), (256, 1024, 1024), (1024, 1920, 1080)]
    if "--big" in sys.argv:
        cases.append((4096, 1920, 1080))
    for grid, W, H in cases:
        origin, dps = scene_params(tris, grid)
        keys, c, e = ref.voxelize(tris, col, emi, origin, float(dps), grid)
        D = int(keys.size)
        mk, mc, me = ref.merge_voxels(keys, c, e)
        U = int(mk.size)
        tree = ref.build_octree(mk, origin, float(dps), grid, dag=True)
        nodes = tree.nodes()
        naive = ref.build_octree(mk, origin, float(dps), grid, dag=False)
        cam = ref.camera_lookat((4, 4, 4), (0, 0, 0), (0, 1, 0), None, W, H)
        t, nm, vi = tree.render_primary(cam, W, H)
        hit = t != orc.FLT_MAX
        order = np.argsort(keys, kind="stable")
        case = {
            "grid": grid, "origin": [float(x) for x in origin], "dps": float(dps), "dps_hex": float(dps).hex(),
            "D": D, "U": U, "naive_nodes": naive.num_nodes, "dag_nodes": tree.num_nodes,
            "fnv_sorted_dumped_keys": fnv(ref, keys[order]),
            "fnv_merged_keys": fnv(ref, mk),
            "fnv_merged_rgb": fnv(ref, np.ascontiguousarray(mc[:, :3])),
            "fnv_merged_emission_rgb": fnv(ref, np.ascontiguousarray(me[:, :3])),
            "fnv_node_mask": fnv(ref, np.ascontiguousarray(nodes["mask"])),
            "fnv_node_children": fnv(ref, np.ascontiguousarray(nodes["children"])),
            "fnv_node_psum": fnv(ref, np.ascontiguousarray(nodes["psum"])),
            "image": [W, H], "camera": {"eye": [4, 4, 4], "center": [0, 0, 0], "fovy_deg": 45},
            "camera_bytes_hex": bytes(cam).hex(),
            "hits": int(hit.sum()),
            "fnv_vindex": fnv(ref, vi),
            "fnv_nmajor": fnv(ref, nm),
            "fnv_t": fnv(ref, t),
            "sum_t_hits": float(np.sum(t[hit].astype(np.float64))),
        }
        golden["cases"][str(grid)] = case
        print(grid, {k: case[k] for k in ("D", "U", "naive_nodes", "dag_nodes", "hits", "fnv_vindex", "sum_t_hits")})
        if grid <= 64:
            small["keys_%d" % grid] = keys
            small["colors_%d" % grid] = c
            small["emissions_%d" % grid] = e
            small["mkeys_%d" % grid] = mk
            small["mcolors_%d" % grid] = mc
            small["memissions_%d" % grid] = me
            small["nodes_%d" % grid] = nodes.view(np.uint8).reshape(-1, 68)
            small["t_%d" % grid] = t
            small["nmajor_%d" % grid] = nm.astype(np.int8)
            small["vindex_%d" % grid] = vi

    # large-grid arithmetic (BASELINE config 4: the 2^10..2^14 sweep) on a bounded input: the first 64 bunny triangles in the
    # full-mesh frame (same origin / dps as the whole bunny at that grid), so Morton keys use 39 / 42 / 45 bits
    golden["subset_cases"] = {}
    NSUB = 64
    for grid in (8192, 16384, 32768):
        origin, dps = scene_params(tris, grid)
        st, sc, se = tris[:NSUB], col[:NSUB], emi[:NSUB]
        keys, c, e = ref.voxelize(st, sc, se, origin, float(dps), grid)
        mk, mc, me = ref.merge_voxels(keys, c, e)
        tree = ref.build_octree(mk, origin, float(dps), grid, dag=True)
        nodes = tree.nodes()
        W, H = 512, 512
        cam = ref.camera_lookat((-2.2, 1.2, 1.0), tuple(float(x) for x in st.reshape(-1, 3).mean(0)), (0, 1, 0), None, W, H)
        t, nm, vi = tree.render_primary(cam, W, H)
        golden["subset_cases"][str(grid)] = {
            "grid": grid, "triangles": NSUB, "D": int(keys.size), "U": int(mk.size), "dag_nodes": tree.num_nodes,
            "fnv_sorted_dumped_keys": fnv(ref, np.sort(keys)), "fnv_merged_keys": fnv(ref, mk),
            "fnv_merged_rgb": fnv(ref, np.ascontiguousarray(mc[:, :3])), "fnv_merged_emission_rgb": fnv(ref, np.ascontiguousarray(me[:, :3])),
            "fnv_node_mask": fnv(ref, np.ascontiguousarray(nodes["mask"])), "fnv_node_children": fnv(ref, np.ascontiguousarray(nodes["children"])),
            "fnv_node_psum": fnv(ref, np.ascontiguousarray(nodes["psum"])),
            "image": [W, H], "camera_bytes_hex": bytes(cam).hex(), "hits": int((t != orc.FLT_MAX).sum()),
            "fnv_vindex": fnv(ref, vi), "fnv_nmajor": fnv(ref, nm), "fnv_t": fnv(ref, t)}
        print("subset", grid, golden["subset_cases"][str(grid)]["D"], golden["subset_cases"][str(grid)]["U"], tree.num_nodes, golden["subset_cases"][str(grid)]["hits"])

    # degenerate / edge-case triangles (SURVEY.md section 7 hard part 3): collinear, point, off-grid, huge
    deg = np.array([
        [[0, 0, 0], [1, 1, 1], [2, 2, 2]],            # collinear
        [[0.5, 0.5, 0.5]] * 3,                         # point
        [[-5, -5, -5], [-4, -5, -5], [-5, -4, -5]],    # outside the grid
        [[0.1, 0.1, 0.1], [0.9, 0.1, 0.1], [0.1, 0.9, 0.1]],   # axis-aligned in-plane
        [[0.1, 0.2, 0.3], [0.9, 0.25, 0.35], [0.2, 0.95, 0.8]],
        [[-3, -3, 0.5], [5, -3, 0.5], [1, 6, 0.5]],    # larger than the grid
    ], np.float32)
    dorigin = np.zeros(3, np.float32)
    for grid in (8, 32):
        ddps = np.float32(1.0 / grid)
        for six in (True, False):
            k, c, e = ref.voxelize(deg, np.ones_like(deg), np.zeros_like(deg), dorigin, float(ddps), grid, six=six)
            small["deg_keys_%d_%d" % (grid, int(six))] = np.sort(k)
            # per-triangle counts
            cnt = [ref.voxelize(deg[i:i + 1], None, None, dorigin, float(ddps), grid, six=six, attributes=False)[0].size
                   for i in range(deg.shape[0])]
            small["deg_counts_%d_%d" % (grid, int(six))] = np.asarray(cnt, np.int64)
    small["deg_tris"] = deg

    # random rays against the 64^3 tree (primary + shadow semantics, negative / zero direction components)
    rng = np.random.default_rng(1234)
    origin, dps = scene_params(tris, 64)
    keys, c, e = ref.voxelize(tris, col, emi, origin, float(dps), 64)
    mk, mc, me = ref.merge_voxels(keys, c, e)
    tree = ref.build_octree(mk, origin, float(dps), 64)
    lo, up = tree.bounds()
    n = 20000
    ro = (lo + (up - lo) * rng.uniform(-0.5, 1.5, (n, 3))).astype(np.float32)
    rd = rng.normal(size=(n, 3)).astype(np.float32)
    rd[::7, 0] = 0.0
    rd[::11, 1] = -0.0
    rd[::13, 2] = 0.0
    rd[5::97] *= np.float32(1e-20)
    t, nm, vi = tree.intersect(ro, rd)
    ts, nms, vis = tree.intersect(ro, rd, shadow=True)
    small["rays_ro"] = ro
    small["rays_rd"] = rd
    small["rays_t"] = t
    small["rays_nmajor"] = nm.astype(np.int8)
    small["rays_vindex"] = vi
    small["rays_shadow_t"] = ts
    small["rays_shadow_vindex"] = vis

    np.savez_compressed(os.path.join(HERE, "small_cases.npz"), **small)
    pt = os.path.join(HERE, "make_golden_pt.py")
    if os.path.exists(pt) and "--no-pt" not in sys.argv:
        import runpy
        golden["pt"] = runpy.run_path(pt)["generate"](ref, tris, col, emi)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(golden, f, indent=1)
    print("written", HERE)


if __name__ == "__main__":
    main()
