"""Deterministic test/bench scenes (synthetic; no dataset access).  Shared by tests/ and bench.py."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def vertex_colors(tris):
    """Same rule as tests/golden/make_golden.py:vertex_colors."""
    p = tris.reshape(-1, 3)
    lo, hi = p.min(0), p.max(0)
    col = ((p - lo) / (hi - lo)).astype(np.float32)
    emi = np.zeros_like(col)
    emi[p[:, 1] > (lo[1] + 0.8 * (hi[1] - lo[1]))] = np.float32(0.75)
    return col.reshape(-1, 3, 3), emi.reshape(-1, 3, 3)


def bunny_scene():
    tris = np.load(os.path.join(GOLDEN, "bunny_tris.npz"))["tris"]
    col, emi = vertex_colors(tris)
    return tris, col, emi


def scene_params(tris, grid):
    """origin = bbox lower, dps = max extent / gridRes (voxRT.cpp:188-190)."""
    p = tris.reshape(-1, 3)
    lo = p.min(0)
    hi = p.max(0)
    size = hi - lo
    dps = np.float32(np.float32(max(size[0], size[1], size[2])) / np.float32(grid))
    return lo.astype(np.float32), dps


def _hash01(a):
    """Cheap deterministic integer hash of float32 bit patterns -> [0,1) float32 (position-keyed noise)."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    h = (u[..., 0] * np.uint64(0x9E3779B1) ^ u[..., 1] * np.uint64(0x85EBCA77) ^ u[..., 2] * np.uint64(0xC2B2AE3D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    return (h.astype(np.float64) / 4294967296.0).astype(np.float32)


def subdivide(tris, levels, noise=0.0):
    """Midpoint subdivision (x4 triangles per level) with position-hash displacement of the new
    vertices (keyed on the edge midpoint so shared edges stay watertight)."""
    t = np.ascontiguousarray(tris, np.float32)
    for lv in range(levels):
        a, b, c = t[:, 0], t[:, 1], t[:, 2]
        ab, bc, ca = (a + b) * np.float32(0.5), (b + c) * np.float32(0.5), (c + a) * np.float32(0.5)
        if noise > 0.0:
            for m, (p, q) in ((ab, (a, b)), (bc, (b, c)), (ca, (c, a))):
                ln = np.linalg.norm(p - q, axis=1).astype(np.float32)
                d = np.stack([_hash01(m), _hash01(m[:, ::-1]), _hash01(m * np.float32(1.7))], 1) - np.float32(0.5)
                m += d * (ln * np.float32(noise))[:, None]
        t = np.concatenate([np.stack([a, ab, ca], 1), np.stack([ab, b, bc], 1), np.stack([ca, bc, c], 1), np.stack([ab, bc, ca], 1)], 0)
    return np.ascontiguousarray(t, np.float32)


def dragon_standin(levels=5, noise=0.15):
    """Synthetic stand-in for the missing xyzrgb_dragon blob: the bunny
    midpoint-subdivided `levels` times (5110 * 4^levels triangles) with hash-noise displacement and
    position-derived vertex colours.  If $HAKO_ASSETS/xyzrgb_dragon.ply exists it is NOT read here;
    see massivevoxelraytracing_b200.io for the PLY path."""
    tris, _, _ = bunny_scene()
    t = subdivide(tris, levels, noise)
    col, emi = vertex_colors(t)
    return t, col, emi


def procedural_hdri(w=256, h=128, seed=0):
    """Procedural lat-long sky + sun (RGBA32F, float32 [h,w,4]) standing in for the missing monks_forest_2k.hdr
    (SURVEY.md section 8d): gradient sky, a bright sun lobe, hash noise so no two texels tie."""
    y = (np.arange(h, dtype=np.float32) + np.float32(0.5)) / np.float32(h)
    x = (np.arange(w, dtype=np.float32) + np.float32(0.5)) / np.float32(w)
    X, Y = np.meshgrid(x, y)
    theta = Y * np.float32(np.pi)
    phi = X * np.float32(2 * np.pi)
    d = np.stack([np.cos(phi) * np.sin(theta), np.cos(theta), np.sin(phi) * np.sin(theta)], -1).astype(np.float32)
    sun = np.asarray([0.5, 0.7, 0.3], np.float32)
    sun /= np.linalg.norm(sun)
    c = np.clip((d * sun).sum(-1), 0, 1).astype(np.float32)
    sky = np.float32(0.25) + np.float32(0.75) * np.clip(d[..., 1], 0, 1)
    rng = np.random.default_rng(seed)
    noise = rng.random((h, w), dtype=np.float32) * np.float32(0.05)
    img = np.zeros((h, w, 4), np.float32)
    img[..., 0] = sky * np.float32(0.6) + np.float32(40.0) * c ** 64 + noise
    img[..., 1] = sky * np.float32(0.8) + np.float32(36.0) * c ** 64 + noise
    img[..., 2] = sky * np.float32(1.0) + np.float32(30.0) * c ** 64 + noise
    img[..., 3] = 1.0
    return np.ascontiguousarray(img)
