"""Dense-grid Amanatides-Woo comparator for the octree traversal (the reference's own comparator, IntersectorEmbree,
needs an Embree binary that is not in the checkout).  Shared by the oracle test (CPU) and the CUDA-path test (GPU)."""
import numpy as np

FLT_MAX = np.float32(3.402823466e38)


def dda_first_hits(occ, lower, dps, ro, rd):
    """Amanatides-Woo march over the DENSE occupancy grid (float64): first occupied cell per ray, -1 when none.
    An independent comparator for the octree traversal (the reference's comparator, IntersectorEmbree,
    needs an Embree binary that is not in the checkout)."""
    n = occ.shape[0]
    ro = ro.astype(np.float64)
    rd = rd.astype(np.float64)
    lo = lower.astype(np.float64)
    up = lo + float(dps) * n
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / rd
        t0 = (lo - ro) * inv
        t1 = (up - ro) * inv
    tmin = np.minimum(t0, t1).max(1)
    tmax = np.maximum(t0, t1).min(1)
    alive = (tmax >= np.maximum(tmin, 0.0)) & (tmax > 0)
    tstart = np.maximum(tmin, 0.0)
    p = ro + rd * (tstart[:, None] + 1e-9)
    cell = np.clip(np.floor((p - lo) / float(dps)).astype(np.int64), 0, n - 1)
    step = np.where(rd > 0, 1, -1)
    nxt = lo + (cell + (step > 0)) * float(dps)
    with np.errstate(divide="ignore", invalid="ignore"):
        tnext = np.where(rd != 0, (nxt - ro) * inv, np.inf)
        tdelta = np.where(rd != 0, float(dps) * np.abs(inv), np.inf)
    hit = np.full((ro.shape[0], 3), -1, np.int64)
    done = ~alive
    for _ in range(3 * n + 3):
        inside = ((cell >= 0) & (cell < n)).all(1) & ~done
        c = np.clip(cell, 0, n - 1)
        o = occ[c[:, 0], c[:, 1], c[:, 2]] & inside
        hit[o] = cell[o]
        done |= o | ~inside
        if done.all():
            break
        ax = np.argmin(tnext, 1)
        r = np.arange(ro.shape[0])
        cell[r, ax] += step[r, ax]
        tnext[r, ax] += tdelta[r, ax]
    return hit


def compact3(x):
    """every third bit of a Morton key (morton.hpp:101-115 inverse)"""
    x = x & np.uint64(0x1249249249249249)
    for s, m in ((2, 0x10c30c30c30c30c3), (4, 0x100f00f00f00f00f), (8, 0x1f0000ff0000ff), (16, 0x1f00000000ffff), (32, 0x1fffff)):
        x = (x ^ (x >> np.uint64(s))) & np.uint64(m)
    return x.astype(np.int64)


def check_against_dda(intersect, mk, origin, dps, grid, n=20000, seed=7):
    """intersect(ro, rd) -> (t, nMajor, vIndex) with the reference's conventions, on the octree of the sorted unique keys `mk`.
    Asserts agreement with the float64 DDA on > 99.9 % of `n` random rays (the rest pass through a cell corner, where the
    float64 march and the fp32 parametric order legitimately differ); returns (rays compared, both-hit rays)."""
    xyz = np.stack([compact3(mk), compact3(mk >> np.uint64(1)), compact3(mk >> np.uint64(2))], 1)
    occ = np.zeros((grid,) * 3, bool)
    occ[xyz[:, 0], xyz[:, 1], xyz[:, 2]] = True
    rng = np.random.default_rng(seed)
    lo = np.asarray(origin, np.float32)
    up = lo + np.float32(dps) * grid
    ro = (lo + (up - lo) * rng.uniform(-0.6, 1.6, (n, 3))).astype(np.float32)
    target = (lo + (up - lo) * rng.uniform(0.1, 0.9, (n, 3))).astype(np.float32)
    rd = (target - ro).astype(np.float32)
    t, nm, vi = intersect(ro, rd)
    dda = dda_first_hits(occ, lo, dps, ro, rd)
    hit_o = t != FLT_MAX
    hit_d = dda[:, 0] >= 0
    # rays that start inside an occupied voxel: the parametric traversal reports the next surface it ENTERS (0 < t), the DDA the cell itself
    start_cell = np.floor((ro.astype(np.float64) - lo) / float(dps)).astype(np.int64)
    inside_start = ((start_cell >= 0) & (start_cell < grid)).all(1)
    sc = np.clip(start_cell, 0, grid - 1)
    starts_in_voxel = inside_start & occ[sc[:, 0], sc[:, 1], sc[:, 2]]
    ok = ~starts_in_voxel
    assert (hit_o == hit_d)[ok].mean() > 0.999
    both = ok & hit_o & hit_d
    assert (xyz[vi[both]] == dda[both]).all(1).mean() > 0.999
    return int(ok.sum()), int(both.sum())
