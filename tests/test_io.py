"""Input / output helpers around the hot path: .hdr decode, PNG writer, OBJ / PLY readers + trianglesFlattened,
the dps rule, the voxel-face PLY export; on the GPU: bounding-box reduction and the RTCamp frame loop."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from scenes import scene_params

REF = "/root/reference"


@pytest.fixture(scope="module")
def hk():
    import massivevoxelraytracing_b200 as h
    return h


# ---------------------------------------------------------------- Radiance .hdr
def _rgbe_expected(rgbe):
    e = rgbe[..., 3].astype(np.int32)
    f = np.where(e != 0, np.ldexp(np.float32(1.0), e - 136), 0.0).astype(np.float32)
    out = np.ones(rgbe.shape[:2] + (4,), np.float32)
    out[..., :3] = rgbe[..., :3].astype(np.float32) * f[..., None]
    return out


def _rle_scanline(row):
    """new-style RLE: 2 2 hi lo, then each channel as runs (count > 128) / literals"""
    w = row.shape[0]
    out = bytearray([2, 2, w >> 8, w & 255])
    for c in range(4):
        v = row[:, c]
        x = 0
        while x < w:
            run = 1
            while x + run < w and run < 127 and v[x + run] == v[x]:
                run += 1
            if run >= 3:
                out += bytes([128 + run, int(v[x])])
                x += run
            else:
                lit = 1
                while x + lit < w and lit < 128 and not (x + lit + 2 < w and v[x + lit] == v[x + lit + 1] == v[x + lit + 2]):
                    lit += 1
                out += bytes([lit]) + bytes(int(t) for t in v[x:x + lit])
                x += lit
    return bytes(out)


@pytest.mark.parametrize("w,h,rle", [(64, 32, True), (40, 7, False), (5, 3, False), (300, 11, True)])
def test_hdr_decode(hk, tmp_path, w, h, rle):
    rng = np.random.default_rng(w * 131 + h)
    rgbe = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    rgbe[::3, ::5, 3] = 0                      # black pixels (exponent 0)
    rgbe[1::4, :, 0] = 77                      # long runs
    rgbe[:, w // 2:, 2] = 9
    body = b"".join(_rle_scanline(rgbe[y]) for y in range(h)) if rle else rgbe.tobytes()
    if not rle and w >= 8:
        rgbe[0, 0, :3] = (1, 1, 1)             # a flat file must not start with the RLE marker 2 2
        body = rgbe.tobytes()
    p = tmp_path / "t.hdr"
    p.write_bytes(b"#?RADIANCE\nSOFTWARE=test\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w) + body)
    got = hk.io.load_hdr(p)
    assert got.shape == (h, w, 4)
    assert np.array_equal(got.view(np.uint32), _rgbe_expected(rgbe).view(np.uint32))


def test_hdr_errors(hk, tmp_path):
    p = tmp_path / "bad.hdr"
    p.write_bytes(b"P6\n")
    with pytest.raises(hk.HakoError):
        hk.io.load_hdr(p)
    p.write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 4 +X 16\n" + bytes([2, 2, 0, 16, 200]))
    with pytest.raises(hk.HakoError):
        hk.io.load_hdr(p)
    with pytest.raises(hk.HakoError):
        hk.io.load_hdr(tmp_path / "missing.hdr")


@pytest.mark.skipif(not os.path.exists(REF + "/libs/prlib/libs/stb/stb_image.h"), reason="needs the reference checkout (stb_image.h, monks_forest_s.hdr)")
def test_hdr_matches_stbi_loadf(hk, tmp_path):
    """the reference decodes through stbi_loadf(.., 4) (libs/prlib/src/prp.cpp:353-366): compile stb_image.h where it lies and compare
    on the reference's own environment map"""
    src = tmp_path / "stb_drv.c"
    src.write_text('#define STB_IMAGE_IMPLEMENTATION\n#define STBI_ONLY_HDR\n#include "stb_image.h"\n'
                   'float* drv_load(const char* f, int* w, int* h) { int c; return stbi_loadf(f, w, h, &c, 4); }\n')
    so = tmp_path / "libstb_drv.so"
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-I" + REF + "/libs/prlib/libs/stb", str(src), "-o", str(so), "-lm"], check=True)
    import ctypes as C
    lib = C.CDLL(str(so))
    lib.drv_load.restype = C.POINTER(C.c_float)
    w, h = C.c_int(), C.c_int()
    f = REF + "/bin/monks_forest_s.hdr"
    ptr = lib.drv_load(f.encode(), C.byref(w), C.byref(h))
    ref = np.ctypeslib.as_array(ptr, shape=(h.value, w.value, 4)).copy()
    got = hk.io.load_hdr(f)
    assert got.shape == ref.shape and np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_hdr_matches_stbi_golden(hk, tmp_path):
    """the same comparison without the reference checkout: tests/golden/hdr_golden.npz holds small synthetic .hdr files and what the
    reference's decoder (stbi_loadf( .., 4 ), libs/prlib/src/prp.cpp:353-366) returned for them (tests/golden/make_hdr_golden.py)"""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hdr_golden.npz"))
    cases = sorted(k for k in g.files if k.startswith("file"))
    assert len(cases) == 5
    for k in cases:
        p = tmp_path / (k + ".hdr")
        p.write_bytes(g[k].tobytes())
        ref = g["image" + k[4:]]
        got = hk.io.load_hdr(p)
        assert got.shape == ref.shape and np.array_equal(got.view(np.uint32), ref.view(np.uint32)), k


# ---------------------------------------------------------------- PNG
def _read_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + data)
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", data[:10])
            assert (depth, ctype) == (8, 6)
        if typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * 4 + 1)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4)


@pytest.mark.parametrize("w,h", [(1, 1), (37, 19), (640, 360)])
def test_png_roundtrip(hk, tmp_path, w, h):
    img = np.random.default_rng(w).integers(0, 256, (h, w, 4), dtype=np.uint8)
    hk.io.save_png(tmp_path / "a.png", img)
    assert np.array_equal(_read_png(tmp_path / "a.png"), img)


# ---------------------------------------------------------------- meshes
def test_mesh_obj(hk, tmp_path):
    p = tmp_path / "m.obj"
    p.write_text("# comment\nv 0 0 0 1 0 0\nv 1 0 0 0 1 0\nv 1 1 0 0 0 1\nv 0 1 0.5 1 1 1\nvn 0 0 1\n"
                 "f 1 2 3\nf 1/1/1 3/2/1 4/3/1\nf -4 -3 -2 -1\n")
    pos, col, emi = hk.io.load_mesh(p)
    v = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0.5]], np.float32)
    c = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 1]], np.float32)
    idx = np.array([[0, 1, 2], [0, 2, 3], [0, 1, 2], [0, 2, 3]])
    assert np.array_equal(pos, v[idx]) and np.array_equal(col, c[idx]) and (emi == 0).all()
    p.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    pos, col, emi = hk.io.load_mesh(p)
    assert pos.shape == (1, 3, 3) and (col == 1).all()          # no Cd -> colour 1 (voxUtil.hpp:49-52)
    p.write_text("v 0 0 0\nf 1 2 3\n")
    with pytest.raises(hk.HakoError):
        hk.io.load_mesh(p)


@pytest.mark.parametrize("fmt", ["ascii", "binary_little_endian", "binary_big_endian"])
def test_mesh_ply(hk, tmp_path, fmt):
    rng = np.random.default_rng(5)
    v = rng.normal(size=(6, 3)).astype(np.float32)
    c = rng.integers(0, 256, (6, 3), dtype=np.uint8)
    faces = [[0, 1, 2], [2, 3, 4, 5], [5, 0, 3]]
    hdr = ("ply\nformat %s 1.0\ncomment test\nelement vertex 6\nproperty float x\nproperty float y\nproperty float z\nproperty float confidence\n"
           "property uchar red\nproperty uchar green\nproperty uchar blue\nelement face 3\nproperty list uchar int vertex_indices\nend_header\n" % fmt)
    if fmt == "ascii":
        body = "".join("%r %r %r 0.5 %d %d %d\n" % (float(a[0]), float(a[1]), float(a[2]), b[0], b[1], b[2]) for a, b in zip(v, c))
        body += "".join("%d %s\n" % (len(f), " ".join(map(str, f))) for f in faces)
        data = (hdr + body).encode()
    else:
        e = "<" if fmt.endswith("little_endian") else ">"
        data = hdr.encode()
        for a, b in zip(v, c):
            data += struct.pack(e + "4f3B", a[0], a[1], a[2], 0.5, *[int(t) for t in b])
        for f in faces:
            data += struct.pack(e + "B%di" % len(f), len(f), *f)
    p = tmp_path / "m.ply"
    p.write_bytes(data)
    pos, col, emi = hk.io.load_mesh(p)
    idx = np.array([[0, 1, 2], [2, 3, 4], [2, 4, 5], [5, 0, 3]])
    assert np.array_equal(pos, v[idx])
    assert np.array_equal(col, (c.astype(np.float32) * np.float32(1.0 / 255.0))[idx])
    assert (emi == 0).all()


@pytest.mark.skipif(not os.path.exists(REF + "/scenes/bunny.obj"), reason="needs the reference checkout")
def test_mesh_loader_reads_the_reference_bunny(hk, bunny):
    pos, col, emi = hk.io.load_mesh(REF + "/scenes/bunny.obj")
    assert np.array_equal(pos.view(np.uint32), bunny[0].view(np.uint32))        # the committed fixture = the reference's triangle soup
    assert (col == 1).all() and (emi == 0).all()


def test_scene_params_rule(hk, bunny):
    tris = bunny[0]
    p = tris.reshape(-1, 3)
    for grid in (16, 256, 4096, 32768):
        origin, dps = hk.io.scene_params(p.min(0), p.max(0), grid)
        o2, d2 = scene_params(tris, grid)
        assert np.array_equal(origin, o2) and np.float32(dps).tobytes() == np.float32(d2).tobytes()


def test_voxel_mesh_export(hk, oracle, bunny, tmp_path):
    tris, col, emi = bunny
    grid = 16
    origin, dps = scene_params(tris, grid)
    keys, c, e = oracle.voxelize(tris, col, emi, origin, float(dps), grid)
    mk = oracle.merge_voxels(keys, c, e)[0]
    f = tmp_path / "vox.ply"
    hk.io.save_voxel_mesh(f, mk, origin, dps)
    b = f.read_bytes()
    end = b.index(b"end_header\n") + 11
    hdr = b[:end].decode()
    nv = int(hdr.split("element vertex ")[1].split("\n")[0])
    nf = int(hdr.split("element face ")[1].split("\n")[0])
    assert nv == mk.size * 8 and "format binary_little_endian 1.0" in hdr
    pts = np.frombuffer(b, np.float32, nv * 3, end).reshape(-1, 8, 3)
    fa = np.frombuffer(b, np.uint8, nf * 17, end + nv * 12).reshape(nf, 17)
    assert (fa[:, 0] == 4).all() and len(b) == end + nv * 12 + nf * 17
    # exposed faces from a dense occupancy grid
    def compact(x):
        x = x & np.uint64(0x1249249249249249)
        for s, m in ((2, 0x10c30c30c30c30c3), (4, 0x100f00f00f00f00f), (8, 0x1f0000ff0000ff), (16, 0x1f00000000ffff), (32, 0x1fffff)):
            x = (x ^ (x >> np.uint64(s))) & np.uint64(m)
        return x.astype(np.int64)
    x, y, z = compact(mk), compact(mk >> np.uint64(1)), compact(mk >> np.uint64(2))
    occ = np.zeros((grid + 2,) * 3, bool)
    occ[x + 1, y + 1, z + 1] = True
    exposed = sum(int((~occ[x + 1 + dx, y + 1 + dy, z + 1 + dz]).sum()) for dx, dy, dz in ((1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)))
    assert nf == exposed
    lower = np.stack([origin[0] + x.astype(np.float32) * np.float32(dps), origin[1] + y.astype(np.float32) * np.float32(dps),
                      origin[2] + z.astype(np.float32) * np.float32(dps)], 1).astype(np.float32)
    assert np.array_equal(pts[:, 0], lower)
    assert np.allclose(pts[:, 6] - pts[:, 0], dps)


# ---------------------------------------------------------------- GPU: bounds + frame loop
@pytest.mark.gpu
def test_bounds_device(hk, bunny):
    import torch
    ctx = hk.Context(0)
    tris = bunny[0]
    t = torch.from_numpy(tris.reshape(-1, 3).copy()).cuda()
    lo, up = hk.io.bounds_device(ctx, t.data_ptr(), t.shape[0])
    p = tris.reshape(-1, 3)
    assert np.array_equal(lo, p.min(0)) and np.array_equal(up, p.max(0))
    lo, up = hk.io.bounds_device(ctx, t.data_ptr(), 0)
    assert (lo == np.float32(3.402823466e38)).all() and (up == -np.float32(3.402823466e38)).all()      # voxUtil.hpp:68-69
    ctx.close()


@pytest.mark.gpu
def test_frame_sequence_matches_direct_calls(hk, bunny, tmp_path):
    """RTCamp frame loop: animated resolution, PNG per frame written by the pool threads == the same frames through the plain
    PathTracer calls.  This compares the library with itself on purpose: it checks the ORCHESTRATION (build of frame f + 1 overlapped with
    the steps of frame f, resolve, asynchronous download, PNG writer threads); what a step computes is pinned elsewhere
    (tests/test_gpu_pt.py against the oracle, tests/test_gpu_refgpu.py against the reference's own kernels)"""
    from scenes import procedural_hdri
    tris, col, emi = bunny
    ctx = hk.Context(0)
    W, H = 160, 96
    env = procedural_hdri(128, 64)
    cam = hk.camera_lookat((4, 4, 4), (0, 0, 0), width=W, height=H, focus=6.0, lensR=0.01)
    pt = hk.PathTracer(ctx).setup()
    pt.resizeFrameBufferIfNeeded(None, W, H)
    pt.loadHDRI(None, env)
    seq = hk.io.FrameSequence(pt, nBuffers=2, nWriters=2)
    grids = [32, 64, 128, 64, 256]
    for f, grid in enumerate(grids):
        origin, dps = scene_params(tris, grid)
        tm = seq.render_frame(tris, col, emi, origin, dps, grid, cam, iterations=2, maxDepth=4, png=tmp_path / ("%03d.png" % f), timings=(f == 1))
        assert tm is None or (tm[0] > 0 and tm[1] > 0)
    seq.wait()
    seq.close()
    pt2 = hk.PathTracer(ctx).setup()
    pt2.resizeFrameBufferIfNeeded(None, W, H)
    pt2.loadHDRI(None, env)
    pt2.maxDepth = 4
    for f, grid in enumerate(grids):
        origin, dps = scene_params(tris, grid)
        pt2.updateScene(tris, col, emi, None, origin, float(dps), grid)
        pt2.clearFrameBuffer()
        for _ in range(2):
            pt2.step(None, cam)
        assert np.array_equal(_read_png(tmp_path / ("%03d.png" % f)), pt2.toImageAsync()), f
    pt2.cleanUp()
    pt.cleanUp()
    ctx.close()
