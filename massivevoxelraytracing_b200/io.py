"""Host-side mirror of the reference's input / output helpers over the C ABI: mesh loading + trianglesFlattened
+ getBoundingBox + the dps rule (voxUtil.hpp:8-77, voxRT.cpp:188-190), the Radiance .hdr decode behind PathTracer::loadHDRI,
saveAsPngUncompressed, saveVoxelsAsMesh (voxMesh.cpp:111-219) and the RTCamp frame loop (RTCamp.cpp:122-196).
Every function is one call into libhako_b200.so."""
import ctypes as C

import numpy as np

from ._abi import Camera, check, load
from .intersector import _f32, _ptr


def load_mesh(path):
    """OBJ / PLY -> (vertices[T,3,3], vcolors[T,3,3], vemissions[T,3,3]) float32: what trianglesFlattened hands to build()."""
    lib = load()
    h = C.c_void_p()
    check(lib.hako_mesh_load(str(path).encode(), C.byref(h)))
    try:
        nv, nt, hc = C.c_uint64(), C.c_uint64(), C.c_int()
        check(lib.hako_mesh_info(h, C.byref(nv), C.byref(nt), C.byref(hc)))
        pos = np.empty((nt.value, 3, 3), np.float32)
        col = np.empty_like(pos)
        emi = np.empty_like(pos)
        check(lib.hako_mesh_flatten(h, _ptr(pos), _ptr(col), _ptr(emi)))
    finally:
        lib.hako_mesh_destroy(h)
    return pos, col, emi


def bounds_device(ctx, pos_device_ptr, n_vertices, stream=None):
    """getBoundingBox over device-resident vertices -> (lower[3], upper[3])"""
    lo, up = np.empty(3, np.float32), np.empty(3, np.float32)
    check(ctx.lib.hako_bounds_device(ctx.h, stream, C.c_void_p(pos_device_ptr), int(n_vertices), _ptr(lo), _ptr(up)))
    return lo, up


def scene_params(lower, upper, gridRes):
    """origin = bbox lower, dps = max extent / gridRes (voxRT.cpp:188-190)"""
    lo, up = _f32(lower), _f32(upper)
    origin = np.empty(3, np.float32)
    dps = C.c_float()
    check(load().hako_scene_params(_ptr(lo), _ptr(up), int(gridRes), _ptr(origin), C.byref(dps)))
    return origin, np.float32(dps.value)


def load_hdr(path):
    """Radiance .hdr -> float32 [h, w, 4] as stbi_loadf(.., 4) returns it"""
    lib = load()
    p, w, h = C.c_void_p(), C.c_int(), C.c_int()
    check(lib.hako_image_load_hdr(str(path).encode(), C.byref(p), C.byref(w), C.byref(h)))
    try:
        n = w.value * h.value * 4
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n,)).copy().reshape(h.value, w.value, 4)
    finally:
        lib.hako_image_free(p)
    return out


def save_png(path, rgba):
    a = np.ascontiguousarray(rgba, np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4
    check(load().hako_image_save_png(str(path).encode(), _ptr(a), a.shape[1], a.shape[0]))


def save_voxel_mesh(path, sorted_mortons, origin, dps):
    k = np.ascontiguousarray(sorted_mortons, np.uint64)
    o = _f32(origin)
    check(load().hako_voxel_mesh_save_ply(str(path).encode(), _ptr(k), k.size, _ptr(o), float(dps)))


class FrameSequence:
    """RTCamp.cpp:122-196: build -> clear -> N steps -> resolve -> pooled frame buffer -> PNG on a writer thread."""

    def __init__(self, pt, nBuffers=4, nWriters=2):
        self.pt = pt
        self.lib = pt.lib
        h = C.c_void_p()
        check(self.lib.hako_seq_create(pt.h, int(nBuffers), int(nWriters), C.byref(h)))
        self.h = h

    def render_frame(self, vertices, vcolors, vemissions, origin, dps, gridRes, camera, iterations=8, maxDepth=8, png=None, timings=False):
        v, c, e = _f32(vertices), _f32(vcolors), _f32(vemissions)
        o = _f32(origin)
        assert isinstance(camera, Camera)
        um, rm = C.c_float(), C.c_float()
        check(self.lib.hako_seq_render_frame(self.h, _ptr(v), _ptr(c), _ptr(e), v.size // 9, _ptr(o), float(dps), int(gridRes), C.byref(camera),
                                             int(iterations), int(maxDepth), None if png is None else str(png).encode(),
                                             C.byref(um) if timings else None, C.byref(rm) if timings else None))
        return (um.value, rm.value) if timings else None

    def wait(self):
        check(self.lib.hako_seq_wait(self.h))

    def close(self):
        if self.h:
            self.lib.hako_seq_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
