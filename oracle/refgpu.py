"""ctypes binding of oracle/librefgpu.so: the REFERENCE's own GPU kernels (voxKernel.cu compiled ahead of time for sm_100a into
oracle/_ref/voxKernel.sm_100a*.cubin by oracle/Makefile) launched through the CUDA driver API.

TEST INFRASTRUCTURE -- never imported by the product package.  Used by tests/ (to pin the path-tracing half: SAT tables,
renderPT accumulators, renderResolve) and by bench.py's `reference_gpu` block (the same-box GPU baseline).
"""
import ctypes as C
import os

import numpy as np

from .orc import Camera, HERE, NODE68

VARIANTS = {
    "default": "voxKernel.sm_100a.cubin",      # as the reference ships: fmad on, depth < 8
    "nofma": "voxKernel.sm_100a.nofma.cubin",  # -fmad=false
    "d4": "voxKernel.sm_100a.d4.cubin",        # depth < 4 (BASELINE config 3)
}


def cubin_path(variant="default"):
    return os.path.join(HERE, "_ref", VARIANTS[variant])


def available(variant="default"):
    return os.path.exists(os.path.join(HERE, "librefgpu.so")) and os.path.exists(cubin_path(variant))


class RefGpuError(RuntimeError):
    pass


def _lib():
    L = C.CDLL(os.path.join(HERE, "librefgpu.so"))
    L.refgpu_last_error.restype = C.c_char_p
    vp, i, u32, f, fp = C.c_void_p, C.c_int, C.c_uint32, C.c_float, C.POINTER(C.c_float)
    L.refgpu_open.argtypes = [C.c_char_p, i, C.POINTER(vp)]
    L.refgpu_close.argtypes = [vp]
    L.refgpu_close.restype = None
    L.refgpu_set_octree.argtypes = [vp, vp, u32, vp, u32, vp, vp, f, u32]
    L.refgpu_set_attrs.argtypes = [vp, vp, u32]
    L.refgpu_render.argtypes = [vp, C.POINTER(Camera), i, i, i, vp, i, fp, fp]
    L.refgpu_load_hdri.argtypes = [vp, vp, i, i, vp, i, i, fp]
    L.refgpu_get_sats.argtypes = [vp, vp]
    L.refgpu_set_pmj.argtypes = [vp, vp]
    L.refgpu_pt_resize.argtypes = [vp, i, i]
    L.refgpu_pt_clear.argtypes = [vp]
    L.refgpu_pt_step.argtypes = [vp, C.POINTER(Camera), i, fp]
    L.refgpu_pt_read_accum.argtypes = [vp, vp]
    L.refgpu_pt_write_accum.argtypes = [vp, vp]
    L.refgpu_pt_resolve.argtypes = [vp, vp]
    L.refgpu_set_scales.argtypes = [vp, f, f]
    L.refgpu_set_scales.restype = None
    return L


class RefGpu:
    """One loaded module of the reference's kernels + the state PathTracer / voxRTGPU keep around them."""

    def __init__(self, variant="default", device=0):
        self.lib = _lib()
        self.variant = variant
        h = C.c_void_p()
        self._ck(self.lib.refgpu_open(cubin_path(variant).encode(), int(device), C.byref(h)))
        self.h = h
        self.hdri_size = None
        self.w = self.h_ = 0

    def _ck(self, r):
        if r != 0:
            raise RefGpuError("%d: %s" % (r, (self.lib.refgpu_last_error() or b"").decode()))

    def close(self):
        if getattr(self, "h", None):
            self.lib.refgpu_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _cam(cam):
        """accepts oracle.orc.Camera or the package's Camera (identical 60-byte layout)"""
        c = Camera()
        assert C.sizeof(cam) == 60
        C.memmove(C.byref(c), C.byref(cam), 60)
        return c

    def set_octree(self, nodes68, attrs, lower, upper, dps, has_emission):
        n = np.ascontiguousarray(nodes68)
        assert n.dtype.itemsize == NODE68.itemsize
        a = np.ascontiguousarray(attrs)
        assert a.dtype.itemsize * (a.size // max(a.shape[0], 1)) == 8 or a.size == 0
        lo, up = np.ascontiguousarray(lower, np.float32), np.ascontiguousarray(upper, np.float32)
        self.n_voxels = a.shape[0]
        self._ck(self.lib.refgpu_set_octree(self.h, n.ctypes.data, n.shape[0], a.ctypes.data, a.shape[0], lo.ctypes.data, up.ctypes.data,
                                            float(dps), int(bool(has_emission))))

    def set_attrs(self, attrs):
        a = np.ascontiguousarray(attrs)
        self._ck(self.lib.refgpu_set_attrs(self.h, a.ctypes.data, a.shape[0]))

    def render(self, cam, w, h, show_vertex_color=False, reps=1):
        """`render` (voxKernel.cu:437-483) -> (rgba uint8 [h, w, 4], best ms, mean ms)"""
        out = np.empty((h, w, 4), np.uint8)
        best, mean = C.c_float(), C.c_float()
        c = self._cam(cam)
        self._ck(self.lib.refgpu_render(self.h, C.byref(c), w, h, int(bool(show_vertex_color)), out.ctypes.data, int(reps), C.byref(best), C.byref(mean)))
        return out, best.value, mean.value

    def render_ids(self, cam, w, h):
        """hit voxel ranks read out of the reference's `render` kernel: the colour bytes of voxel i are set to the bytes of i and the
        frame is rendered with showVertexColor; a miss writes (0,0,0,255) = 0xFF000000, which no rank reaches (< 2^24 voxels asserted)"""
        assert self.n_voxels < 0xFF000000
        enc = np.zeros((self.n_voxels, 2), np.uint32)
        enc[:, 0] = np.arange(self.n_voxels, dtype=np.uint32)
        self.set_attrs(enc.view(np.uint8).reshape(-1, 8))
        img, _, _ = self.render(cam, w, h, True)
        ids = img.reshape(h, w, 4).view(np.uint32).reshape(h, w).copy()
        ids[ids == 0xFF000000] = 0xFFFFFFFF
        return ids

    def load_hdri(self, pixels, primary=None):
        p = np.ascontiguousarray(pixels, np.float32)
        q = None if primary is None else np.ascontiguousarray(primary, np.float32)
        ms = C.c_float()
        self._ck(self.lib.refgpu_load_hdri(self.h, p.ctypes.data, p.shape[1], p.shape[0], None if q is None else q.ctypes.data,
                                           0 if q is None else q.shape[1], 0 if q is None else q.shape[0], C.byref(ms)))
        self.hdri_size = (p.shape[1], p.shape[0])
        return ms.value

    def sats(self):
        w, h = self.hdri_size
        out = np.empty(7 * w * h, np.uint32)
        self._ck(self.lib.refgpu_get_sats(self.h, out.ctypes.data))
        return out

    def set_pmj(self, table):
        t = np.ascontiguousarray(table, np.float32)
        assert t.size == 128 * 4096 * 2
        self._ck(self.lib.refgpu_set_pmj(self.h, t.ctypes.data))

    def set_scales(self, emission_scale=7.5, hdri_scale=1.75):
        self.lib.refgpu_set_scales(self.h, float(emission_scale), float(hdri_scale))

    def pt_resize(self, w, h):
        self._ck(self.lib.refgpu_pt_resize(self.h, int(w), int(h)))
        self.w, self.h_ = int(w), int(h)

    def pt_clear(self):
        self._ck(self.lib.refgpu_pt_clear(self.h))

    def pt_step(self, cam, iteration):
        """`renderPT` (voxKernel.cu:610-777): 16 spp; returns the device time of the launch in ms"""
        ms = C.c_float()
        c = self._cam(cam)
        self._ck(self.lib.refgpu_pt_step(self.h, C.byref(c), int(iteration), C.byref(ms)))
        return ms.value

    def pt_accum(self):
        out = np.empty((self.h_, self.w, 4), np.float32)
        self._ck(self.lib.refgpu_pt_read_accum(self.h, out.ctypes.data))
        return out

    def pt_write_accum(self, acc):
        a = np.ascontiguousarray(acc, np.float32)
        assert a.size == self.w * self.h_ * 4
        self._ck(self.lib.refgpu_pt_write_accum(self.h, a.ctypes.data))

    def pt_resolve(self):
        out = np.empty((self.h_, self.w, 4), np.uint8)
        self._ck(self.lib.refgpu_pt_resolve(self.h, out.ctypes.data))
        return out
