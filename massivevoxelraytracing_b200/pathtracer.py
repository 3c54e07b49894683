"""Host-side mirror of PathTracer (PathTracer.hpp:14-170) over the C ABI: setup / cleanUp / resizeFrameBufferIfNeeded /
clearFrameBuffer / loadHDRI / updateScene / step / resolve / toImageAsync, with the reference's names and call order
(voxPTGPU.cpp:88-91,169,177-193; RTCamp.cpp:111-123,147-169).  Nothing here computes: every method is one call into
libhako_b200.so.

Differences a caller sees: no kernel source / include dir / isNvidia arguments (ahead-of-time sm_100a code), no 1 GiB
stack pool; `step` takes the camera either as a ready `Camera` (CameraPinhole) or as look-at parameters (what
pr::GetCameraMatrix derives from a pr::Camera3D); `loadHDRI` takes RGBA32F pixels (decode the .hdr with any reader) and
optional separate primary-ray pixels; `step(..., rows=(y0, y1))` and `iteration=` expose the tile / spp sharding used
across GPUs.
"""
import ctypes as C

import numpy as np

from ._abi import Camera, check
from .camera import camera_lookat
from .intersector import IntersectorOctreeGPU, _f32, _ptr, _stream


class PathTracer:
    def __init__(self, ctx):
        self.ctx = ctx
        self.lib = ctx.lib
        self.h = None
        self.m_intersectorOctreeGPU = IntersectorOctreeGPU(ctx)
        self.m_width = 0
        self.m_height = 0
        self.m_steps = 0
        self.maxDepth = 8  # voxKernel.cu:691

    # PathTracer::setup, PathTracer.hpp:43-69 (PMJ table generation + upload)
    def setup(self, stream=None):
        self.cleanUp()
        h = C.c_void_p()
        check(self.lib.hako_pt_create(self.ctx.h, _stream(stream), C.byref(h)))
        self.h = h
        return self

    def cleanUp(self):
        if self.h:
            self.lib.hako_pt_destroy(self.h)
            self.h = None
        self.m_intersectorOctreeGPU.cleanUp()

    def __del__(self):
        try:
            self.cleanUp()
        except Exception:
            pass

    def getSteps(self):
        return self.m_steps

    def getNumberOfVoxels(self):
        return int(self.m_intersectorOctreeGPU.m_numberOfVoxels)

    def getOctreeBytes(self):
        return int(self.m_intersectorOctreeGPU.m_numberOfNodes) * 68  # sizeof(OctreeNode), PathTracer.hpp:38-41

    def resizeFrameBufferIfNeeded(self, stream, width, height):
        if (width, height) != (self.m_width, self.m_height):
            self.m_steps = 0  # a new buffer is cleared, PathTracer.hpp:84-88
        check(self.lib.hako_pt_resize(self.h, _stream(stream), int(width), int(height)))
        self.m_width, self.m_height = int(width), int(height)

    def clearFrameBuffer(self, stream=None):
        self.m_steps = 0
        check(self.lib.hako_pt_clear(self.h, _stream(stream)))

    def loadHDRI(self, stream, pixels, pixelsPrimary=None):
        """pixels: a Radiance .hdr file name (PathTracer.hpp:104-116 takes file names) or float32 [h, w, 4] (RGBA32F, as
        pr::Image2DRGBA32::loadFromHDR yields); builds the 7 SATs."""
        if isinstance(pixels, (str, bytes)) or hasattr(pixels, "__fspath__"):
            from .io import load_hdr
            pixels = load_hdr(pixels)
        if isinstance(pixelsPrimary, (str, bytes)) or hasattr(pixelsPrimary, "__fspath__"):
            from .io import load_hdr
            pixelsPrimary = load_hdr(pixelsPrimary)
        p = _f32(pixels)
        assert p.ndim == 3 and p.shape[2] == 4
        q = None if pixelsPrimary is None else _f32(pixelsPrimary)
        check(self.lib.hako_pt_load_hdri(self.h, _stream(stream), _ptr(p), p.shape[1], p.shape[0], _ptr(q),
                                         0 if q is None else q.shape[1], 0 if q is None else q.shape[0]))

    def updateScene(self, vertices, vcolors, vemissions, stream, origin, dps, gridRes):
        self.m_intersectorOctreeGPU.build(vertices, vcolors, vemissions, stream, origin, dps, gridRes)

    def step(self, stream, camera, focus=1.0, lensR=0.0, rows=None, iteration=None):
        """One call = 16 spp over rows [y0, y1) (default: the whole frame); increments m_steps like the reference."""
        if not isinstance(camera, Camera):
            camera = camera_lookat(width=self.m_width, height=self.m_height, focus=focus, lensR=lensR, **camera)
        y0, y1 = (0, self.m_height) if rows is None else rows
        it = self.m_steps if iteration is None else int(iteration)
        check(self.lib.hako_pt_step(self.h, self.m_intersectorOctreeGPU.h, _stream(stream), C.byref(camera), it, int(y0), int(y1), int(self.maxDepth)))
        self.m_steps += 1

    def resolve(self, stream=None):
        check(self.lib.hako_pt_resolve(self.h, _stream(stream)))

    def toImageAsync(self, stream=None):
        """resolve + D2H -> uint8 [h, w, 4] (returns after the copy has landed; the reference's caller syncs the stream)."""
        out = np.empty((self.m_height, self.m_width, 4), np.uint8)
        check(self.lib.hako_pt_read_rgba8(self.h, _stream(stream), _ptr(out)))
        return out

    # ---- beyond the reference surface: parity / multi-GPU plumbing ----
    def accumulators(self, stream=None):
        out = np.empty((self.m_height, self.m_width, 4), np.float32)
        check(self.lib.hako_pt_read_accum(self.h, _stream(stream), _ptr(out)))
        return out

    def accum_device(self):
        p, b = C.c_void_p(), C.c_uint64()
        check(self.lib.hako_pt_accum_device(self.h, C.byref(p), C.byref(b)))
        return p.value or 0, b.value

    def ray_count(self, stream=None, reset=False):
        v = C.c_uint64()
        check(self.lib.hako_pt_ray_count(self.h, _stream(stream), C.byref(v), int(reset)))
        return int(v.value)

    def trace_steps(self, stream=None, reset=False):
        """node visits of the secondary rays traced by the wavefront's trace waves since the last reset"""
        v = C.c_uint64()
        check(self.lib.hako_pt_trace_steps(self.h, _stream(stream), C.byref(v), int(reset)))
        return int(v.value)

    def pmj_table(self, stream=None):
        out = np.empty(128 * 4096 * 2, np.float32)
        check(self.lib.hako_pt_pmj_table(self.h, _stream(stream), _ptr(out)))
        return out

    def hdri_sats(self, w, h, stream=None):
        out = np.empty(7 * w * h, np.uint32)
        check(self.lib.hako_pt_hdri_sats(self.h, _stream(stream), _ptr(out)))
        return out
