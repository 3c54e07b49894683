"""Host-side mirror of the reference's operator surface for the massive-voxel hot path, over the C ABI.

    IntersectorOctreeGPU   IntersectorOctreeGPU.hpp:21-275   build / cleanUp / intersect (+ the `render` kernel, voxKernel.cu:437-483)
    IntersectorOctree      IntersectorOctree.hpp:214-266     build / buildDAGReference / intersect from a sorted unique Morton list

Method names, argument order and conventions follow the reference (caller presets t = FLT_MAX; a miss leaves the
outputs untouched; gridRes must be a power of two).  The `Shader*` argument of the reference's build() is gone -- kernels
are compiled ahead of time.  Host data are numpy arrays; methods with a `_device` suffix take torch CUDA tensors
and only enqueue work on the given stream.  All arithmetic happens in libhako_b200.so; nothing here computes.
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import BuildStats, Camera, OctreeInfo, check

FLT_MAX = np.float32(3.402823466e38)

NODE68 = np.dtype([("mask", np.uint8), ("_pad", np.uint8, (3,)), ("children", np.uint32, (8,)), ("psum", np.uint32, (8,))])
VOXEL_ATTRIB = np.dtype([("color", np.uint8, (4,)), ("emission", np.uint8, (4,))])
assert NODE68.itemsize == 68 and VOXEL_ATTRIB.itemsize == 8


def _ptr(a):
    """address of a numpy array / torch tensor / None"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, np.float32)
    return a if shape is None else a.reshape(shape)


def _stream(stream):
    if stream is None:
        return None
    return getattr(stream, "cuda_stream", stream)


class Context:
    """One per device (replaces oroCtxCreate + the nvrtc Shader, voxRTGPU.cpp:25-68)."""

    def __init__(self, device=0):
        self.lib = _abi.load()
        h = C.c_void_p()
        check(self.lib.hako_ctx_create(int(device), C.byref(h)))
        self.h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.lib.hako_ctx_destroy(self.h)
            self.h = None

    def synchronize(self, stream=None):
        check(self.lib.hako_ctx_synchronize(self.h, _stream(stream)))

    @property
    def launch_count(self):
        return int(self.lib.hako_ctx_launch_count(self.h))

    def reserve(self, nbytes, stream=None):
        """grow the context's memory pool to nbytes now instead of during the first build"""
        check(self.lib.hako_ctx_reserve(self.h, _stream(stream), int(nbytes)))

    def l2_probe(self, set_bytes=0, reps=20, stream=None):
        """(L2 size in bytes, measured L2 read bandwidth in GB/s over a resident set) -- the roof the traversal is held against"""
        l2, gbs = C.c_uint64(), C.c_float()
        check(self.lib.hako_ctx_l2_probe(self.h, _stream(stream), int(set_bytes), int(reps), C.byref(l2), C.byref(gbs)))
        return int(l2.value), float(gbs.value)

    def build_stats(self):
        s = BuildStats()
        check(self.lib.hako_ctx_build_stats(self.h, C.byref(s)))
        return s.as_dict()

    # ---- build stages (BASELINE config 4 sweep / stage-level parity) ----
    def voxelize(self, vertices, vcolors, vemissions, origin, dps, gridRes, sixSeparating=True, stream=None):
        """voxCount + voxelize (voxKernel.cu:58-166): -> (keys u64[D], attribs VOXEL_ATTRIB[D]), unordered."""
        pos = _f32(vertices).reshape(-1)
        ntri = pos.size // 9
        col = _f32(vcolors).reshape(-1)
        emi = _f32(vemissions).reshape(-1)
        org = _f32(origin)
        D = C.c_uint64()
        check(self.lib.hako_voxelize(self.h, _stream(stream), _ptr(pos), _ptr(col), _ptr(emi), ntri, _ptr(org), float(dps), int(gridRes),
                                     int(bool(sixSeparating)), None, None, 0, C.byref(D)))
        keys = np.empty(D.value, np.uint64)
        attrs = np.empty(D.value, VOXEL_ATTRIB)
        D2 = C.c_uint64()
        check(self.lib.hako_voxelize(self.h, _stream(stream), _ptr(pos), _ptr(col), _ptr(emi), ntri, _ptr(org), float(dps), int(gridRes),
                                     int(bool(sixSeparating)), _ptr(keys), _ptr(attrs), D.value, C.byref(D2)))
        assert D2.value == D.value
        return keys, attrs

    def sort_pairs(self, keys, values=None, beginBit=0, endBit=64, stream=None):
        """thrs::RadixSort::sortPairs / sortKeys (tinyhipradixsort.hpp:849-944): stable LSD sort, host arrays, returns copies."""
        k = np.array(keys, np.uint64)
        v = None if values is None else np.array(values, np.uint64)
        check(self.lib.hako_sort_pairs(self.h, _stream(stream), _ptr(k), _ptr(v), k.size, int(beginBit), int(endBit)))
        return (k, v) if v is not None else k

    def sort_pairs_device(self, keys, values=None, beginBit=0, endBit=64, stream=None):
        """in place on torch CUDA int64/uint64 tensors"""
        check(self.lib.hako_sort_pairs_device(self.h, _stream(stream), _ptr(keys), _ptr(values), keys.numel(), int(beginBit), int(endBit)))

    def unique_merge(self, keys, attrs, stream=None):
        """`unique` kernel (voxKernel.cu:170-243) == mergeVoxels (voxRT.cpp:14-51) on sorted input."""
        k = np.array(keys, np.uint64)
        a = np.array(attrs, VOXEL_ATTRIB)
        U = C.c_uint64()
        he = C.c_uint32()
        check(self.lib.hako_unique_merge(self.h, _stream(stream), _ptr(k), _ptr(a), k.size, C.byref(U), C.byref(he)))
        return k[:U.value].copy(), a[:U.value].copy(), bool(he.value)


class _OctreeBase:
    def __init__(self, ctx):
        self.ctx = ctx
        self.lib = ctx.lib
        self.h = None

    # IntersectorOctreeGPU::cleanUp, IntersectorOctreeGPU.hpp:26-38
    def cleanUp(self):
        if self.h:
            self.lib.hako_octree_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.cleanUp()
        except Exception:
            pass

    def _adopt(self, handle):
        self.cleanUp()
        self.h = handle

    def info(self):
        i = OctreeInfo()
        check(self.lib.hako_octree_info_get(self.h, C.byref(i)))
        return i

    @property
    def m_numberOfNodes(self):
        return self.info().numberOfNodes

    @property
    def m_numberOfVoxels(self):
        return self.info().numberOfVoxels

    @property
    def m_lower(self):
        return np.array(self.info().lower, np.float32)

    @property
    def m_upper(self):
        return np.array(self.info().upper, np.float32)

    def hasEmission(self):
        return bool(self.info().hasEmission)

    def getMemoryConsumption(self):
        return int(self.info().deviceBytes)

    def export(self, nodes=True, mortons=False, attrs=False, stream=None):
        """download in the reference layout: OctreeNode[ ] (68 B), sorted unique Morton keys, VoxelAttirb[ ]"""
        i = self.info()
        n = np.zeros(i.numberOfNodes, NODE68) if nodes else None
        m = np.zeros(i.numberOfVoxels, np.uint64) if mortons else None
        a = np.zeros(i.numberOfVoxels, VOXEL_ATTRIB) if attrs else None
        check(self.lib.hako_octree_export(self.h, _stream(stream), _ptr(n), _ptr(m), _ptr(a)))
        return n, m, a

    def import_nodes(self, nodes68, lower, upper, dps, attrs=None, stream=None):
        """upload a tree built elsewhere (canonical layout, masks embedded) -- parity tooling"""
        nodes68 = np.ascontiguousarray(nodes68)
        assert nodes68.dtype.itemsize == 68
        attrs = None if attrs is None else np.ascontiguousarray(attrs)
        h = C.c_void_p()
        lo, up = _f32(lower), _f32(upper)  # named: a temporary would be freed (and its slot reused) before the call runs
        assert lo.size == 3 and up.size == 3
        check(self.lib.hako_octree_import(self.ctx.h, _stream(stream), _ptr(nodes68), nodes68.size, _ptr(attrs),
                                          0 if attrs is None else attrs.shape[0], _ptr(lo), _ptr(up), float(dps), C.byref(h)))
        self._adopt(h)
        return self

    # IntersectorOctreeGPU::intersect (:243-251) / IntersectorOctree::intersect (:248-257), batched
    def intersect(self, ro, rd, t=None, nMajor=None, vIndex=None, isShadowRay=False, stream=None):
        ro = _f32(ro).reshape(-1, 3)
        rd = _f32(rd).reshape(-1, 3)
        n = ro.shape[0]
        t = np.full(n, FLT_MAX, np.float32) if t is None else t
        nMajor = np.full(n, -1, np.int32) if nMajor is None else nMajor
        vIndex = np.full(n, 0xFFFFFFFF, np.uint32) if vIndex is None else vIndex
        check(self.lib.hako_intersect_batch(self.h, _stream(stream), _ptr(ro), _ptr(rd), n, int(bool(isShadowRay)), _ptr(t), _ptr(nMajor), _ptr(vIndex)))
        return t, nMajor, vIndex

    def intersect_device(self, ro, rd, t, nMajor, vIndex, isShadowRay=False, stream=None):
        check(self.lib.hako_intersect_batch_device(self.h, _stream(stream), _ptr(ro), _ptr(rd), ro.numel() // 3, int(bool(isShadowRay)),
                                                   _ptr(t), _ptr(nMajor), _ptr(vIndex)))

    # `render` kernel (voxKernel.cu:437-483) / renderLine (voxRT.cpp:307-347), rows [y0, y1)
    def render(self, camera, width, height, y0=0, y1=None, showVertexColor=False, rgba=True, ids=True, stream=None):
        y1 = height if y1 is None else y1
        rows = y1 - y0
        out_rgba = np.empty((rows, width, 4), np.uint8) if rgba else None
        t = np.empty((rows, width), np.float32) if ids else None
        nm = np.empty((rows, width), np.int32) if ids else None
        vi = np.empty((rows, width), np.uint32) if ids else None
        check(self.lib.hako_render_primary(self.h, _stream(stream), C.byref(camera), width, height, y0, y1, int(bool(showVertexColor)),
                                           _ptr(out_rgba), _ptr(t), _ptr(nm), _ptr(vi)))
        return out_rgba, t, nm, vi

    def render_views(self, cameras, width, height, showVertexColor=False, rgba=True, ids=True, stream=None, out=None):
        """a run of whole frames in one call (the orbit loop of voxRTGPU.cpp:150-208, headless); frame i+1 is traced while
        frame i is copied out.  `out` = (rgba, t, nMajor, vIndex) host arrays to fill (e.g. pinned), else fresh arrays."""
        n = len(cameras)
        cams = (Camera * max(n, 1))(*cameras)
        if out is None:
            out = (np.empty((n, height, width, 4), np.uint8) if rgba else None,
                   np.empty((n, height, width), np.float32) if ids else None,
                   np.empty((n, height, width), np.int32) if ids else None,
                   np.empty((n, height, width), np.uint32) if ids else None)
        check(self.lib.hako_render_primary_views(self.h, _stream(stream), cams, n, width, height, int(bool(showVertexColor)),
                                                 _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3])))
        return out

    def render_device(self, camera, width, height, y0=0, y1=None, showVertexColor=False, rgba=None, t=None, nMajor=None, vIndex=None, stream=None):
        y1 = height if y1 is None else y1
        check(self.lib.hako_render_primary_device(self.h, _stream(stream), C.byref(camera), width, height, y0, y1, int(bool(showVertexColor)),
                                                  _ptr(rgba), _ptr(t), _ptr(nMajor), _ptr(vIndex)))

    def device_arrays(self):
        """[(address, bytes)] of nodes / masks / attributes resident in HBM (what the NCCL broadcast moves)"""
        p = [C.c_void_p() for _ in range(3)]
        b = [C.c_uint64() for _ in range(3)]
        check(self.lib.hako_octree_device_arrays(self.h, C.byref(p[0]), C.byref(b[0]), C.byref(p[1]), C.byref(b[1]), C.byref(p[2]), C.byref(b[2])))
        return [(p[i].value or 0, b[i].value) for i in range(3)]


class IntersectorOctreeGPU(_OctreeBase):
    """IntersectorOctreeGPU.hpp:21-275."""

    def build(self, vertices, vcolors, vemissions, stream=None, origin=(0, 0, 0), dps=1.0, gridRes=256):
        """triangles (9 floats each, as trianglesFlattened voxUtil.hpp:8-65) -> voxelize -> sort -> unique -> octree"""
        pos = _f32(vertices).reshape(-1)
        col = _f32(vcolors).reshape(-1)
        emi = _f32(vemissions).reshape(-1)
        assert pos.size % 9 == 0 and col.size == pos.size and emi.size == pos.size
        h = C.c_void_p()
        org = _f32(origin)
        assert org.size == 3
        check(self.lib.hako_octree_build(self.ctx.h, _stream(stream), _ptr(pos), _ptr(col), _ptr(emi), pos.size // 9, _ptr(org), float(dps),
                                         int(gridRes), C.byref(h)))
        self._adopt(h)
        return self

    def build_device(self, vertices, vcolors, vemissions, stream=None, origin=(0, 0, 0), dps=1.0, gridRes=256):
        """same with torch CUDA float32 tensors already resident in HBM (no H2D inside)"""
        h = C.c_void_p()
        org = _f32(origin)
        assert org.size == 3
        check(self.lib.hako_octree_build_device(self.ctx.h, _stream(stream), _ptr(vertices), _ptr(vcolors), _ptr(vemissions), vertices.numel() // 9,
                                                _ptr(org), float(dps), int(gridRes), C.byref(h)))
        self._adopt(h)
        return self


class IntersectorOctree(_OctreeBase):
    """IntersectorOctree.hpp:214-266 (the CPU class's API; the work happens on the GPU)."""

    def _build(self, mortonVoxels, origin, dps, gridRes, dag, attrs=None, stream=None):
        m = np.ascontiguousarray(mortonVoxels, np.uint64)
        attrs = None if attrs is None else np.ascontiguousarray(attrs)
        h = C.c_void_p()
        org = _f32(origin)
        assert org.size == 3
        check(self.lib.hako_octree_build_from_morton(self.ctx.h, _stream(stream), _ptr(m), _ptr(attrs), m.size, _ptr(org), float(dps), int(gridRes),
                                                     int(dag), C.byref(h)))
        self._adopt(h)
        return self

    def build(self, mortonVoxels, origin, dps, gridRes, attrs=None, stream=None):
        """buildOctreeNaive (IntersectorOctree.hpp:125-212): no dedupe, nVoxelsPSum left 0"""
        return self._build(mortonVoxels, origin, dps, gridRes, 0, attrs, stream)

    def buildDAGReference(self, mortonVoxels, origin, dps, gridRes, attrs=None, stream=None):
        """buildOctreeDAGReference (IntersectorOctree.hpp:11-123): DAG with the reference's node numbering"""
        return self._build(mortonVoxels, origin, dps, gridRes, 1, attrs, stream)
