"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE -- never imported by the product package).

Loads either ``oracle/liboracle.so`` (our restatement, kind="port") or
``oracle/_ref/libhako_ref.so`` (the reference's own headers, kind="reference"); both export the ABI of
``oracle/hako_oracle.h``.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
FLT_MAX = np.float32(3.402823466e38)

u64p = np.ctypeslib.ndpointer(np.uint64, flags="C")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")

NODE68 = np.dtype([("mask", np.uint8), ("_pad", np.uint8, (3,)), ("children", np.uint32, (8,)), ("psum", np.uint32, (8,))])
assert NODE68.itemsize == 68


class Camera(C.Structure):
    """== reference CameraPinhole (renderCommon.hpp:77-83), 60 bytes."""
    _fields_ = [("o", C.c_float * 3), ("front", C.c_float * 3), ("up", C.c_float * 3), ("right", C.c_float * 3),
                ("tanHthetaY", C.c_float), ("lensR", C.c_float), ("focus", C.c_float)]


def build(quiet=True):
    """Compile liboracle.so (and _ref/libhako_ref.so when /root/reference is present)."""
    subprocess.run(["make", "-C", HERE] + (["-s"] if quiet else []), check=True)


def ref_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libhako_ref.so"))


class Oracle:
    def __init__(self, kind="port"):
        path = os.path.join(HERE, "liboracle.so") if kind == "port" else os.path.join(HERE, "_ref", "libhako_ref.so")
        if kind == "port" and not os.path.exists(path):
            build()
        self.kind = kind
        L = self.lib = C.CDLL(path)
        L.orc_impl_name.restype = C.c_char_p
        L.orc_morton_encode.restype = C.c_uint64
        L.orc_morton_encode.argtypes = [C.c_uint32] * 3
        L.orc_morton_encode_naive.restype = C.c_uint64
        L.orc_morton_encode_naive.argtypes = [C.c_uint32] * 3
        L.orc_morton_decode.argtypes = [C.c_uint64] + [C.POINTER(C.c_uint32)] * 3
        L.orc_sort_bits_morton.argtypes = [C.c_uint32]
        L.orc_voxelize.restype = C.c_uint64
        L.orc_voxelize.argtypes = [f32p, f32p, f32p, C.c_uint32, f32p, C.c_float, C.c_int, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_merge_voxels.restype = C.c_uint64
        L.orc_merge_voxels.argtypes = [u64p, u8p, u8p, C.c_uint64]
        for f in (L.orc_octree_build_dag, L.orc_octree_build_naive):
            f.restype = C.c_void_p
            f.argtypes = [u64p, C.c_uint64, f32p, C.c_float, C.c_int]
        L.orc_octree_free.argtypes = [C.c_void_p]
        L.orc_octree_num_nodes.restype = C.c_uint32
        L.orc_octree_num_nodes.argtypes = [C.c_void_p]
        L.orc_octree_bounds.argtypes = [C.c_void_p, f32p, f32p]
        L.orc_octree_copy_nodes.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_intersect_batch.argtypes = [C.c_void_p, f32p, f32p, C.c_uint64, C.c_int, f32p, i32p, u32p, C.c_void_p]
        L.orc_camera_lookat.argtypes = [f32p, f32p, f32p, C.c_float, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float,
                                        C.c_float, C.POINTER(Camera)]
        L.orc_camera_shoot.argtypes = [C.POINTER(Camera), C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, f32p, f32p]
        L.orc_camera_shoot_thin_lens.argtypes = [C.POINTER(Camera), C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int,
                                                 C.c_float, C.c_float, f32p, f32p]
        L.orc_render_primary.argtypes = [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         f32p, i32p, u32p, C.c_void_p]
        L.orc_fnv1a64.restype = C.c_uint64
        L.orc_fnv1a64.argtypes = [C.c_void_p, C.c_uint64]
        self._bind_pt()

    def _bind_pt(self):
        L = self.lib
        if not hasattr(L, "orc_murmur_combine"):
            return
        L.orc_murmur_combine.restype = C.c_uint32
        L.orc_murmur_combine.argtypes = [C.c_uint32, u32p, C.c_int]
        L.orc_pmj_table.argtypes = [f32p]
        L.orc_pmj_sample2d.argtypes = [f32p, C.c_uint32, C.c_uint32, C.c_uint32, f32p]
        L.orc_pcg32_stream.argtypes = [C.c_uint64, C.c_uint64, C.c_int, u32p]
        L.orc_nested_uniform_scramble.restype = C.c_uint32
        L.orc_nested_uniform_scramble.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_scramble_f32.restype = C.c_float
        L.orc_scramble_f32.argtypes = [C.c_float, C.c_uint32]
        L.orc_hdri_importance_sample.argtypes = [f32p, u32p, C.c_int, C.c_int, C.c_float, f32p, C.c_int,
                                                 C.c_float, C.c_float, C.c_float, C.c_float, f32p, f32p, f32p, u32p]
        L.orc_sample_lambertian.argtypes = [C.c_float, C.c_float, f32p, f32p]
        if not hasattr(L, "orc_pt_render"):
            return  # the reference library: device-only pieces are not available on the host
        L.orc_hdri_build_sats.argtypes = [f32p, C.c_int, C.c_int, u32p]
        L.orc_hdri_sample_nearest.argtypes = [f32p, C.c_int, C.c_int, C.c_float, f32p, f32p]
        L.orc_pt_render.argtypes = [C.c_void_p, u8p, C.c_uint32, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                    C.c_void_p, C.c_int, C.c_int, f32p, C.c_int, f32p, C.c_void_p, C.c_int]
        L.orc_pt_resolve.argtypes = [f32p, C.c_uint64, u8p]

    # ---- sampler / environment (path-tracing half) ----
    def murmur(self, seed, values):
        v = np.ascontiguousarray(values, np.uint32).reshape(-1)
        return int(self.lib.orc_murmur_combine(seed, v, v.size))

    def pcg32(self, seed, stream, n):
        out = np.zeros(n, np.uint32)
        self.lib.orc_pcg32_stream(seed, stream, n, out)
        return out

    def pmj_table(self):
        out = np.zeros(128 * 4096 * 2, np.float32)
        self.lib.orc_pmj_table(out)
        return out

    def pmj_sample2d(self, table, sample_idx, dim, stream):
        xy = np.zeros(2, np.float32)
        self.lib.orc_pmj_sample2d(table, sample_idx, dim, stream, xy)
        return xy

    def hdri_build_sats(self, rgba, w, h):
        rgba = np.ascontiguousarray(rgba, np.float32).reshape(-1)
        sats = np.zeros(7 * w * h, np.uint32)
        self.lib.orc_hdri_build_sats(rgba, w, h, sats)
        return sats

    def hdri_importance_sample(self, rgba, sats, w, h, scale, n, axis_aligned, u):
        d = np.zeros(3, np.float32)
        l = np.zeros(3, np.float32)
        p = np.zeros(1, np.float32)
        xy = np.zeros(2, np.uint32)
        self.lib.orc_hdri_importance_sample(np.ascontiguousarray(rgba, np.float32).reshape(-1), sats, w, h, scale,
                                            np.asarray(n, np.float32), int(axis_aligned), u[0], u[1], u[2], u[3], d, l, p, xy)
        return d, l, float(p[0]), xy

    def hdri_sample_nearest(self, rgba, w, h, scale, d):
        out = np.zeros(3, np.float32)
        self.lib.orc_hdri_sample_nearest(np.ascontiguousarray(rgba, np.float32).reshape(-1), w, h, scale,
                                         np.asarray(d, np.float32), out)
        return out

    def sample_lambertian(self, a, b, ng):
        out = np.zeros(3, np.float32)
        self.lib.orc_sample_lambertian(a, b, np.asarray(ng, np.float32), out)
        return out

    def pt_render(self, tree, colors, emissions, has_emission, cam, w, h, iteration, pmj, accum, hdri=None, sats=None,
                  hdri_wh=(0, 0), hdri_scale=1.75, primary=None, primary_wh=(0, 0), max_depth=8, y0=0, y1=None, threads=0):
        """One renderPT step (16 spp) accumulated into ``accum`` (float32 [h*w*4]); returns rays traced."""
        attrs = np.ascontiguousarray(np.concatenate([np.asarray(colors, np.uint8).reshape(-1, 4),
                                                     np.asarray(emissions, np.uint8).reshape(-1, 4)], axis=1)).reshape(-1)
        st = np.zeros(1, np.uint64)
        threads = threads or (os.cpu_count() or 1)
        hp = np.ascontiguousarray(hdri, np.float32).reshape(-1) if hdri is not None else None
        pp = np.ascontiguousarray(primary, np.float32).reshape(-1) if primary is not None else None
        self.lib.orc_pt_render(tree.h, attrs, int(has_emission), C.byref(cam), w, h, y0, h if y1 is None else y1, iteration,
                               hp.ctypes.data if hp is not None else None, sats.ctypes.data if sats is not None else None,
                               hdri_wh[0], hdri_wh[1], hdri_scale, pp.ctypes.data if pp is not None else None,
                               primary_wh[0], primary_wh[1], pmj, max_depth, accum, st.ctypes.data, threads)
        return int(st[0])

    def pt_resolve(self, accum):
        accum = np.ascontiguousarray(accum, np.float32).reshape(-1)
        out = np.zeros(accum.size, np.uint8)
        self.lib.orc_pt_resolve(accum, accum.size // 4, out)
        return out

    # ---- morton ----
    def morton_encode(self, x, y, z):
        return int(self.lib.orc_morton_encode(x, y, z))

    def morton_encode_naive(self, x, y, z):
        return int(self.lib.orc_morton_encode_naive(x, y, z))

    def morton_decode(self, m):
        x, y, z = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.lib.orc_morton_decode(m, C.byref(x), C.byref(y), C.byref(z))
        return x.value, y.value, z.value

    # ---- voxelize ----
    def voxelize(self, pos, col, emi, origin, dps, grid_res, six=True, attributes=True):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1)
        ntri = pos.size // 9
        col = np.ascontiguousarray(col, np.float32).reshape(-1) if col is not None else np.ones_like(pos)
        emi = np.ascontiguousarray(emi, np.float32).reshape(-1) if emi is not None else np.zeros_like(pos)
        origin = np.ascontiguousarray(origin, np.float32)
        n = self.lib.orc_voxelize(pos, col, emi, ntri, origin, dps, grid_res, int(six), None, None, None, 0)
        keys = np.empty(n, np.uint64)
        colors = np.empty((n, 4), np.uint8)
        emissions = np.empty((n, 4), np.uint8)
        n2 = self.lib.orc_voxelize(pos, col, emi, ntri, origin, dps, grid_res, int(six),
                                   keys.ctypes.data, colors.ctypes.data if attributes else None,
                                   emissions.ctypes.data if attributes else None, n)
        assert n2 == n
        return keys, colors, emissions

    def merge_voxels(self, keys, colors, emissions):
        keys = np.array(keys, np.uint64)
        colors = np.array(colors, np.uint8).reshape(-1, 4)
        emissions = np.array(emissions, np.uint8).reshape(-1, 4)
        u = self.lib.orc_merge_voxels(keys, colors.reshape(-1), emissions.reshape(-1), keys.size)
        return keys[:u].copy(), colors[:u].copy(), emissions[:u].copy()

    # ---- octree ----
    def build_octree(self, mortons, origin, dps, grid_res, dag=True):
        mortons = np.ascontiguousarray(mortons, np.uint64)
        origin = np.ascontiguousarray(origin, np.float32)
        f = self.lib.orc_octree_build_dag if dag else self.lib.orc_octree_build_naive
        return OracleOctree(self, f(mortons, mortons.size, origin, dps, grid_res))

    def camera_lookat(self, eye, center, up=(0, 1, 0), fovy=None, width=1920, height=1080, znear=0.01, zfar=1000.0,
                      focus=1.0, lens_r=0.0):
        cam = Camera()
        if fovy is None:
            fovy = float(np.float32(np.radians(np.float32(45.0))))  # glm::radians(45.0f)
        self.lib.orc_camera_lookat(np.asarray(eye, np.float32), np.asarray(center, np.float32), np.asarray(up, np.float32),
                                   fovy, width, height, znear, zfar, focus, lens_r, C.byref(cam))
        return cam

    def camera_shoot(self, cam, x, y, ox, oy, w, h):
        ro = np.zeros(3, np.float32)
        rd = np.zeros(3, np.float32)
        self.lib.orc_camera_shoot(C.byref(cam), x, y, ox, oy, w, h, ro, rd)
        return ro, rd

    def camera_shoot_thin_lens(self, cam, x, y, ox, oy, w, h, u0, u1):
        ro = np.zeros(3, np.float32)
        rd = np.zeros(3, np.float32)
        self.lib.orc_camera_shoot_thin_lens(C.byref(cam), x, y, ox, oy, w, h, u0, u1, ro, rd)
        return ro, rd

    def fnv1a64(self, arr):
        arr = np.ascontiguousarray(arr)
        return int(self.lib.orc_fnv1a64(arr.ctypes.data, arr.nbytes))


class OracleOctree:
    def __init__(self, orc, handle):
        self.orc = orc
        self.h = handle

    def __del__(self):
        if self.h:
            self.orc.lib.orc_octree_free(self.h)
            self.h = None

    @property
    def num_nodes(self):
        return int(self.orc.lib.orc_octree_num_nodes(self.h))

    def bounds(self):
        lo = np.zeros(3, np.float32)
        up = np.zeros(3, np.float32)
        self.orc.lib.orc_octree_bounds(self.h, lo, up)
        return lo, up

    def nodes(self):
        out = np.zeros(self.num_nodes, NODE68)
        self.orc.lib.orc_octree_copy_nodes(self.h, out.ctypes.data)
        return out

    def intersect(self, ro, rd, shadow=False, stats=False):
        ro = np.ascontiguousarray(ro, np.float32).reshape(-1, 3)
        rd = np.ascontiguousarray(rd, np.float32).reshape(-1, 3)
        n = ro.shape[0]
        t = np.full(n, FLT_MAX, np.float32)
        nmajor = np.full(n, -1, np.int32)
        vindex = np.full(n, 0xFFFFFFFF, np.uint32)
        st = np.zeros(4, np.uint64)
        self.orc.lib.orc_intersect_batch(self.h, ro.reshape(-1), rd.reshape(-1), n, int(shadow), t, nmajor, vindex,
                                         st.ctypes.data if stats else None)
        return (t, nmajor, vindex, st) if stats else (t, nmajor, vindex)

    def render_primary(self, cam, w, h, y0=0, y1=None, threads=1, stats=False):
        y1 = h if y1 is None else y1
        t = np.full(w * h, FLT_MAX, np.float32)
        nmajor = np.full(w * h, -1, np.int32)
        vindex = np.full(w * h, 0xFFFFFFFF, np.uint32)
        st = np.zeros(4, np.uint64)
        self.orc.lib.orc_render_primary(self.h, C.byref(cam), w, h, y0, y1, threads, t, nmajor, vindex,
                                        st.ctypes.data if stats else None)
        return (t, nmajor, vindex, st) if stats else (t, nmajor, vindex)
