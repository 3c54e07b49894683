"""ctypes binding of libhako_b200.so (the C ABI declared in include/hako.h).

This module is plumbing only: it loads the ahead-of-time compiled sm_100a library that sits next to this file
and attaches argument types.  There is NO fallback: if the library is missing, or a symbol of include/hako.h is
not exported, importing fails loudly (the product path never routes through oracle/ or any CPU code).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# HAKO_LIB_PATH: measurement tooling only (a variant build of the same sources, e.g. tools/ A/B probes)
LIB_PATH = os.environ.get("HAKO_LIB_PATH") or os.path.join(HERE, "libhako_b200.so")

HAKO_OK, HAKO_E_INVALID, HAKO_E_CUDA, HAKO_E_LIMIT, HAKO_E_STATE = 0, -1, -2, -3, -4


class HakoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("hako error %d: %s" % (code, msg))
        self.code = code


class Camera(C.Structure):
    """== CameraPinhole members (renderCommon.hpp:77-83), 60 bytes."""
    _fields_ = [("o", C.c_float * 3), ("front", C.c_float * 3), ("up", C.c_float * 3), ("right", C.c_float * 3),
                ("tanHthetaY", C.c_float), ("lensR", C.c_float), ("focus", C.c_float)]


class OctreeInfo(C.Structure):
    _fields_ = [("numberOfNodes", C.c_uint32), ("numberOfVoxels", C.c_uint32), ("lower", C.c_float * 3), ("upper", C.c_float * 3),
                ("dps", C.c_float), ("emissionScale", C.c_float), ("hasEmission", C.c_uint32), ("gridRes", C.c_int32),
                ("deviceBytes", C.c_uint64), ("wideIndex", C.c_uint32), ("reserved", C.c_uint32)]


class BuildStats(C.Structure):
    _fields_ = [("nTriangles", C.c_uint64), ("dumpedVoxels", C.c_uint64), ("uniqueVoxels", C.c_uint64),
                ("numberOfNodes", C.c_uint32), ("sortPasses", C.c_uint32),
                ("ms_h2d", C.c_float), ("ms_voxelize", C.c_float), ("ms_sort", C.c_float), ("ms_unique", C.c_float),
                ("ms_octree", C.c_float), ("ms_total", C.c_float), ("hostSyncs", C.c_uint32), ("kernelLaunches", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


vp = C.c_void_p
pp = C.POINTER(C.c_void_p)
u64 = C.c_uint64
u32 = C.c_uint32
i32 = C.c_int
f32 = C.c_float
pu64 = C.POINTER(C.c_uint64)
pu32 = C.POINTER(C.c_uint32)

# name -> (restype, argtypes).  Every function include/hako.h declares is listed; tests/test_abi.py checks the
# header, this table and the exported symbols against each other.
SIGNATURES = {
    "hako_last_error": (C.c_char_p, []),
    "hako_version": (i32, []),
    "hako_ctx_create": (i32, [i32, pp]),
    "hako_ctx_destroy": (None, [vp]),
    "hako_ctx_synchronize": (i32, [vp, vp]),
    "hako_ctx_build_stats": (i32, [vp, C.POINTER(BuildStats)]),
    "hako_ctx_launch_count": (u64, [vp]),
    "hako_ctx_l2_probe": (i32, [vp, vp, u64, i32, pu64, C.POINTER(C.c_float)]),
    "hako_ctx_reserve": (i32, [vp, vp, u64]),
    "hako_octree_build": (i32, [vp, vp, vp, vp, vp, u32, vp, f32, i32, pp]),
    "hako_octree_build_device": (i32, [vp, vp, vp, vp, vp, u32, vp, f32, i32, pp]),
    "hako_octree_build_from_morton": (i32, [vp, vp, vp, vp, u64, vp, f32, i32, i32, pp]),
    "hako_octree_import": (i32, [vp, vp, vp, u32, vp, u32, vp, vp, f32, pp]),
    "hako_octree_info_get": (i32, [vp, C.POINTER(OctreeInfo)]),
    "hako_octree_export": (i32, [vp, vp, vp, vp, vp]),
    "hako_octree_destroy": (None, [vp]),
    "hako_octree_destroy_async": (None, [vp, vp]),
    "hako_octree_device_arrays": (i32, [vp, pp, pu64, pp, pu64, pp, pu64]),
    "hako_octree_create_replica": (i32, [vp, C.POINTER(OctreeInfo), pp]),
    "hako_voxelize": (i32, [vp, vp, vp, vp, vp, u32, vp, f32, i32, i32, vp, vp, u64, pu64]),
    "hako_voxelize_sorted": (i32, [vp, vp, vp, vp, vp, u32, vp, f32, i32, pp]),
    "hako_pairs_info": (i32, [vp, pu64, pp, pp]),
    "hako_pairs_destroy": (None, [vp]),
    "hako_sort_unique_device": (i32, [vp, vp, vp, vp, u64, i32, pu64, pu32]),
    "hako_octree_build_from_morton_device": (i32, [vp, vp, vp, vp, u64, vp, f32, i32, i32, i32, pp]),
    "hako_octree_create_for_voxels": (i32, [vp, vp, u64, vp, f32, i32, pp, pp, pp]),
    "hako_octree_build_nodes": (i32, [vp, vp, i32, i32]),
    "hako_sort_pairs_device": (i32, [vp, vp, vp, vp, u64, i32, i32]),
    "hako_sort_pairs": (i32, [vp, vp, vp, vp, u64, i32, i32]),
    "hako_unique_merge": (i32, [vp, vp, vp, vp, u64, pu64, pu32]),
    "hako_unique_merge_noalloc": (i32, [vp, vp, vp, vp, u64, vp, vp, vp]),
    "hako_intersect_batch": (i32, [vp, vp, vp, vp, u64, i32, vp, vp, vp]),
    "hako_intersect_batch_device": (i32, [vp, vp, vp, vp, u64, i32, vp, vp, vp]),
    "hako_render_primary": (i32, [vp, vp, C.POINTER(Camera), i32, i32, i32, i32, i32, vp, vp, vp, vp]),
    "hako_render_primary_device": (i32, [vp, vp, C.POINTER(Camera), i32, i32, i32, i32, i32, vp, vp, vp, vp]),
    "hako_render_primary_views": (i32, [vp, vp, C.POINTER(Camera), i32, i32, i32, i32, vp, vp, vp, vp]),
    "hako_camera_from_matrices": (i32, [vp, vp, f32, f32, C.POINTER(Camera)]),
    "hako_camera_lookat": (i32, [vp, vp, vp, f32, i32, i32, f32, f32, f32, f32, C.POINTER(Camera)]),
    "hako_pt_create": (i32, [vp, vp, pp]),
    "hako_pt_destroy": (None, [vp]),
    "hako_pt_resize": (i32, [vp, vp, i32, i32]),
    "hako_pt_clear": (i32, [vp, vp]),
    "hako_pt_load_hdri": (i32, [vp, vp, vp, i32, i32, vp, i32, i32]),
    "hako_pt_step": (i32, [vp, vp, vp, C.POINTER(Camera), i32, i32, i32, i32]),
    "hako_pt_resolve": (i32, [vp, vp]),
    "hako_pt_read_rgba8": (i32, [vp, vp, vp]),
    "hako_pt_read_rgba8_async": (i32, [vp, vp, vp]),
    "hako_pt_read_accum": (i32, [vp, vp, vp]),
    "hako_pt_accum_device": (i32, [vp, pp, pu64]),
    "hako_pt_ray_count": (i32, [vp, vp, pu64, i32]),
    "hako_pt_trace_steps": (i32, [vp, vp, pu64, i32]),
    "hako_pt_pmj_table": (i32, [vp, vp, vp]),
    "hako_pt_hdri_sats": (i32, [vp, vp, vp]),
    "hako_pt_rgba8_device": (i32, [vp, pp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "hako_pt_context": (vp, [vp]),
    "hako_mesh_load": (i32, [C.c_char_p, pp]),
    "hako_mesh_destroy": (None, [vp]),
    "hako_mesh_info": (i32, [vp, pu64, pu64, C.POINTER(C.c_int)]),
    "hako_mesh_flatten": (i32, [vp, vp, vp, vp]),
    "hako_bounds_device": (i32, [vp, vp, vp, u64, vp, vp]),
    "hako_scene_params": (i32, [vp, vp, i32, vp, C.POINTER(C.c_float)]),
    "hako_image_load_hdr": (i32, [C.c_char_p, pp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "hako_image_free": (None, [vp]),
    "hako_image_save_png": (i32, [C.c_char_p, vp, i32, i32]),
    "hako_seq_create": (i32, [vp, i32, i32, pp]),
    "hako_seq_render_frame": (i32, [vp, vp, vp, vp, u32, vp, f32, i32, C.POINTER(Camera), i32, i32, C.c_char_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "hako_seq_wait": (i32, [vp]),
    "hako_seq_destroy": (None, [vp]),
    "hako_voxel_mesh_save_ply": (i32, [C.c_char_p, vp, u64, vp, f32]),
    "hako_dist_unique_id": (i32, [vp]),
    "hako_dist_init": (i32, [vp, vp, i32, i32, pp]),
    "hako_dist_destroy": (None, [vp]),
    "hako_dist_info": (i32, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "hako_dist_barrier": (i32, [vp, vp]),
    "hako_dist_max_f32": (i32, [vp, vp, C.POINTER(C.c_float)]),
    "hako_dist_broadcast_octree": (i32, [vp, vp, i32, vp, pp]),
    "hako_dist_reduce_accum": (i32, [vp, vp, vp, i32]),
    "hako_dist_reduce_f32": (i32, [vp, vp, vp, u64, i32]),
    "hako_dist_broadcast_bytes": (i32, [vp, vp, vp, u64, i32]),
    "hako_dist_build_sharded": (i32, [vp, vp, vp, vp, vp, u32, vp, f32, i32, pp, C.POINTER(C.c_float)]),
}

_lib = None


def load():
    """dlopen libhako_b200.so and bind every entry point.  No GPU is needed to load (only to call)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `make -C massivevoxelraytracing_b200/csrc` "
                          "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    missing = [n for n in SIGNATURES if not hasattr(lib, n)]
    if missing:
        raise ImportError("libhako_b200.so does not export: %s" % ", ".join(missing))
    for name, (res, args) in SIGNATURES.items():
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
    _lib = lib
    return lib


def check(code):
    if code != 0:
        raise HakoError(code, load().hako_last_error().decode("utf-8", "replace"))
