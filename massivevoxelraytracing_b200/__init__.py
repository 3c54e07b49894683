"""massivevoxelraytracing_b200 -- B200-native (sm_100a, ahead-of-time CUDA) drop-in for the GPU hot path of
Ushio/MassiveVoxelRayTracing ("Hako"): voxelize -> Morton sort/unique -> sparse-voxel-octree (DAG) build ->
Revelles parametric ray traversal -> wavefront path tracing.

The product is `libhako_b200.so` (C ABI in include/hako.h).  This package is the thin host-side mirror of the
reference's operator surface (IntersectorOctreeGPU / IntersectorOctree / PathTracer) over that ABI.
Importing it loads the library and fails loudly when it is missing -- there is no CPU fallback.
"""
from . import _abi
from ._abi import Camera, HakoError

_abi.load()

from .intersector import (FLT_MAX, NODE68, VOXEL_ATTRIB, Context, IntersectorOctree, IntersectorOctreeGPU)  # noqa: E402
from .camera import camera_from_matrices, camera_lookat  # noqa: E402
from .pathtracer import PathTracer  # noqa: E402
from . import io  # noqa: E402

__all__ = ["Camera", "HakoError", "Context", "IntersectorOctree", "IntersectorOctreeGPU", "PathTracer", "FLT_MAX", "NODE68", "VOXEL_ATTRIB",
           "camera_lookat", "camera_from_matrices"]
