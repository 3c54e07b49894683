"""Camera setup: CameraPinhole::initFromPerspective (renderCommon.hpp:21-35) behind pr::GetCameraMatrix
(libs/prlib/src/pr.cpp:1099-1191).  Host-side math done in the C++ library so float rounding matches the reference."""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import Camera, check


def _f3(v):
    return np.ascontiguousarray(v, np.float32)


def camera_lookat(eye, center, up=(0.0, 1.0, 0.0), fovy=None, width=1920, height=1080, zNear=0.01, zFar=1000.0, focus=1.0, lensR=0.0):
    """voxRT.cpp:69-72 default camera: origin (4,4,4) -> (0,0,0), fovy = glm::radians(45)."""
    if fovy is None:
        fovy = float(np.float32(np.radians(np.float32(45.0))))
    cam = Camera()
    e, c, u = _f3(eye), _f3(center), _f3(up)
    check(_abi.load().hako_camera_lookat(e.ctypes.data, c.ctypes.data, u.ctypes.data, fovy, int(width), int(height), zNear, zFar, focus, lensR, C.byref(cam)))
    return cam


def camera_from_matrices(view, proj, focus=1.0, lensR=0.0):
    """view / proj: 4x4 column-major (glm::value_ptr order) float32"""
    cam = Camera()
    v, p = _f3(view).reshape(16), _f3(proj).reshape(16)
    check(_abi.load().hako_camera_from_matrices(v.ctypes.data, p.ctypes.data, focus, lensR, C.byref(cam)))
    return cam
