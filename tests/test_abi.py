"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every function include/hako.h
declares, the ctypes table matches the header, and calls fail loudly (no CPU fallback) when no device is usable."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "hako.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hako_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from massivevoxelraytracing_b200 import _abi
    lib = _abi.load()
    names = _header_functions()
    assert len(names) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", _abi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (hako_[a-z0-9_]+)", out))
    assert set(names) <= exported, sorted(set(names) - exported)
    assert set(names) == set(_abi.SIGNATURES), (sorted(set(names) - set(_abi.SIGNATURES)), sorted(set(_abi.SIGNATURES) - set(names)))
    for n in names:
        assert hasattr(lib, n)
    assert lib.hako_version() >= 100


def test_struct_layouts_match_reference():
    from massivevoxelraytracing_b200 import _abi, intersector
    assert C.sizeof(_abi.Camera) == 60                 # CameraPinhole, renderCommon.hpp:77-83
    assert intersector.NODE68.itemsize == 68           # OctreeNode, voxCommon.hpp:133-138
    assert intersector.VOXEL_ATTRIB.itemsize == 8      # VoxelAttirb, voxCommon.hpp:121-125


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import massivevoxelraytracing_b200 as hk
    with pytest.raises(hk.HakoError) as e:
        hk.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    """the oracle is test infrastructure: nothing under the package (or include/) may reference it"""
    pkg = os.path.join(ROOT, "massivevoxelraytracing_b200")
    for base, _, files in os.walk(pkg):
        if "build" in base.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", "Makefile")):
                text = open(os.path.join(base, f)).read()
                assert "liboracle" not in text and "hako_oracle" not in text and "from oracle" not in text and "import oracle" not in text, f


def test_camera_matches_oracle(oracle):
    import numpy as np
    import massivevoxelraytracing_b200 as hk
    for eye, ctr, w, h, fovy in (((4, 4, 4), (0, 0, 0), 1920, 1080, None), ((1, 2, 3), (0.5, 0, -1), 640, 480, 0.6), ((-3, 0.5, 9), (0, 1, 0), 33, 77, 1.2)):
        a = hk.camera_lookat(eye, ctr, (0, 1, 0), fovy, w, h, focus=2.5, lensR=0.1)
        b = oracle.camera_lookat(eye, ctr, (0, 1, 0), fovy, w, h, focus=2.5, lens_r=0.1)
        assert bytes(a) == bytes(b)
    # initFromPerspective from explicit matrices (column-major identity view, simple projection)
    view = np.eye(4, dtype=np.float32)
    view[3, :3] = (1, 2, 3)            # column 3 = translation (column-major: row index of numpy = column)
    proj = np.zeros((4, 4), np.float32)
    proj[1, 1] = 2.0
    cam = hk.camera_from_matrices(view, proj, 1.5, 0.25)
    assert tuple(cam.o) == (-1.0, -2.0, -3.0) and tuple(cam.front) == (-0.0, -0.0, -1.0) and cam.tanHthetaY == 0.5
    assert cam.focus == 1.5 and cam.lensR == 0.25


def build_shim_demo(out_dir):
    """compile tests/cpp/shim_demo.cpp against include/*.hpp and link it with libhako_b200.so"""
    exe = os.path.join(str(out_dir), "shim_demo")
    pkg = os.path.join(ROOT, "massivevoxelraytracing_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "shim_demo.cpp"),
                    "-o", exe, "-pthread", "-L" + pkg, "-lhako_b200", "-Wl,-rpath," + pkg], check=True)
    return exe


def test_cpp_shims_compile_and_fail_loudly_without_gpu(tmp_path):
    """the C++ drop-in headers (IntersectorOctreeGPU.hpp / PathTracer.hpp) build with plain g++; with no device the
    demo exits through hako::Error (no CPU fallback)"""
    import torch
    exe = build_shim_demo(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present (the demo itself runs in tests/test_gpu_shim.py)")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr


def test_reference_gpu_launcher_loads_and_fails_loudly_without_a_driver():
    """oracle/librefgpu.so (the driver-API launcher of the reference's own kernels; test infrastructure) exports its ABI and, on a
    machine without libcuda, refuses to open with a message instead of crashing"""
    import ctypes as C
    import subprocess
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    path = os.path.join(ROOT, "oracle", "librefgpu.so")
    lib = C.CDLL(path)
    for name in ("refgpu_open", "refgpu_close", "refgpu_set_octree", "refgpu_set_attrs", "refgpu_render", "refgpu_load_hdri", "refgpu_get_sats", "refgpu_set_pmj",
                 "refgpu_pt_resize", "refgpu_pt_clear", "refgpu_pt_step", "refgpu_pt_read_accum", "refgpu_pt_write_accum", "refgpu_pt_resolve", "refgpu_set_scales",
                 "refgpu_last_error"):
        assert hasattr(lib, name), name
    import torch
    if not torch.cuda.is_available():
        lib.refgpu_last_error.restype = C.c_char_p
        h = C.c_void_p()
        rc = lib.refgpu_open(b"/nonexistent.cubin", 0, C.byref(h))
        assert rc != 0 and not h.value and lib.refgpu_last_error()
