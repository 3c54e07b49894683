"""The C++ host side: a voxRTGPU/voxPTGPU-style caller written against include/IntersectorOctreeGPU.hpp and
include/PathTracer.hpp runs end to end on the GPU (build, primary rays, IntersectorOctree rebuild from Morton keys,
path tracing, resolve, error behaviour)."""
import subprocess

import pytest

from test_abi import build_shim_demo

pytestmark = pytest.mark.gpu


def test_cpp_shim_demo_runs(tmp_path):
    exe = build_shim_demo(tmp_path)
    r = subprocess.run([exe, str(tmp_path / "out.ppm")], capture_output=True, text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "shim demo ok" in r.stdout
    assert (tmp_path / "out.ppm").stat().st_size > 256 * 192 * 3
