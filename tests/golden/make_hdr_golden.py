#!/usr/bin/env python
"""Golden vectors for the Radiance .hdr decoder, produced by the decoder the REFERENCE uses: stbi_loadf( .., 4 ) of the stb_image.h vendored
in /root/reference/libs/prlib/libs/stb (libs/prlib/src/prp.cpp:353-366), compiled where it lies.

Run in the build container only (needs /root/reference):  python tests/golden/make_hdr_golden.py

Output: hdr_golden.npz -- for each case the bytes of a small synthetic .hdr file (new-style RLE and flat scanlines, black pixels, long runs,
extreme exponents; written by this script, not reference data) and the float32 RGBA image stbi_loadf returns for it.  The test
tests/test_io.py::test_hdr_matches_stbi_golden decodes the same bytes with hako_load_hdr and compares bit for bit, on any machine."""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from test_io import _rle_scanline  # noqa: E402  (the test module's own RLE writer)

STB = "/root/reference/libs/prlib/libs/stb"


def hdr_bytes(rgbe, rle):
    h, w = rgbe.shape[:2]
    body = b"".join(_rle_scanline(rgbe[y]) for y in range(h)) if rle else rgbe.tobytes()
    return b"#?RADIANCE\nSOFTWARE=hako-tests\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w) + body


def main():
    tmp = tempfile.mkdtemp()
    src = os.path.join(tmp, "stb_drv.c")
    open(src, "w").write('#define STB_IMAGE_IMPLEMENTATION\n#define STBI_ONLY_HDR\n#include "stb_image.h"\n'
                         'float* drv_load(const char* f, int* w, int* h) { int c; return stbi_loadf(f, w, h, &c, 4); }\n')
    so = os.path.join(tmp, "libstb_drv.so")
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-I" + STB, src, "-o", so, "-lm"], check=True)
    lib = C.CDLL(so)
    lib.drv_load.restype = C.POINTER(C.c_float)
    out = {}
    for i, (w, h, rle) in enumerate([(64, 32, True), (40, 7, False), (9, 5, True), (300, 11, True), (128, 64, True)]):
        rng = np.random.default_rng(1000 + i)
        rgbe = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        rgbe[::3, ::5, 3] = 0                     # black pixels (exponent 0)
        rgbe[1::4, :, 0] = 77                     # long runs
        rgbe[:, w // 2:, 2] = 9
        rgbe[0, :4, 3] = (1, 2, 254, 255)         # extreme exponents
        if not rle:
            rgbe[0, 0, :3] = (1, 1, 1)            # a flat file must not start with the RLE marker 2 2
        if w < 8:
            rle = False                           # stb reads scanlines shorter than 8 pixels as flat
        data = hdr_bytes(rgbe, rle)
        path = os.path.join(tmp, "c%d.hdr" % i)
        open(path, "wb").write(data)
        ww, hh = C.c_int(), C.c_int()
        ptr = lib.drv_load(path.encode(), C.byref(ww), C.byref(hh))
        assert ptr and (ww.value, hh.value) == (w, h)
        out["file%d" % i] = np.frombuffer(data, np.uint8)
        out["image%d" % i] = np.ctypeslib.as_array(ptr, shape=(h, w, 4)).copy()
    np.savez_compressed(os.path.join(HERE, "hdr_golden.npz"), **out)
    print("wrote hdr_golden.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
