#!/usr/bin/env python3
"""tests/golden/make_golden_basic.py — golden images of BasicRayTracingWithC++/main.cpp FROM THE REFERENCE'S OWN SOURCE.

Runs only where oracle/_ref/libref_basic_det.so exists (oracle/build_ref.py, needs /root/reference): that library is the
reference's main.cpp with a reproducible counter-based random stream plugged into its randf() (oracle/ref_basic_wrap.cpp
lists the edits).  basic.npz holds the per-pixel sum images (float64, row 0 = top) for a few (width, height, samples, seed).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CASES = [(48, 48, 6, 7), (33, 20, 3, 1), (64, 40, 2, 12345)]


def ref_basic_render(w, h, samples, seed):
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_basic_det.so"))
    L.ref_basic_render.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint, C.c_void_p]
    img = np.zeros((h, w, 3), np.float64)
    L.ref_basic_render(w, h, samples, seed, img.ctypes.data)
    return img


def main():
    out = {}
    for w, h, n, seed in CASES:
        out["w%d_h%d_n%d_s%d" % (w, h, n, seed)] = ref_basic_render(w, h, n, seed)
    np.savez_compressed(os.path.join(HERE, "basic.npz"), **out)
    print("basic.npz:", {k: (v.shape, float(v.mean())) for k, v in out.items()})


if __name__ == "__main__":
    main()
