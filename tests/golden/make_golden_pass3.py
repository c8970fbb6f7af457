#!/usr/bin/env python3
"""tests/golden/make_golden_pass3.py — golden display bytes FROM THE REFERENCE'S pass3.fsh.

Runs only where /root/reference exists (oracle/build_ref.py -> oracle/_ref/libref_pass3.so = shaders/pass3.fsh compiled as
C++ through oracle/glsl_compat.h).  Input: the golden framebuffers of tests/golden/images.npz (<kind>_f4, produced by the
reference's pass1.fsh) plus one synthetic framebuffer with values a path tracer can produce and a display pass must survive
(0, denormal, > 1, 1e4, inf, NaN).  Output pass3.npz: <name>_display = uint8 [H, W, 4], row 0 = TOP row, i.e. what the
reference's window shows: fragColor through GL's RGBA8 conversion, round(clamp(c, 0, 1) * 255) (NaN -> 0), which is the one
step that is the GL pipeline's and not the shader's."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
f32p = C.POINTER(C.c_float)


def pass3_display(L, fb):
    fb = np.ascontiguousarray(fb, np.float32)
    H, W = fb.shape[:2]
    out = np.empty_like(fb)
    L.ref_pass3_render(fb.ctypes.data_as(f32p), W, H, out.ctypes.data_as(f32p))
    v = np.nan_to_num(out[..., :3], nan=0.0, posinf=1.0, neginf=0.0)
    q = (np.clip(v, np.float32(0), np.float32(1)) * np.float32(255.0) + np.float32(0.5)).astype(np.int32).astype(np.uint8)
    disp = np.concatenate([q, np.full((H, W, 1), 255, np.uint8)], axis=2)
    return disp[::-1].copy()                                   # window rows run top to bottom; the framebuffer's row 0 is the bottom


def special_framebuffer():
    rng = np.random.default_rng(11)
    fb = (rng.random((24, 40, 4), dtype=np.float32) ** 4 * np.float32(30.0)).astype(np.float32)
    fb[..., 3] = 1.0
    fb[0, :8, :3] = [[0, 0, 0], [1e-42, 1e-42, 1e-42], [1, 1, 1], [1e4, 1e4, 1e4], [np.inf, 1, 1], [np.nan, 0.5, 0.5], [0.5, 2.0, 8.0], [1.5, 1.5, 1.5]]
    return fb


def main():
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_pass3.so"))
    imgs = np.load(os.path.join(HERE, "images.npz"))
    out = {}
    for kind in ("lambert", "disney", "is"):
        out[kind + "_display"] = pass3_display(L, imgs[kind + "_f4"])
    fb = special_framebuffer()
    out["special_fb"] = fb
    out["special_display"] = pass3_display(L, fb)
    np.savez_compressed(os.path.join(HERE, "pass3.npz"), **out)
    print("pass3.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
