"""ctypes mirror of the Basic-tracer entry points of libpt.so (BasicRayTracingWithC++/main.cpp on the GPU, include/pt.h).

    shapes = default_scene()                      # the reference's Cornell box literal, main.cpp:304-353
    image, ms = render(shapes, 256, 256, 200)     # per-pixel SUM over the samples, float64 [H, W, 3], row 0 = top
    rgb8 = imshow(image)                          # the reference's gamma + 8-bit conversion, main.cpp:169-190
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import renderer as R


class PtBasicShape(C.Structure):
    _fields_ = [("kind", C.c_int), ("p1", C.c_float * 3), ("p2", C.c_float * 3), ("p3", C.c_float * 3), ("radius", C.c_double),
                ("color", C.c_float * 3), ("isEmissive", C.c_int), ("specularRate", C.c_double), ("roughness", C.c_double),
                ("refractRate", C.c_double), ("refractAngle", C.c_double), ("refractRoughness", C.c_double)]


def _lib():
    L = R.lib()
    if not getattr(L, "_basic_ready", False):
        L.pt_basic_default_scene.restype, L.pt_basic_default_scene.argtypes = C.c_int, [C.POINTER(PtBasicShape), C.c_int]
        L.pt_basic_render.restype = C.c_int
        L.pt_basic_render.argtypes = [C.c_int, C.POINTER(PtBasicShape), C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.POINTER(C.c_double)]
        L.pt_basic_imshow.restype, L.pt_basic_imshow.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.pt_basic_last_error.restype, L.pt_basic_last_error.argtypes = C.c_char_p, []
        L._basic_ready = True
    return L


def default_scene():
    L = _lib()
    n = L.pt_basic_default_scene(None, 0)
    arr = (PtBasicShape * n)()
    L.pt_basic_default_scene(arr, n)
    return arr


def render(shapes, width, height, samples, seed=1, device=0):
    """-> (float64 [H, W, 3] sum image, device milliseconds).  Raises PtError without a CUDA device (no CPU fallback)."""
    L = _lib()
    image = np.zeros((height, width, 3), np.float64)
    ms = C.c_double(0.0)
    rc = L.pt_basic_render(device, shapes, len(shapes), width, height, samples, seed, image.ctypes.data, C.byref(ms))
    if rc != 0:
        raise R.PtError("pt_basic_render: %s" % L.pt_basic_last_error().decode())
    return image, ms.value


def imshow(image):
    L = _lib()
    img = np.ascontiguousarray(image, np.float64)
    out = np.empty(img.shape, np.uint8)
    rc = L.pt_basic_imshow(img.ctypes.data, img.shape[1], img.shape[0], out.ctypes.data)
    if rc != 0:
        raise R.PtError("pt_basic_imshow: %s" % L.pt_basic_last_error().decode())
    return out
