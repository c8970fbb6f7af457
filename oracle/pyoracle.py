"""oracle/pyoracle.py — TEST INFRASTRUCTURE: ctypes access to the CPU oracle (libptoracle.so) and, when it was
built in this container, to the reference's own shader source compiled as C++ (oracle/_ref/).

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_f32p = C.POINTER(C.c_float)
COUNTER_NAMES = ["rays", "ref_inner", "ref_leaf", "ref_tris", "ref_closer", "cull_inner", "cull_tris", "cull_hits",
                 "max_sp", "over256", "texels", "wide_max_sp", "model_mismatch"]


def build(force=False):
    """Compiles oracle/ptoracle.cpp (parity flags: -O2 -ffp-contract=off) -> oracle/libptoracle.so."""
    src, out = os.path.join(HERE, "ptoracle.cpp"), os.path.join(HERE, "libptoracle.so")
    deps = [src, os.path.join(HERE, "sobol_v.inc"), os.path.join(ROOT, "include", "ptmath.h")]
    if force or not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC",
                               "-shared", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), src, "-o", out])
    return out


def _fp(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libptoracle.so")
        if not os.path.exists(path):
            build()
        _lib = C.CDLL(path)
        _lib.oracle_render.restype = C.c_int
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def render(variant, max_bounce, tris, nodes, hdr, hdrCache, width, height, eye, cam, first_frame, n_frames,
           accum=None, math=1, counters=False, region=None, env_sampler=0):
    """Returns (accum [H,W,4] float32, primary_ids [H,W] int32 of the last frame, counters dict or None).
    env_sampler = 1: variant 2 picks its light samples by per-sample CDF binary search (the product's ENV_CDF_SEARCH)
    instead of the reference's hdrCache table lookup."""
    lib().oracle_set_env_sampler(int(env_sampler))
    tris = np.ascontiguousarray(tris, np.float32); nodes = np.ascontiguousarray(nodes, np.float32)
    hdr = np.ascontiguousarray(hdr, np.float32)
    hc = np.ascontiguousarray(hdrCache, np.float32) if hdrCache is not None else None
    eye = np.ascontiguousarray(eye, np.float32); cam = np.ascontiguousarray(cam, np.float32)
    if accum is None:
        accum = np.zeros((height, width, 4), np.float32)
    ids = np.full((height, width), -1, np.int32)
    cnt = (C.c_ulonglong * 13)() if counters else None
    x0, y0, x1, y1 = region if region else (0, 0, width, height)
    rc = lib().oracle_render(variant, max_bounce, math, _fp(tris), len(tris), _fp(nodes), len(nodes), _fp(hdr), _fp(hc),
                             hdr.shape[1], hdr.shape[0], width, height, _fp(eye), _fp(cam), C.c_uint(first_frame),
                             C.c_uint(n_frames), _fp(accum), ids.ctypes.data_as(C.POINTER(C.c_int)), cnt, x0, y0, x1, y1)
    lib().oracle_set_env_sampler(0)
    if rc == -2:
        raise OverflowError("oracle: a ray needed more than 16384 traversal stack entries")
    if rc != 0:
        raise RuntimeError("oracle_render rejected its arguments")
    return accum, ids, (dict(zip(COUNTER_NAMES, [int(v) for v in cnt])) if counters else None)


def env_cdf_sample(hdr, xi):
    """Texels (x, y) the CDF-search sampler picks for the random pairs xi [n, 2]."""
    hdr = np.ascontiguousarray(hdr, np.float32); xi = np.ascontiguousarray(xi, np.float32)
    out = np.empty((len(xi), 2), np.int32)
    lib().oracle_env_cdf_sample(_fp(hdr), hdr.shape[1], hdr.shape[0], len(xi), _fp(xi), out.ctypes.data_as(C.POINTER(C.c_int)))
    return out


def env_pdf(hdr):
    """calculateHdrCache's per-texel pdf (its B channel), [H, W]."""
    hdr = np.ascontiguousarray(hdr, np.float32)
    out = np.empty(hdr.shape[:2], np.float32)
    lib().oracle_env_pdf(_fp(hdr), hdr.shape[1], hdr.shape[0], _fp(out))
    return out


def wang_hash_seq(px, py, frame, n):
    h = np.empty(n, np.uint32); r = np.empty(n, np.float32)
    lib().oracle_wang_hash_seq(px, py, C.c_uint(frame), n, h.ctypes.data_as(C.POINTER(C.c_uint)), _fp(r))
    return h, r


def sobol_vec2(i, b):
    uv = np.empty(2, np.float32)
    lib().oracle_sobol_vec2(C.c_uint(i), C.c_uint(b), _fp(uv))
    return uv


def num_threads():
    return lib().oracle_num_threads()


# ---- the reference's own sources, compiled (present only when oracle/build_ref.py ran in this container) ----
REF_DIR = os.path.join(HERE, "_ref")
REF_SHADER = {0: "ogl", 1: "disney", 2: "is"}


def ref_shader_available(variant, math="ptmath"):
    return os.path.exists(os.path.join(REF_DIR, "libref_shader_%s_%s.so" % (REF_SHADER[variant], math)))


def ref_shader_threads(variant, math="ptmath"):
    """OpenMP threads the compiled reference shader runs on."""
    L = C.CDLL(os.path.join(REF_DIR, "libref_shader_%s_%s.so" % (REF_SHADER[variant], math)))
    return int(L.ref_shader_num_threads()) if hasattr(L, "ref_shader_num_threads") else num_threads()


def ref_shader_render(variant, tris, nodes, hdr, hdrCache, width, height, eye, cam, first_frame, n_frames, accum=None,
                      math="ptmath", region=None):
    L = C.CDLL(os.path.join(REF_DIR, "libref_shader_%s_%s.so" % (REF_SHADER[variant], math)))
    tris = np.ascontiguousarray(tris, np.float32); nodes = np.ascontiguousarray(nodes, np.float32)
    hdr = np.ascontiguousarray(hdr, np.float32)
    hc = np.ascontiguousarray(hdrCache, np.float32) if hdrCache is not None else None
    eye = np.ascontiguousarray(eye, np.float32); cam = np.ascontiguousarray(cam, np.float32)
    if accum is None:
        accum = np.zeros((height, width, 4), np.float32)
    x0, y0, x1, y1 = region if region else (0, 0, width, height)
    L.ref_shader_render(_fp(tris), len(tris), _fp(nodes), len(nodes), _fp(hdr), _fp(hc), hdr.shape[1], hdr.shape[0],
                        width, height, _fp(eye), _fp(cam), C.c_uint(first_frame), C.c_uint(n_frames), _fp(accum),
                        x0, y0, x1, y1)
    return accum
