"""Shared test helpers."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_CAMERA = (10.0, 25.0, 4.0)
GOLDEN_W, GOLDEN_H = 48, 32
KINDS = {"lambert": 0, "disney": 1, "is": 2}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def from_bits(b):
    return np.array(b, dtype=np.uint32).view(np.float32)


def rrmse(a, b):
    a = np.asarray(a)[..., :3].astype(np.float64)
    b = np.asarray(b)[..., :3].astype(np.float64)
    return float(np.sqrt(((a - b) ** 2).sum() / max(1e-300, (b ** 2).sum())))


def golden_scene(kind):
    from opengl_ray_tracing_b200 import scenes
    return scenes.build(kind, "1k", hdr_res=(128, 64), camera=GOLDEN_CAMERA, use_cache=False)


def small_scene(kind, size="1k", hdr_res=(128, 64), camera=GOLDEN_CAMERA):
    from opengl_ray_tracing_b200 import scenes
    return scenes.build(kind, size, hdr_res=hdr_res, camera=camera, use_cache=False)


def oracle_render(sc, W, H, first, n, accum=None, math=1, counters=False, region=None, max_bounce=None):
    from oracle import pyoracle as O
    return O.render(sc.variant, max_bounce or sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam,
                    first, n, accum=accum, math=math, counters=counters, region=region)
