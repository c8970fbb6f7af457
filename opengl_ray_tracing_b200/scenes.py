"""Synthetic scenes for BASELINE.json's configs — the reference's own scene literals with seeded procedural
meshes standing in for its missing models/*.obj.

Every scene is assembled through the host scene library exactly as the reference's main() does it:
readObj-style mesh insertion -> buildBVHwithSAH(…, 8) -> encoded arrays -> HDR map (+ calculateHdrCache).
Built scenes are cached on disk (the SAH build is the slow part at >= 1M triangles).

    lambert   config 2: OpenglRayTracing/main.cpp:650-670   (mesh + ground quad + emissive quad), variant 0
    disney    config 3: DisneyBRDF/main.cpp:720-727         (one metallic clear-coated mesh),      variant 1
    is        config 4: ImportanceSampling_…/main.cpp:756-771 (gold mesh + glossy floor),           variant 2
    full      config 5: ten meshes + floor + light,                                                variant 0
"""
from __future__ import annotations

import hashlib
import os
from dataclasses import dataclass

import numpy as np

from . import scene as S

CACHE_DIR = os.environ.get("PT_SCENE_CACHE", os.path.join(os.path.dirname(os.path.abspath(__file__)), "build", "scene_cache"))

# icosphere frequency n -> 20 n^2 triangles
FREQ_FOR = {"1k": 7, "10k": 22, "100k": 71, "1M": 224, "10M": 707}


@dataclass
class BuiltScene:
    name: str
    variant: int
    max_bounce: int
    tris: np.ndarray        # [T, 36]
    nodes: np.ndarray       # [N, 12]
    hdr: np.ndarray         # [h, w, 3]
    hdrCache: np.ndarray | None
    eye: np.ndarray
    cam: np.ndarray

    @property
    def nTriangles(self):
        return len(self.tris)

    @property
    def nNodes(self):
        return len(self.nodes)


def _T(rot, trans, scale):
    return S.getTransformMatrix(rot, trans, scale)


def _populate(sc: S.Scene, kind: str, freq: int):
    if kind == "lambert":
        sc.addIcosphere(freq, 0.15, 1, S.Material(baseColor=(0, 1, 1)), _T((0, 0, 0), (0.3, -0.4, 0), (1.5, 1.5, 1.5)), True)
        sc.addQuad(S.Material(baseColor=(0.725, 0.71, 0.68)), _T((0, 0, 0), (0, -1.4, 0), (18.83, 0.01, 18.83)))
        sc.addQuad(S.Material(baseColor=(1, 1, 1), emissive=(20, 20, 20)), _T((0, 0, 0), (0.0, 1.38, -0.0), (0.7, 0.01, 0.7)))
    elif kind == "disney":
        m = S.Material(baseColor=(0.75, 0.7, 0.15), roughness=0.15, metallic=1.0, specular=0.5, clearcoat=1.0)
        sc.addIcosphere(freq, 0.15, 2, m, _T((0, 0, 0), (0, -0.4, 0), (1.75, 1.75, 1.75)), True)
    elif kind == "is":
        m = S.Material(baseColor=(1, 0.73, 0.25), roughness=0.5, specular=1.0, metallic=1.0, clearcoat=1.0, clearcoatGloss=0.0)
        sc.addIcosphere(freq, 0.15, 3, m, _T((0, 0, 0), (0, -0.5, 0), (0.75, 0.75, 0.75)), True)
        # the reference mutates `m` in place (M:766-769): clearcoat stays 1 on the floor
        f = S.Material(baseColor=(1, 1, 1), roughness=0.01, metallic=0.1, specular=1.0, clearcoat=1.0, clearcoatGloss=0.0)
        sc.addQuad(f, _T((0, 0, 0), (0, -0.5, 0), (13.0, 0.01, 13.0)))
    elif kind == "full":
        pos = [(-1.6, -0.6, -1.2), (0.0, -0.6, -1.4), (1.6, -0.6, -1.2), (-2.0, -0.6, 0.2), (-0.7, -0.6, 0.0),
               (0.7, -0.6, 0.0), (2.0, -0.6, 0.2), (-1.2, -0.6, 1.3), (0.0, -0.6, 1.5), (1.2, -0.6, 1.3)]
        for k, p in enumerate(pos):
            col = (0.3 + 0.07 * k, 1.0 - 0.06 * k, 0.4 + 0.05 * ((k * 3) % 10))
            sc.addIcosphere(freq, 0.15, 10 + k, S.Material(baseColor=col), _T((0, 36.0 * k, 0), p, (0.6, 0.6, 0.6)), True)
        sc.addQuad(S.Material(baseColor=(0.725, 0.71, 0.68)), _T((0, 0, 0), (0, -1.0, 0), (18.83, 0.01, 18.83)))
        sc.addQuad(S.Material(baseColor=(1, 1, 1), emissive=(20, 20, 20)), _T((0, 0, 0), (0.0, 2.2, 0.0), (1.4, 0.01, 1.4)))
    else:
        raise ValueError("unknown scene kind %r" % kind)


KINDS = {"lambert": (0, 2), "disney": (1, 5), "is": (2, 2), "full": (0, 2)}
LEAF_N = 8                                                       # buildBVH*(…, 8), M:787-788


def _builder_fingerprint() -> str:
    """What a cached scene depends on besides (kind, frequency, builder): the scene literals of _populate above and the
    sources of the host library that generates the meshes and builds the trees.  A change to any of them changes the key."""
    import inspect
    h = hashlib.sha1(inspect.getsource(_populate).encode())
    here = os.path.dirname(os.path.abspath(__file__))
    for name in ("pts_bvh.cpp", "pts_meshgen.cpp", "pts_scene.cpp", "pts_scene.h", "pts_math.h"):
        with open(os.path.join(here, "csrc", "scene", name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:12]


def build(kind: str, size: str = "10k", hdr_res=(1024, 512), hdr_path: str | None = None, threads: int | None = None,
          camera=(0.0, 0.0, 4.0), use_cache: bool = True, builder: str = "sah") -> BuiltScene:
    """kind in KINDS; size in FREQ_FOR (per-mesh triangle count; 'full' has ten meshes of size/10… see below)."""
    variant, bounces = KINDS[kind]
    freq = FREQ_FOR[size] if isinstance(size, str) else int(size)
    if kind == "full" and isinstance(size, str):
        freq = max(2, int(round(freq / np.sqrt(10.0))))          # ten meshes share the triangle budget
    threads = threads or min(16, os.cpu_count() or 1)
    os.makedirs(CACHE_DIR, exist_ok=True)
    # builder "sah" = buildBVHwithSAH (what the reference's main() calls, M:788); "median" = buildBVH, the alternative the
    # reference keeps commented out one line above (M:787) — seconds instead of hours at 10 M triangles
    key = hashlib.sha1(("%s-%d-%s-leaf%d-%s" % (kind, freq, builder, LEAF_N, _builder_fingerprint())).encode()).hexdigest()[:16]
    path = os.path.join(CACHE_DIR, "%s_%d_%s.ptsc" % (kind, freq, key))
    sc = S.Scene()
    if use_cache and os.path.exists(path):
        sc.load(path)
    else:
        _populate(sc, kind, freq)
        if builder == "sah":
            sc.buildBVHwithSAH(LEAF_N, threads)                  # M:788
        else:
            sc.buildBVH(LEAF_N, threads)                         # M:787
        if use_cache:
            tmp = "%s.%d.tmp" % (path, os.getpid())              # atomic: another rank may be waiting to load this file
            sc.save(tmp)
            os.replace(tmp, path)
    tris, nodes = sc.triangles_encoded(), sc.nodes_encoded()
    sc.close()
    if hdr_path:
        w, h, hdr = S.HDRLoader.load(hdr_path)
    else:
        w, h = hdr_res
        hdr = S.synthHdr(w, h, 1)
    cache = S.calculateHdrCache(hdr, w, h) if variant == 2 else None
    eye, cam = S.camera(*camera)
    return BuiltScene("%s-%s" % (kind, size), variant, bounces, tris, nodes, hdr, cache, eye, cam)
