#!/usr/bin/env python3
"""tests/golden/make_golden.py — regenerates the committed golden vectors FROM THE REFERENCE'S OWN SOURCES.

Runs only where /root/reference exists (oracle/build_ref.py must have produced oracle/_ref/).  Everything written
here comes out of code compiled from the reference tree:
  kats.json    rand() sequences and sobolVec2 / CranleyPattersonRotation values from pass1.fsh (as float32 bit
               patterns), getTransformMatrix / camera matrices from IS/main.cpp + glm, SHA-256 of the node and
               triangle arrays produced by buildBVH / buildBVHwithSAH on seeded synthetic meshes, SHA-256 of
               calculateHdrCache's output
  images.npz   accumulated RGBA32F framebuffers of the three pass1.fsh shaders (compiled as C++ with the ptmath
               transcendental set) on the small synthetic scenes of opengl_ray_tracing_b200.scenes, after 1 and 4
               frames, 48x32 pixels
The inputs (meshes, HDR map) are produced by libptscene's seeded generators, which are deterministic.
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from opengl_ray_tracing_b200 import scene as S, scenes  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


def bits(a):
    return [int(v) for v in np.asarray(a, np.float32).view(np.uint32).ravel()]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


IMG_W, IMG_H = 48, 32
IMG_SCENES = {"lambert": 0, "disney": 1, "is": 2}
CAMERA = (10.0, 25.0, 4.0)


def golden_scene(kind):
    return scenes.build(kind, "1k", hdr_res=(128, 64), camera=CAMERA, use_cache=False)


def main():
    RH = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_host.so"))
    RS = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_shader_is_libm.so"))
    kats = {"rand": [], "sobol_cp": [], "transform": [], "camera": [], "bvh": [], "hdr_cache": []}
    for (px, py, frame) in [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1), (1919, 1079, 4095), (37, 11, 7)]:
        out = np.empty(8, np.float32)
        RS.ref_shader_rand_seq(px, py, 1920, 1080, C.c_uint(frame), 8, fp(out))
        kats["rand"].append({"px": px, "py": py, "frame": frame, "bits": bits(out)})
    for i in range(1, 17):
        for b in range(4):
            out = np.empty(4, np.float32)
            RS.ref_shader_sobol_cp(5, 9, 1920, 1080, C.c_uint(i), C.c_uint(b), fp(out))
            kats["sobol_cp"].append({"i": i, "b": b, "px": 5, "py": 9, "bits": bits(out)})
    rng = np.random.default_rng(7)
    for k in range(6):
        r = (rng.uniform(-180, 180, 3) if k else np.zeros(3)).astype(np.float32)
        t = rng.normal(size=3).astype(np.float32)
        s = rng.uniform(0.01, 20, 3).astype(np.float32)
        out = np.empty(16, np.float32)
        RH.ref_transform_matrix(fp(r), fp(t), fp(s), fp(out))
        kats["transform"].append({"rot": bits(r), "trans": bits(t), "scale": bits(s), "bits": bits(out)})
    RH.ref_camera.argtypes = [C.c_float] * 3 + [f32p, f32p]
    for up, rot, rr in [(0.0, 0.0, 4.0), (10.0, 25.0, 4.0), (-35.5, 200.25, 2.5)]:
        e, c = np.empty(3, np.float32), np.empty(16, np.float32)
        RH.ref_camera(up, rot, rr, fp(e), fp(c))
        kats["camera"].append({"up": up, "rot": rot, "r": rr, "eye": bits(e), "cam": bits(c)})
    # BVH builders on a seeded mesh (+ the two quads of the Lambert scene)
    RH.ref_build_bvh.restype = C.c_int
    for freq, leaf in [(5, 8), (16, 8), (16, 2), (30, 8)]:
        for sah in (0, 1):
            sc = S.Scene()
            scenes._populate(sc, "lambert", freq)
            tris = sc.triangles_encoded().copy()
            T = len(tris)
            nodes = np.empty((2 * T + 2, 12), np.float32)
            nn = RH.ref_build_bvh(fp(tris), T, leaf, sah, fp(nodes), len(nodes))
            kats["bvh"].append({"freq": freq, "leaf": leaf, "sah": sah, "nTriangles": T, "nNodes": nn,
                                "nodes_sha256": sha(nodes[:nn]), "tris_sha256": sha(tris)})
    for (w, h, seed) in [(128, 64, 1), (256, 128, 2)]:
        hdr = S.synthHdr(w, h, seed)
        out = np.empty_like(hdr)
        RH.ref_hdr_cache(fp(hdr), w, h, fp(out))
        kats["hdr_cache"].append({"w": w, "h": h, "seed": seed, "hdr_sha256": sha(hdr), "cache_sha256": sha(out)})
    with open(os.path.join(HERE, "kats.json"), "w") as f:
        json.dump(kats, f, indent=1)
    # framebuffers from the reference shaders
    imgs = {}
    for kind, variant in IMG_SCENES.items():
        sc = golden_scene(kind)
        acc = None
        done = 0
        for upto in (1, 4):
            acc = O.ref_shader_render(variant, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, IMG_W, IMG_H, sc.eye, sc.cam, done, upto - done,
                                      accum=acc, math="ptmath")
            done = upto
            imgs["%s_f%d" % (kind, upto)] = acc.copy()
        imgs["%s_inputs_sha256" % kind] = np.frombuffer(bytes.fromhex(sha(sc.tris) + sha(sc.nodes) + sha(sc.hdr)), np.uint8)
    np.savez_compressed(os.path.join(HERE, "images.npz"), **imgs)
    print("golden vectors written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
