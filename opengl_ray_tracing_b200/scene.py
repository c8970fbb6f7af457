"""Host-side scene API — a ctypes mirror of libptscene.so (include/ptscene.h).

The names follow the reference's own host functions (ImportanceSampling_LowDiscrepancySequence/main.cpp,
"M:"): ``Material`` (M:26-41), ``getTransformMatrix`` (M:243-259), ``readObj`` (M:262-373), ``buildBVH`` /
``buildBVHwithSAH`` (M:377-552), ``HDRLoader.load`` (hdrloader.cpp:29-97), ``calculateHdrCache``
(M:555-652) and the camera lines of ``display()`` (M:671-674).  All the arithmetic is in C++
(csrc/scene/*.cpp); this module only moves numpy arrays across the C ABI.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libptscene.so")
_lib = None

TRI_FLOATS = 36
NODE_FLOATS = 12


class PtsMaterial(C.Structure):
    _fields_ = [("emissive", C.c_float * 3), ("baseColor", C.c_float * 3),
                ("subsurface", C.c_float), ("metallic", C.c_float), ("specular", C.c_float),
                ("specularTint", C.c_float), ("roughness", C.c_float), ("anisotropic", C.c_float),
                ("sheen", C.c_float), ("sheenTint", C.c_float), ("clearcoat", C.c_float),
                ("clearcoatGloss", C.c_float), ("IOR", C.c_float), ("transmission", C.c_float)]


def lib():
    """Loads libptscene.so (built by __graft_entry__.build()); raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError("libptscene.so is not built: run `python __graft_entry__.py build` "
                           "(or opengl_ray_tracing_b200/build.py)")
    L = C.CDLL(_LIB_PATH)
    f32p, u32p, vp, ip = C.POINTER(C.c_float), C.POINTER(C.c_uint), C.c_void_p, C.POINTER(C.c_int)
    mp = C.POINTER(PtsMaterial)
    sig = {
        "pts_material_default": (None, [mp]),
        "pts_get_transform_matrix": (None, [f32p, f32p, f32p, f32p]),
        "pts_camera": (None, [C.c_float, C.c_float, C.c_float, f32p, f32p]),
        "pts_scene_create": (vp, []),
        "pts_scene_destroy": (None, [vp]),
        "pts_read_obj": (C.c_int, [vp, C.c_char_p, mp, f32p, C.c_int]),
        "pts_add_mesh": (C.c_int, [vp, f32p, C.c_int, u32p, C.c_int, mp, f32p, C.c_int]),
        "pts_add_triangles": (C.c_int, [vp, f32p, C.c_int]),
        "pts_add_icosphere": (C.c_int, [vp, C.c_int, C.c_float, C.c_uint, mp, f32p, C.c_int]),
        "pts_add_quad": (C.c_int, [vp, mp, f32p]),
        "pts_write_obj_icosphere": (C.c_int, [C.c_char_p, C.c_int, C.c_float, C.c_uint]),
        "pts_triangle_count": (C.c_int, [vp]),
        "pts_node_count": (C.c_int, [vp]),
        "pts_build_bvh": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
        "pts_triangles_encoded": (f32p, [vp]),
        "pts_nodes_encoded": (f32p, [vp]),
        "pts_scene_save": (C.c_int, [vp, C.c_char_p]),
        "pts_scene_load": (C.c_int, [vp, C.c_char_p]),
        "pts_hdr_load": (C.c_int, [C.c_char_p, ip, ip, C.POINTER(f32p)]),
        "pts_hdr_write": (C.c_int, [C.c_char_p, f32p, C.c_int, C.c_int]),
        "pts_hdr_cache": (C.c_int, [f32p, C.c_int, C.c_int, f32p]),
        "pts_hdr_synth": (C.c_int, [C.c_int, C.c_int, C.c_uint, f32p]),
        "pts_free": (None, [vp]),
        "pts_write_pfm": (C.c_int, [C.c_char_p, f32p, C.c_int, C.c_int]),
        "pts_write_ppm_tonemapped": (C.c_int, [C.c_char_p, f32p, C.c_int, C.c_int]),
        "pts_write_png": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
        "pts_last_error": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTS = ["pts_material_default", "pts_get_transform_matrix", "pts_camera", "pts_scene_create",
           "pts_scene_destroy", "pts_read_obj", "pts_add_mesh", "pts_add_triangles", "pts_add_icosphere",
           "pts_add_quad", "pts_write_obj_icosphere", "pts_triangle_count", "pts_node_count",
           "pts_build_bvh", "pts_triangles_encoded", "pts_nodes_encoded", "pts_scene_save",
           "pts_scene_load", "pts_hdr_load", "pts_hdr_write", "pts_hdr_cache", "pts_hdr_synth", "pts_free",
           "pts_write_pfm", "pts_write_ppm_tonemapped", "pts_write_png", "pts_last_error"]


class SceneError(RuntimeError):
    pass


def _check(rc, what):
    if rc < 0:
        raise SceneError("%s failed (%d): %s" % (what, rc, lib().pts_last_error().decode("utf-8", "replace")))
    return rc


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


@dataclass
class Material:
    """struct Material, M:26-41 (same defaults)."""
    emissive: tuple = (0.0, 0.0, 0.0)
    baseColor: tuple = (1.0, 1.0, 1.0)
    subsurface: float = 0.0
    metallic: float = 0.0
    specular: float = 0.0
    specularTint: float = 0.0
    roughness: float = 0.0
    anisotropic: float = 0.0
    sheen: float = 0.0
    sheenTint: float = 0.0
    clearcoat: float = 0.0
    clearcoatGloss: float = 0.0
    IOR: float = 1.0
    transmission: float = 0.0

    def to_c(self) -> PtsMaterial:
        m = PtsMaterial()
        m.emissive[:] = self.emissive
        m.baseColor[:] = self.baseColor
        for k in ("subsurface", "metallic", "specular", "specularTint", "roughness", "anisotropic", "sheen",
                  "sheenTint", "clearcoat", "clearcoatGloss", "IOR", "transmission"):
            setattr(m, k, getattr(self, k))
        return m

    def to_floats(self) -> np.ndarray:
        return np.frombuffer(bytes(self.to_c()), dtype=np.float32).copy()


def getTransformMatrix(rotateCtrl, translateCtrl, scaleCtrl) -> np.ndarray:
    """M:243-259.  Returns the 16 floats column-major (glm::value_ptr order)."""
    r, t, s, out = _f32(rotateCtrl), _f32(translateCtrl), _f32(scaleCtrl), np.empty(16, np.float32)
    lib().pts_get_transform_matrix(_fp(r), _fp(t), _fp(s), _fp(out))
    return out


def camera(upAngle=0.0, rotateAngle=0.0, r=4.0):
    """display(), M:671-674 -> (eye[3], cameraRotate[16] column-major). Defaults = M:147-149."""
    eye, cam = np.empty(3, np.float32), np.empty(16, np.float32)
    lib().pts_camera(upAngle, rotateAngle, r, _fp(eye), _fp(cam))
    return eye, cam


class Scene:
    """std::vector<Triangle> triangles + std::vector<BVHNode> nodes of the reference's main()."""

    def __init__(self):
        self._h = lib().pts_scene_create()
        if not self._h:
            raise MemoryError("pts_scene_create")

    def close(self):
        if getattr(self, "_h", None):
            lib().pts_scene_destroy(self._h)
            self._h = None

    __del__ = close

    # -- geometry -------------------------------------------------------------------------------
    def readObj(self, filepath, material: Material, trans, smoothNormal: bool) -> int:
        t = _f32(trans)
        return _check(lib().pts_read_obj(self._h, os.fsencode(filepath), C.byref(material.to_c()), _fp(t),
                                         int(smoothNormal)), "readObj")

    def addMesh(self, vertices, indices, material: Material, trans, smoothNormal: bool) -> int:
        v, t = _f32(vertices).reshape(-1, 3), _f32(trans)
        i = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        return _check(lib().pts_add_mesh(self._h, _fp(v), len(v), i.ctypes.data_as(C.POINTER(C.c_uint)), len(i),
                                         C.byref(material.to_c()), _fp(t), int(smoothNormal)), "addMesh")

    def addTriangles(self, tris) -> int:
        a = _f32(tris).reshape(-1, TRI_FLOATS)
        return _check(lib().pts_add_triangles(self._h, _fp(a), len(a)), "addTriangles")

    def addIcosphere(self, frequency, amplitude, seed, material: Material, trans, smoothNormal=True) -> int:
        t = _f32(trans)
        return _check(lib().pts_add_icosphere(self._h, frequency, amplitude, seed, C.byref(material.to_c()),
                                              _fp(t), int(smoothNormal)), "addIcosphere")

    def addQuad(self, material: Material, trans) -> int:
        t = _f32(trans)
        return _check(lib().pts_add_quad(self._h, C.byref(material.to_c()), _fp(t)), "addQuad")

    # -- BVH ------------------------------------------------------------------------------------
    def buildBVH(self, n=8, threads=1) -> int:
        """M:377-428 (+ dummy node 0).  Returns nNodes."""
        return _check(lib().pts_build_bvh(self._h, n, 0, threads), "buildBVH")

    def buildBVHwithSAH(self, n=8, threads=1) -> int:
        """M:431-552 (+ dummy node 0).  Returns nNodes."""
        return _check(lib().pts_build_bvh(self._h, n, 1, threads), "buildBVHwithSAH")

    @property
    def nTriangles(self) -> int:
        return lib().pts_triangle_count(self._h)

    @property
    def nNodes(self) -> int:
        return lib().pts_node_count(self._h)

    def triangles_encoded(self) -> np.ndarray:
        """Triangle_encoded[nTriangles] as float32 [T, 36] (M:793-812). A copy."""
        n = self.nTriangles
        if n == 0:
            return np.zeros((0, TRI_FLOATS), np.float32)
        p = lib().pts_triangles_encoded(self._h)
        return np.ctypeslib.as_array(p, shape=(n, TRI_FLOATS)).copy()

    def nodes_encoded(self) -> np.ndarray:
        """BVHNode_encoded[nNodes] as float32 [N, 12] (M:815-821). A copy."""
        n = self.nNodes
        if n == 0:
            return np.zeros((0, NODE_FLOATS), np.float32)
        p = lib().pts_nodes_encoded(self._h)
        return np.ctypeslib.as_array(p, shape=(n, NODE_FLOATS)).copy()

    def save(self, path):
        _check(lib().pts_scene_save(self._h, os.fsencode(path)), "scene_save")

    def load(self, path):
        _check(lib().pts_scene_load(self._h, os.fsencode(path)), "scene_load")


class HDRLoader:
    """hdrloader.h: ``HDRLoader::load`` -> (width, height, cols[h, w, 3])."""

    @staticmethod
    def load(fileName):
        w, h, data = C.c_int(), C.c_int(), C.POINTER(C.c_float)()
        _check(lib().pts_hdr_load(os.fsencode(fileName), C.byref(w), C.byref(h), C.byref(data)), "HDRLoader.load")
        try:
            arr = np.ctypeslib.as_array(data, shape=(h.value, w.value, 3)).copy()
        finally:
            lib().pts_free(data)
        return w.value, h.value, arr


def calculateHdrCache(HDR, width, height) -> np.ndarray:
    """M:555-652 -> float32 [h, w, 3] = (x/w, y/h, pdf)."""
    src = _f32(HDR).reshape(height, width, 3)
    out = np.empty((height, width, 3), np.float32)
    _check(lib().pts_hdr_cache(_fp(src), width, height, _fp(out)), "calculateHdrCache")
    return out


def synthHdr(width, height, seed=1) -> np.ndarray:
    out = np.empty((height, width, 3), np.float32)
    _check(lib().pts_hdr_synth(width, height, seed, _fp(out)), "synthHdr")
    return out


def writeHdr(path, data):
    d = _f32(data)
    _check(lib().pts_hdr_write(os.fsencode(path), _fp(d), d.shape[1], d.shape[0]), "writeHdr")


def writePfm(path, rgba):
    d = _f32(rgba)
    _check(lib().pts_write_pfm(os.fsencode(path), _fp(d), d.shape[1], d.shape[0]), "writePfm")


def writePng(path, pixels):
    """pixels: uint8 [H, W, 3 or 4], top row first (what PathTracer.readback_display returns)."""
    d = np.ascontiguousarray(pixels, dtype=np.uint8)
    if d.ndim != 3 or d.shape[2] not in (3, 4):
        raise ValueError("writePng: need uint8 [H, W, 3|4]")
    _check(lib().pts_write_png(os.fsencode(path), d.ctypes.data, d.shape[1], d.shape[0], d.shape[2]), "writePng")


def writePpmTonemapped(path, rgba):
    d = _f32(rgba)
    _check(lib().pts_write_ppm_tonemapped(os.fsencode(path), _fp(d), d.shape[1], d.shape[0]), "writePpm")
