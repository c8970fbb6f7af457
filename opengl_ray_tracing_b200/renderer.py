"""ctypes mirror of libpt.so (include/pt.h): the CUDA path tracer behind the reference's draw loop.

``PathTracer`` plays the role of the reference's ``RenderPass pass1`` + ``display()``
(ImportanceSampling_LowDiscrepancySequence/main.cpp:76-133, 659-709): upload once, then per frame set the
camera uniforms and draw.  All compute is in the CUDA library; there is no CPU fallback — if libpt.so is
missing or no CUDA device is visible, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PT_LIBPT: load another BUILD of the same CUDA library (tools/build_variant.py: kernel A/B measurements)
_LIB_PATH = os.environ.get("PT_LIBPT") or os.path.join(_HERE, "libpt.so")
_lib = None

LAMBERT, DISNEY_UNIFORM, DISNEY_IS_MIS = 0, 1, 2
MODE_FAST, MODE_PARITY_REF, MODE_PARITY_WIDE, MODE_PARITY_FAST, MODE_FAST_COUNT = 0, 1, 2, 3, 4
DEFAULT_BOUNCES = {LAMBERT: 2, DISNEY_UNIFORM: 5, DISNEY_IS_MIS: 2}

ACC_MEAN, ACC_SUM = 0, 1
COMBINE_SUM, COMBINE_DISJOINT, COMBINE_DISPLAY, COMBINE_ASYNC, COMBINE_PIPELINED = 0, 1, 4, 8, 16
COMM_ID_BYTES = 128
MEGAKERNEL, WAVEFRONT, PERSISTENT = 0, 1, 2
ENV_TABLE, ENV_CDF_SEARCH = 0, 1
EXPORTS = ["pt_create", "pt_destroy", "pt_upload_scene", "pt_set_camera", "pt_set_region", "pt_set_band_interleave", "pt_set_tile_interleave", "pt_render_spp", "pt_render_frames",
           "pt_clear", "pt_readback", "pt_readback_async", "pt_readback_display", "pt_readback_display_async", "pt_readback_primary_ids", "pt_enable_cost_map", "pt_readback_cost_map", "pt_framebuffer_device_ptr", "pt_stream",
           "pt_scale_framebuffer", "pt_synchronize", "pt_get_stats", "pt_reset_stats", "pt_get_ray_histogram", "pt_calculate_hdr_cache", "pt_calculate_hdr_cache_noalloc", "pt_measure_l2_read_gbs", "pt_kernel_sass_hash", "pt_comm_get_unique_id", "pt_comm_init", "pt_comm_info", "pt_comm_destroy", "pt_combine", "pt_combine_fence", "pt_last_error",
           "pt_build_info"]


class PtConfig(C.Structure):
    _fields_ = [("variant", C.c_int), ("width", C.c_int), ("height", C.c_int), ("max_bounce", C.c_int),
                ("device", C.c_int), ("mode", C.c_int), ("pipeline", C.c_int), ("env_sampler", C.c_int)]


class PtStats(C.Structure):
    _fields_ = [("rays", C.c_ulonglong), ("samples", C.c_ulonglong), ("render_ms", C.c_double),
                ("launches", C.c_ulonglong), ("max_stack_depth", C.c_int), ("n_inner_nodes", C.c_int),
                ("n_materials", C.c_int), ("fast_tree_depth", C.c_int), ("n_fast_inner", C.c_int),
                ("upload_build_ms", C.c_double), ("inner_visits", C.c_ulonglong), ("tri_tests", C.c_ulonglong),
                ("leaf_candidates", C.c_ulonglong), ("tie_retraces", C.c_ulonglong), ("shaded_hits", C.c_ulonglong),
                ("env_texels", C.c_ulonglong)]


class PtError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise PtError("libpt.so is not built (python __graft_entry__.py build); there is no CPU fallback")
    L = C.CDLL(_LIB_PATH)
    vp, f32p = C.c_void_p, C.POINTER(C.c_float)
    sig = {
        "pt_create": (C.c_int, [C.POINTER(PtConfig), C.POINTER(vp)]),
        "pt_destroy": (C.c_int, [vp]),
        "pt_upload_scene": (C.c_int, [vp, f32p, C.c_int, f32p, C.c_int, f32p, f32p, C.c_int, C.c_int]),
        "pt_set_camera": (C.c_int, [vp, f32p, f32p]),
        "pt_set_region": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int]),
        "pt_set_band_interleave": (C.c_int, [vp, C.c_int, C.c_int]),
        "pt_set_tile_interleave": (C.c_int, [vp, C.c_int, C.c_int]),
        "pt_render_spp": (C.c_int, [vp, C.c_uint, C.c_uint]),
        "pt_render_frames": (C.c_int, [vp, C.c_uint, C.c_uint, C.c_uint, C.c_int]),
        "pt_clear": (C.c_int, [vp]),
        "pt_readback": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_readback_async": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_readback_primary_ids": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_readback_display": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_readback_display_async": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_enable_cost_map": (C.c_int, [vp, C.c_int]),
        "pt_readback_cost_map": (C.c_int, [vp, vp, C.c_size_t]),
        "pt_framebuffer_device_ptr": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "pt_stream": (C.c_int, [vp, C.POINTER(vp)]),
        "pt_scale_framebuffer": (C.c_int, [vp, C.c_float]),
        "pt_synchronize": (C.c_int, [vp]),
        "pt_get_stats": (C.c_int, [vp, C.POINTER(PtStats)]),
        "pt_reset_stats": (C.c_int, [vp]),
        "pt_get_ray_histogram": (C.c_int, [vp, C.POINTER(C.c_ulonglong), C.c_int]),
        "pt_last_error": (C.c_char_p, []),
        "pt_build_info": (C.c_char_p, []),
        "pt_kernel_sass_hash": (C.c_char_p, []),
        "pt_comm_get_unique_id": (C.c_int, [vp, C.c_size_t]),
        "pt_comm_init": (C.c_int, [vp, C.c_int, C.c_int, vp, C.c_size_t]),
        "pt_comm_info": (C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "pt_comm_destroy": (C.c_int, [vp]),
        "pt_combine": (C.c_int, [vp, C.c_int, C.c_uint, vp, vp]),
        "pt_combine_fence": (C.c_int, [vp]),
        "pt_calculate_hdr_cache": (C.c_int, [C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int, C.POINTER(C.c_float)]),
        "pt_calculate_hdr_cache_noalloc": (C.c_int, [C.c_int, vp, C.c_int, C.c_int, vp, vp]),
        "pt_measure_l2_read_gbs": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _check(rc, what):
    if rc != 0:
        raise PtError("%s failed (%d): %s" % (what, rc, lib().pt_last_error().decode("utf-8", "replace")))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


class PathTracer:
    """pass1 of the reference (shader variant ``variant``) on CUDA device ``device``."""

    def __init__(self, width, height, variant=DISNEY_IS_MIS, max_bounce=0, device=0, mode=MODE_FAST, pipeline=PERSISTENT,
                 env_sampler=ENV_TABLE):
        self.width, self.height, self.variant, self.mode = width, height, variant, mode
        self.pipeline = pipeline
        cfg = PtConfig(variant, width, height, max_bounce, device, mode, pipeline, env_sampler)
        h = C.c_void_p()
        _check(lib().pt_create(C.byref(cfg), C.byref(h)), "pt_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().pt_destroy(self._h)
            self._h = None

    __del__ = close

    def upload_scene(self, tris, nodes, hdr, hdrCache=None):
        """glBufferData x2 + glTexImage2D x2 (M:825-853).  tris [T,36], nodes [N,12], hdr/hdrCache [h,w,3]."""
        t, n, hd = _f32(tris).reshape(-1, 36), _f32(nodes).reshape(-1, 12), _f32(hdr)
        hc = _f32(hdrCache) if hdrCache is not None else None
        assert hd.ndim == 3 and hd.shape[2] == 3, "hdr must be [h, w, 3]"
        if hc is not None:
            assert hc.shape == hd.shape, "hdrCache must have the HDR map's shape"
        _check(lib().pt_upload_scene(self._h, _fp(t), len(t), _fp(n), len(n), _fp(hd), _fp(hc), hd.shape[1], hd.shape[0]),
               "pt_upload_scene")

    def set_camera(self, eye, cameraRotate):
        e, c = _f32(eye).reshape(3), _f32(cameraRotate).reshape(16)
        _check(lib().pt_set_camera(self._h, _fp(e), _fp(c)), "pt_set_camera")

    def set_region(self, x0, y0, x1, y1):
        _check(lib().pt_set_region(self._h, x0, y0, x1, y1), "pt_set_region")

    def set_band_interleave(self, world, rank):
        """This context renders 8-row band b iff b % world == rank (tile split in one launch)."""
        _check(lib().pt_set_band_interleave(self._h, world, rank), "pt_set_band_interleave")

    def set_tile_interleave(self, world, rank):
        """This context renders the 8x4 pixel tile (tx, strip) iff (tx + strip) % world == rank (tile split in one launch)."""
        _check(lib().pt_set_tile_interleave(self._h, world, rank), "pt_set_tile_interleave")

    def render_spp(self, first_frame, n_frames):
        """frameCounter = first_frame .. first_frame+n_frames-1; asynchronous."""
        _check(lib().pt_render_spp(self._h, first_frame, n_frames), "pt_render_spp")

    def render_frames(self, first_frame, n_frames, stride=1, accumulate=ACC_MEAN):
        """frameCounter = first_frame + k*stride; ACC_SUM adds colours instead of keeping the running mean."""
        _check(lib().pt_render_frames(self._h, first_frame, n_frames, stride, accumulate), "pt_render_frames")

    def clear(self):
        _check(lib().pt_clear(self._h), "pt_clear")

    def synchronize(self):
        _check(lib().pt_synchronize(self._h), "pt_synchronize")

    def readback(self, out=None) -> np.ndarray:
        """float32 [H, W, 4]; row 0 = bottom row."""
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        _check(lib().pt_readback(self._h, out.ctypes.data, out.nbytes), "pt_readback")
        return out

    def readback_into(self, host_ptr, nbytes):
        _check(lib().pt_readback(self._h, host_ptr, nbytes), "pt_readback")

    def readback_async(self, host_ptr, nbytes):
        """pt_readback_async: snapshot now, copy on a second stream; the (pinned) buffer is valid after synchronize()."""
        _check(lib().pt_readback_async(self._h, host_ptr, nbytes), "pt_readback_async")

    def readback_display_async(self, host_ptr, nbytes):
        """pt_readback_display_async: display pass now, copy on a second stream; the (pinned) buffer is valid after synchronize()."""
        _check(lib().pt_readback_display_async(self._h, host_ptr, nbytes), "pt_readback_display_async")

    def readback_display(self) -> np.ndarray:
        """pass3.fsh tone map + 8-bit conversion on the device: uint8 [H, W, 4], row 0 = top row."""
        out = np.empty((self.height, self.width, 4), np.uint8)
        _check(lib().pt_readback_display(self._h, out.ctypes.data, out.nbytes), "pt_readback_display")
        return out

    def readback_primary_ids(self) -> np.ndarray:
        out = np.empty((self.height, self.width), np.int32)
        _check(lib().pt_readback_primary_ids(self._h, out.ctypes.data, out.nbytes), "pt_readback_primary_ids")
        return out

    def enable_cost_map(self, on=True):
        _check(lib().pt_enable_cost_map(self._h, int(on)), "pt_enable_cost_map")

    def readback_cost_map(self) -> np.ndarray:
        """uint32 [H, W]: SM cycles each pixel's thread spent in the last launch."""
        out = np.empty((self.height, self.width), np.uint32)
        _check(lib().pt_readback_cost_map(self._h, out.ctypes.data, out.nbytes), "pt_readback_cost_map")
        return out

    def framebuffer_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(lib().pt_framebuffer_device_ptr(self._h, C.byref(p), C.byref(n)), "pt_framebuffer_device_ptr")
        return p.value, n.value

    def stream(self):
        s = C.c_void_p()
        _check(lib().pt_stream(self._h, C.byref(s)), "pt_stream")
        return s.value or 0

    def scale_framebuffer(self, scale):
        _check(lib().pt_scale_framebuffer(self._h, scale), "pt_scale_framebuffer")

    # ---- multi-GPU (pt_comm.cu) ----
    def comm_init(self, world, rank, unique_id: bytes):
        """Collective over the `world` contexts; unique_id = comm_unique_id() of rank 0, passed around by the caller."""
        buf = C.create_string_buffer(bytes(unique_id), COMM_ID_BYTES)
        _check(lib().pt_comm_init(self._h, world, rank, C.cast(buf, C.c_void_p), COMM_ID_BYTES), "pt_comm_init")

    def comm_info(self):
        w, r, p = C.c_int(), C.c_int(), C.c_int()
        _check(lib().pt_comm_info(self._h, C.byref(w), C.byref(r), C.byref(p)), "pt_comm_info")
        return {"world": w.value, "rank": r.value, "peer_memory": bool(p.value)}

    def combine(self, flags, total_frames=0, out_ptr=None, display_ptr=None):
        """Collective: all ranks' framebuffers -> one image on rank 0 (pt_combine).  flags = COMBINE_SUM | COMBINE_DISJOINT
        [| COMBINE_DISPLAY], the same on every rank; out_ptr / display_ptr: host addresses on rank 0, or None."""
        _check(lib().pt_combine(self._h, flags, total_frames, out_ptr, display_ptr), "pt_combine")

    def combine_fence(self):
        """The render stream waits (on the device) for every pipelined combine issued so far, on every rank."""
        _check(lib().pt_combine_fence(self._h), "pt_combine_fence")

    def combine_readback(self, kind, total_frames=0, display=False):
        """Rank 0: returns the combined float32 [H, W, 4] image (and the uint8 display image); other ranks: None."""
        flags = kind | (COMBINE_DISPLAY if display else 0)
        if self.comm_info()["rank"] != 0:
            self.combine(flags, total_frames)
            return None
        img = np.empty((self.height, self.width, 4), np.float32)
        disp = np.empty((self.height, self.width, 4), np.uint8) if display else None
        self.combine(flags, total_frames, img.ctypes.data, disp.ctypes.data if display else None)
        return (img, disp) if display else img

    def stats(self) -> dict:
        st = PtStats()
        _check(lib().pt_get_stats(self._h, C.byref(st)), "pt_get_stats")
        return {k: getattr(st, k) for k, _ in PtStats._fields_}

    def ray_histogram(self):
        """MODE_FAST_COUNT: bucket k = rays with 2^(k-1) <= traversal steps < 2^k."""
        b = (C.c_ulonglong * 33)()
        _check(lib().pt_get_ray_histogram(self._h, b, 33), "pt_get_ray_histogram")
        return list(b)

    def reset_stats(self):
        _check(lib().pt_reset_stats(self._h), "pt_reset_stats")


def build_info() -> str:
    return lib().pt_build_info().decode()


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(COMM_ID_BYTES)
    _check(lib().pt_comm_get_unique_id(C.cast(buf, C.c_void_p), COMM_ID_BYTES), "pt_comm_get_unique_id")
    return buf.raw


def kernel_sass_hash() -> str:
    return lib().pt_kernel_sass_hash().decode()


def measure_l2_read_gbs(device=0, window_mb=48, passes=64):
    """(L2 GB/s, HBM GB/s) of the ld.global.nc read probe (pt_probe.cu): the L2 roofline denominator for node/triangle fetches."""
    l2, hbm = C.c_double(), C.c_double()
    _check(lib().pt_measure_l2_read_gbs(device, window_mb, passes, C.byref(l2), C.byref(hbm)), "pt_measure_l2_read_gbs")
    return l2.value, hbm.value


def calculate_hdr_cache_noalloc(d_hdr_ptr, width, height, d_cache_ptr, stream=0, device=0):
    """pt_calculate_hdr_cache_noalloc: DEVICE pointers in and out (height*width*3 float32 each), static scratch inside libpt.so,
    no allocation; width <= 4096, width * height <= 262,144.  Asynchronous on `stream`."""
    _check(lib().pt_calculate_hdr_cache_noalloc(device, d_hdr_ptr, width, height, d_cache_ptr, stream or None), "pt_calculate_hdr_cache_noalloc")


def calculate_hdr_cache(hdr, device=0):
    """calculateHdrCache (M:555-652) on the GPU: hdr float32 [H, W, 3] -> table float32 [H, W, 3], bit for bit the host table."""
    h = np.ascontiguousarray(hdr, dtype=np.float32)
    if h.ndim != 3 or h.shape[2] != 3:
        raise ValueError("calculate_hdr_cache: need float32 [H, W, 3]")
    out = np.empty_like(h)
    f32p = C.POINTER(C.c_float)
    _check(lib().pt_calculate_hdr_cache(device, h.ctypes.data_as(f32p), h.shape[1], h.shape[0], out.ctypes.data_as(f32p)), "pt_calculate_hdr_cache")
    return out
