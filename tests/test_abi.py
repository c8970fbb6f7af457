"""The C-ABI libraries load on a machine without a GPU and export every symbol include/*.h declares."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from helpers import ROOT
from opengl_ray_tracing_b200 import renderer as R, scene as S


def declared(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(%s\w+)\s*\(" % prefix, text)))


def test_libpt_exports_every_declared_symbol():
    names = declared("pt.h", "pt_")
    assert len(names) >= 18
    L = C.CDLL(os.path.join(ROOT, "opengl_ray_tracing_b200", "libpt.so"))
    for n in names:
        assert hasattr(L, n), "libpt.so lacks " + n
    from opengl_ray_tracing_b200 import basic as B
    basic = sorted(n for n in names if n.startswith("pt_basic_"))          # bound by basic.py
    assert basic == ["pt_basic_default_scene", "pt_basic_imshow", "pt_basic_last_error", "pt_basic_render"]
    B._lib()
    assert sorted(R.EXPORTS + basic) == names, "renderer.py / basic.py binding lists are out of sync with include/pt.h"
    assert C.sizeof(B.PtBasicShape) == 104


def test_libptscene_exports_every_declared_symbol():
    names = declared("ptscene.h", "pts_")
    L = C.CDLL(os.path.join(ROOT, "opengl_ray_tracing_b200", "libptscene.so"))
    for n in names:
        assert hasattr(L, n), "libptscene.so lacks " + n
    assert sorted(S.EXPORTS) == names


def test_build_info_and_struct_layouts():
    info = R.build_info()
    assert "sm_100a" in info
    assert C.sizeof(R.PtConfig) == 32 and C.sizeof(R.PtStats) == 112 and C.sizeof(S.PtsMaterial) == 72
    # the cubin inside libpt.so really targets sm_100a
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", os.path.join(ROOT, "opengl_ray_tracing_b200", "libpt.so")],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    assert "sm_100a" in out


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(R.PtError):
        R.PathTracer(16, 16, 0)


def test_argument_validation_needs_no_gpu():
    L = R.lib()
    h = C.c_void_p()
    for bad in (R.PtConfig(7, 16, 16, 0, 0, 0, 0), R.PtConfig(0, 0, 16, 0, 0, 0, 0), R.PtConfig(0, 16, 16, 0, 0, 9, 0),
                R.PtConfig(2, 16, 16, 5, 0, 0, 0), R.PtConfig(0, 16, 16, 0, 0, 0, 7), R.PtConfig(0, 16, 16, 0, 0, 1, 1),
                R.PtConfig(1, 16, 16, 9, 0, 0, 1)):
        assert L.pt_create(C.byref(bad), C.byref(h)) == -1          # PT_EINVAL before any CUDA call
        assert len(L.pt_last_error()) > 0
    assert L.pt_create(None, C.byref(h)) == -1
    assert L.pt_render_spp(None, 0, 1) == -1 and L.pt_readback(None, None, 0) == -1
