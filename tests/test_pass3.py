"""The display pass's golden bytes (tests/golden/pass3.npz) are what the reference's pass3.fsh produces, and the float32
restatement of its one formula (toneMapping(c, 1.5), pass3.fsh:14-24) agrees with it."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from conftest import have_ref
from helpers import ROOT

sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def restated(fb):
    f = np.float32
    c = fb[::-1, :, :3].astype(f)
    with np.errstate(all="ignore"):
        lum = (f(0.3) * c[..., 0] + f(0.6) * c[..., 1]) + f(0.1) * c[..., 2]
        k = f(1.0) + lum / f(1.5)
        v = (c * f(1.0)) / k[..., None]
    v = np.where(np.isnan(v), f(0), v)
    q = (np.clip(v, f(0), f(1)) * f(255.0) + f(0.5)).astype(np.int32).astype(np.uint8)
    return np.concatenate([q, np.full(q.shape[:2] + (1,), 255, np.uint8)], axis=2)


def test_restatement_equals_golden_display_bytes():
    g = np.load(os.path.join(ROOT, "tests", "golden", "pass3.npz"))
    imgs = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    for kind in ("lambert", "disney", "is"):
        assert restated(imgs[kind + "_f4"]).tobytes() == g[kind + "_display"].tobytes()
    assert restated(g["special_fb"]).tobytes() == g["special_display"].tobytes()


@pytest.mark.skipif(not have_ref() or not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_pass3.so")),
                    reason="oracle/_ref (the reference compiled in the container that has /root/reference) is not built")
def test_golden_display_bytes_come_from_the_reference_shader():
    import make_golden_pass3 as M
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_pass3.so"))
    g = np.load(os.path.join(ROOT, "tests", "golden", "pass3.npz"))
    imgs = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    for kind in ("lambert", "disney", "is"):
        assert M.pass3_display(L, imgs[kind + "_f4"]).tobytes() == g[kind + "_display"].tobytes()
    assert M.pass3_display(L, g["special_fb"]).tobytes() == g["special_display"].tobytes()
    assert M.special_framebuffer().tobytes() == g["special_fb"].tobytes()
