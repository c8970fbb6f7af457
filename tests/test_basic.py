"""BasicRayTracingWithC++/main.cpp on the GPU (pt_basic_*, SURVEY.md §8f.4 / BASELINE.json config 1).

Oracle: the reference's own source compiled with a reproducible random stream (oracle/_ref/libref_basic_det.so, built by
oracle/build_ref.py; edits listed in oracle/ref_basic_wrap.cpp).  tests/golden/basic.npz holds images that library
produced (tests/golden/make_golden_basic.py).  The bar is bit-identical float64 sum images."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden_basic import CASES, ref_basic_render  # noqa: E402

from opengl_ray_tracing_b200 import basic as B  # noqa: E402

GOLDEN = np.load(os.path.join(ROOT, "tests", "golden", "basic.npz"))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_basic_det.so")


def test_default_scene_is_the_reference_literal():
    shapes = B.default_scene()
    assert len(shapes) == 17
    assert [s.kind for s in shapes] == [1, 1, 1] + [0] * 14
    assert [round(s.radius, 6) for s in shapes[:3]] == [0.3, 0.4, 0.3]
    s1, s2, s3 = shapes[0], shapes[1], shapes[2]
    assert (s1.specularRate, s1.roughness) == (0.3, 0.1) and (s2.refractRate, s2.refractAngle) == (0.95, 0.1) and s3.specularRate == 0.3
    assert [i for i, s in enumerate(shapes) if s.isEmissive] == [5, 6]
    assert all(s.roughness == 1.0 and s.refractAngle == 1.0 for s in shapes[3:])
    assert tuple(shapes[3].color) == tuple(np.float32([1.0, 1.0, 0.1]))


@pytest.mark.skipif(not os.path.exists(REF_LIB), reason="oracle/_ref not built (needs /root/reference)")
def test_golden_images_come_from_the_reference_build():
    for w, h, n, seed in CASES:
        img = ref_basic_render(w, h, n, seed)
        assert img.tobytes() == GOLDEN["w%d_h%d_n%d_s%d" % (w, h, n, seed)].tobytes()
    assert ref_basic_render(16, 16, 2, 1).tobytes() != ref_basic_render(16, 16, 2, 2).tobytes()


def test_imshow_conversion():
    img = np.array([[[0.0, 1.0, 0.5], [4.0, 1e-3, 0.25]]], np.float64)
    want = np.clip(np.power(img, np.float64(np.float32(1.0) / np.float32(2.2))) * 255, 0, 255).astype(np.uint8)
    assert (B.imshow(img) == want).all()


@pytest.mark.gpu
def test_gpu_image_is_the_reference_image_bit_for_bit():
    shapes = B.default_scene()
    for w, h, n, seed in CASES:
        img, ms = B.render(shapes, w, h, n, seed)
        want = GOLDEN["w%d_h%d_n%d_s%d" % (w, h, n, seed)]
        bad = int((img != want).any(axis=2).sum())
        assert bad == 0, "%d of %d pixels differ (max |diff| %.3g)" % (bad, w * h, float(np.abs(img - want).max()))
        assert ms > 0.0


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(REF_LIB), reason="oracle/_ref not built")
def test_gpu_matches_the_reference_build_on_fresh_cases():
    shapes = B.default_scene()
    for w, h, n, seed in [(96, 64, 4, 99), (17, 9, 9, 3), (128, 128, 2, 2024)]:
        img, _ = B.render(shapes, w, h, n, seed)
        assert img.tobytes() == ref_basic_render(w, h, n, seed).tobytes()


@pytest.mark.gpu
def test_gpu_chunked_samples_and_errors():
    from opengl_ray_tracing_b200 import renderer as R
    shapes = B.default_scene()
    a, _ = B.render(shapes, 40, 30, 5, 11)
    b, _ = B.render(shapes, 40, 30, 5, 11)
    assert a.tobytes() == b.tobytes()                              # reproducible, unlike the reference's shared mt19937
    with pytest.raises(R.PtError):
        B.render(shapes, 0, 30, 5, 11)
    with pytest.raises(R.PtError):
        B.render(shapes, 8, 8, 1, 1, device=99)
    # the light (emissive, colour 1, not scaled by BRIGHTNESS: main.cpp:386-389) sums to `samples`; everything else is O(1)
    img, _ = B.render(shapes, 64, 64, 64, 5)
    assert img.max() == pytest.approx(64.0) and 0.3 < img.mean() < 5.0
