"""The CPU oracle (oracle/ptoracle.cpp): pinned against the reference's own shader source compiled as C++
(oracle/_ref/, live where it exists), against golden framebuffers generated from it, and against the
known-answer values of the shader's RNG / Sobol code."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import have_ref
from helpers import GOLDEN_H, GOLDEN_W, KINDS, ROOT, from_bits, golden_scene, oracle_render, rrmse, small_scene
from oracle import pyoracle as O


def test_rand_known_answers(golden):
    for k in golden[0]["rand"]:
        _, r = O.wang_hash_seq(k["px"], k["py"], k["frame"], 8)
        assert r.view(np.uint32).tolist() == k["bits"], "rand() sequence differs from the shader's (pass1.fsh:73-89)"


def test_sobol_known_answers(golden):
    for k in golden[0]["sobol_cp"]:
        uv = O.sobol_vec2(k["i"], k["b"])
        assert uv.view(np.uint32).tolist() == k["bits"][:2], "sobolVec2 differs (pass1.fsh:101-116)"
        assert ((0.0 <= from_bits(k["bits"][2:])) & (from_bits(k["bits"][2:]) <= 1.0)).all()


def test_sobol_first_dimension_is_van_der_corput():
    for i in range(1, 64):
        g = i ^ (i >> 1)
        expect = int("{:032b}".format(g)[::-1], 2) / 2.0 ** 32
        assert abs(float(O.sobol_vec2(i, 0)[0]) - expect) < 1e-7


@pytest.mark.parametrize("kind", sorted(KINDS))
def test_golden_framebuffers(golden, kind):
    """Bit-identical to the reference shader (compiled as C++, ptmath transcendental set) after 1 and 4 frames."""
    sc = golden_scene(kind)
    imgs = golden[1]
    acc, ids, _ = oracle_render(sc, GOLDEN_W, GOLDEN_H, 0, 1)
    assert acc.tobytes() == imgs[kind + "_f1"].tobytes()
    acc, ids, _ = oracle_render(sc, GOLDEN_W, GOLDEN_H, 1, 3, accum=acc)
    assert acc.tobytes() == imgs[kind + "_f4"].tobytes()
    assert (ids >= 0).any() and (ids < 0).any()


@pytest.mark.parametrize("kind", sorted(KINDS))
@pytest.mark.parametrize("math", ["ptmath", "libm"])
def test_live_vs_reference_shader(kind, math):
    if not have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    sc = small_scene(kind, "10k", hdr_res=(256, 128))
    W, H = 80, 56
    a = O.ref_shader_render(sc.variant, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, 3, math=math)
    b, _, _ = oracle_render(sc, W, H, 0, 3, math=1 if math == "ptmath" else 0)
    assert a.tobytes() == b.tobytes(), "oracle restatement is not bit-identical to the reference shader source"


def test_libm_vs_ptmath_flavours_agree_statistically():
    """The two transcendental sets differ in the last bits only; converged means agree (north star: 0.5 %)."""
    sc = small_scene("is", "1k")
    a, _, _ = oracle_render(sc, 64, 40, 0, 32, math=0)
    b, _, _ = oracle_render(sc, 64, 40, 0, 32, math=1)
    assert abs(a[..., :3].mean() - b[..., :3].mean()) / b[..., :3].mean() < 5e-3


def test_region_and_frame_continuation():
    sc = small_scene("is", "1k")
    W, H = 40, 24
    full, ids, _ = oracle_render(sc, W, H, 0, 3)
    part = np.zeros_like(full)
    for reg in [(0, 0, 17, H), (17, 0, W, 9), (17, 9, W, H)]:
        part, _, _ = oracle_render(sc, W, H, 0, 3, accum=part, region=reg)
    assert part.tobytes() == full.tobytes()
    two, _, _ = oracle_render(sc, W, H, 0, 2)
    two, _, _ = oracle_render(sc, W, H, 2, 1, accum=two)
    assert two.tobytes() == full.tobytes()


@pytest.mark.parametrize("kind", sorted(KINDS))
def test_traversal_models_select_the_same_triangle(kind):
    """The culled wide-traversal model inside the oracle (what the CUDA kernels do) must pick the reference's hit."""
    sc = small_scene(kind, "10k", hdr_res=(64, 32))
    _, _, c = oracle_render(sc, 96, 64, 0, 2, counters=True)
    assert c["model_mismatch"] == 0 and c["rays"] > 0
    assert c["cull_inner"] <= c["ref_inner"] and c["cull_tris"] <= c["ref_tris"]


def test_leaf_box_characterisation_of_the_reference_hit():
    """Tree independence: the reference's closest hit = first strict minimum of t over the triangles whose
    reference LEAF box passes hitAABB > 0 (pt_build.cpp relies on this).  Checked by brute force over all leaves."""
    sc = small_scene("lambert", "1k")
    W, H = 24, 16
    _, ids, _ = oracle_render(sc, W, H, 0, 1)
    nodes, tris = sc.nodes, sc.tris
    leaves = nodes[1:][nodes[1:, 3] > 0]
    # primary rays exactly as the oracle builds them
    eye, cam = sc.eye, sc.cam.reshape(4, 4)          # cam[c] = column c
    mismatches = 0
    for py in range(H):
        for px in range(W):
            _, r = O.wang_hash_seq(px, py, 0, 2)
            f = np.float32
            pixx = f(f(f(f(px) + f(0.5)) / f(W)) * f(2)) - f(1)
            pixy = f(f(f(f(py) + f(0.5)) / f(H)) * f(2)) - f(1)
            vx, vy = f(pixx + f(f(r[0] - f(0.5)) / f(W))), f(pixy + f(f(r[1] - f(0.5)) / f(H)))
            d = (cam[0, :3] * vx + cam[1, :3] * vy) + cam[2, :3] * f(-1.5)
            d = (d * (f(1) / np.sqrt(f((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2])))).astype(np.float32)
            with np.errstate(divide="ignore", invalid="ignore"):
                inv = f(1) / d
                lo = (leaves[:, 6:9] - eye) * inv
                hi = (leaves[:, 9:12] - eye) * inv
            t0 = np.fmax.reduce(np.fmin(lo, hi), axis=1)
            t1 = np.fmin.reduce(np.fmax(lo, hi), axis=1)
            dd = np.where(t1 >= t0, np.where(t0 > 0, t0, t1), -1)
            best_t, best = np.inf, -1
            for leaf in leaves[dd > 0]:
                for i in range(int(leaf[4]), int(leaf[4] + leaf[3])):
                    p1, p2, p3 = tris[i, 0:3].astype(np.float64), tris[i, 3:6].astype(np.float64), tris[i, 6:9].astype(np.float64)
                    N = np.cross(p2 - p1, p3 - p1); N /= np.linalg.norm(N)
                    nd = N @ d
                    if abs(nd) < 1e-5: continue
                    t = (N @ p1 - eye @ N) / nd
                    if t < 0.0005: continue
                    P = eye + d * t
                    s = [np.cross(p2 - p1, P - p1) @ N, np.cross(p3 - p2, P - p2) @ N, np.cross(p1 - p3, P - p3) @ N]
                    if (min(s) > 0 or max(s) < 0) and t < best_t:
                        best_t, best = t, i
            mismatches += int(best != ids[py, px])
    assert mismatches <= 1      # float64 brute force vs float32 shader arithmetic may disagree on a razor-edge ray


def test_overflow_of_the_shader_stack_is_counted_not_hidden():
    sc = small_scene("is", "10k", hdr_res=(64, 32))
    _, _, c = oracle_render(sc, 64, 40, 0, 1, counters=True)
    assert c["over256"] == 0 and c["max_sp"] < 256 and c["wide_max_sp"] <= c["max_sp"]


def test_bad_arguments_are_rejected():
    sc = small_scene("is", "1k")
    with pytest.raises(RuntimeError):
        O.render(2, 5, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, 8, 8, sc.eye, sc.cam, 0, 1)     # > 4 bounces: 8 Sobol dims
    with pytest.raises(RuntimeError):
        O.render(2, 2, sc.tris, sc.nodes, sc.hdr, None, 8, 8, sc.eye, sc.cam, 0, 1)            # variant 2 needs the cache
