"""GPU parity tests proper: libpt.so (through the C ABI) against the oracle and the golden framebuffers.

Bit-exact where the arithmetic is defined to be (parity modes, primary-hit ids in every mode); production (fast)
mode within the north star's tolerance: relative RMSE <= 1e-3 at fixed spp, converged mean within 0.5 %."""
import ctypes as C

import numpy as np
import pytest

from helpers import GOLDEN_H, GOLDEN_W, KINDS, golden_scene, oracle_render, rrmse, small_scene

pytestmark = pytest.mark.gpu

RRMSE_TOL = 1e-3      # north star: per-pixel radiance within 1e-3 relative RMSE at fixed spp
MEAN_TOL = 5e-3       # north star: converged image mean within 0.5 %


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from opengl_ray_tracing_b200 import renderer
    return renderer


def render(R, sc, W, H, first, n, mode, region=None, max_bounce=0, pt=None, close=True, pipeline=None):
    if pt is None:
        pt = R.PathTracer(W, H, sc.variant, max_bounce=max_bounce, mode=mode,
                          pipeline=R.PERSISTENT if pipeline is None else pipeline)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
        pt.set_camera(sc.eye, sc.cam)
    if region:
        pt.set_region(*region)
    pt.render_spp(first, n)
    img, ids = pt.readback(), pt.readback_primary_ids()
    st = pt.stats()
    if close:
        pt.close()
    return img, ids, st, pt


@pytest.mark.parametrize("kind", sorted(KINDS))
def test_golden_framebuffers_all_parity_modes(R, golden, kind):
    """CUDA vs framebuffers produced by the reference's shader source (tests/golden/images.npz): bit-identical."""
    sc = golden_scene(kind)
    for mode in (R.MODE_PARITY_REF, R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST):
        img, _, _, pt = render(R, sc, GOLDEN_W, GOLDEN_H, 0, 1, mode, close=False)
        assert img.tobytes() == golden[1][kind + "_f1"].tobytes(), "mode %d, frame 1" % mode
        img, _, _, _ = render(R, sc, GOLDEN_W, GOLDEN_H, 1, 3, mode, pt=pt)
        assert img.tobytes() == golden[1][kind + "_f4"].tobytes(), "mode %d, frames 1-4" % mode
    img, _, _, _ = render(R, sc, GOLDEN_W, GOLDEN_H, 0, 4, R.MODE_FAST)
    assert rrmse(img, golden[1][kind + "_f4"]) <= RRMSE_TOL


@pytest.mark.parametrize("kind", sorted(KINDS))
def test_vs_oracle_10k(R, kind):
    sc = small_scene(kind, "10k", hdr_res=(512, 256))
    W, H, NF = 200, 120, 3                                   # ragged: not a multiple of the 32x8 CTA tile
    ref, ids_ref, _ = oracle_render(sc, W, H, 0, NF)
    for mode in (R.MODE_PARITY_REF, R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST):
        img, ids, st, _ = render(R, sc, W, H, 0, NF, mode)
        assert img.tobytes() == ref.tobytes(), "mode %d not bit-identical to the oracle" % mode
        assert (ids == ids_ref).all() and st["rays"] > W * H * NF
    img, ids, st, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    assert (ids == ids_ref).all(), "primary-hit triangle ids must be bit-exact in the production mode"
    assert rrmse(img, ref) <= RRMSE_TOL


def test_metric_scene_1M_triangles(R):
    """The metric scene: the reference builder's tree is thousands of levels deep here (the shader's stack[256]
    would be overrun); the kernels traverse their own topology and must still match the oracle bit for bit."""
    sc = small_scene("is", "1M", hdr_res=(1024, 512), camera=(0.0, 0.0, 4.0))
    W, H = 480, 270
    ref, ids_ref, _ = oracle_render(sc, W, H, 0, 2)
    img, ids, st, _ = render(R, sc, W, H, 0, 2, R.MODE_PARITY_FAST)
    assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all()
    assert st["max_stack_depth"] > 256 and st["fast_tree_depth"] < 64
    img, ids, _, _ = render(R, sc, W, H, 0, 2, R.MODE_FAST)
    assert (ids == ids_ref).all() and rrmse(img, ref) <= RRMSE_TOL


def test_primary_ids_full_1080p(R):
    """BASELINE.json's resolution: every one of the 2,073,600 primary-hit ids equals the oracle's."""
    sc = small_scene("lambert", "100k", hdr_res=(256, 128), camera=(0.0, 0.0, 4.0))
    W, H = 1920, 1080
    _, ids_ref, _ = oracle_render(sc, W, H, 0, 1)
    for mode in (R.MODE_FAST, R.MODE_PARITY_WIDE):
        _, ids, _, _ = render(R, sc, W, H, 0, 1, mode)
        assert (ids == ids_ref).all()
    assert (ids_ref >= 0).mean() > 0.2


def test_converged_mean_and_determinism(R):
    sc = small_scene("is", "10k", hdr_res=(512, 256))
    W, H, NF = 160, 96, 64
    a, _, _, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    b, _, _, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    assert a.tobytes() == b.tobytes(), "two runs must give the same bits"
    p, _, _, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST)
    assert abs(a[..., :3].mean() - p[..., :3].mean()) / p[..., :3].mean() <= MEAN_TOL
    assert rrmse(a, p) <= RRMSE_TOL


def test_frame_continuation_and_clear(R):
    sc = small_scene("disney", "1k")
    W, H = 96, 64
    full, _, _, _ = render(R, sc, W, H, 0, 6, R.MODE_PARITY_FAST)
    img, _, _, pt = render(R, sc, W, H, 0, 2, R.MODE_PARITY_FAST, close=False)
    img, _, _, _ = render(R, sc, W, H, 2, 4, R.MODE_PARITY_FAST, pt=pt, close=False)
    assert img.tobytes() == full.tobytes()
    pt.clear()
    img, _, _, _ = render(R, sc, W, H, 0, 6, R.MODE_PARITY_FAST, pt=pt)
    assert img.tobytes() == full.tobytes()


def test_regions_and_band_interleave_tile_the_image(R):
    sc = small_scene("is", "1k")
    W, H, NF = 100, 52, 2
    full, ids_full, _, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    for reg in [(0, 0, 37, H), (37, 0, W, 19), (37, 19, W, H)]:
        pt.set_region(*reg); pt.render_spp(0, NF)
    assert pt.readback().tobytes() == full.tobytes()
    pt.close()
    parts = []
    for rank in range(3):                                    # 3 "GPUs", interleaved 8-row bands, one launch each
        q = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
        q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
        q.set_band_interleave(3, rank); q.render_spp(0, NF)
        parts.append(q.readback()); q.close()
    assert (parts[0] + parts[1] + parts[2]).tobytes() == full.tobytes()      # disjoint: x + 0 is exact
    from opengl_ray_tracing_b200 import dist as D
    for world in (2, 5):                                     # diagonal 8x4 tiles: disjoint, complete, where TilePlan says
        parts, px = [], 0
        for rank in range(world):
            q = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
            q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
            q.set_tile_interleave(world, rank); q.render_spp(0, NF)
            img = q.readback(); px += q.stats()["samples"]; q.close()
            own = np.zeros((H, W), bool)
            for x0, y0, x1, y1 in D.TilePlan(W, H, world, rank).tiles():
                own[y0:y1, x0:x1] = True
            assert (img[~own] == 0).all() and (img[own][:, 3] == 1).all()
            parts.append(img)
        assert sum(parts[1:], parts[0]).tobytes() == full.tobytes()
        assert px == W * H * NF


def test_sample_split_sum_matches_running_mean(R):
    sc = small_scene("is", "1k")
    W, H, NF = 64, 40, 8
    full, _, _, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    acc = np.zeros_like(full)
    for rank in range(2):
        q = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
        q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
        q.render_frames(rank, NF // 2, 2, R.ACC_SUM)
        acc += q.readback(); q.close()
    acc[..., :3] /= NF
    assert rrmse(acc, full) < 1e-5


def test_single_leaf_scene_and_one_pixel(R):
    from opengl_ray_tracing_b200 import scene as S, scenes
    sc0 = S.Scene()
    sc0.addQuad(S.Material(baseColor=(0.8, 0.3, 0.2)), S.getTransformMatrix((40, 0, 0), (0, 0, 0), (3, 1, 3)))
    sc0.buildBVHwithSAH(8)                                    # 2 triangles -> the root is itself a leaf
    assert sc0.nNodes == 2
    eye, cam = S.camera()
    hdr = S.synthHdr(64, 32, 1)
    sc = scenes.BuiltScene("quad", 0, 2, sc0.triangles_encoded(), sc0.nodes_encoded(), hdr, None, eye, cam)
    for (W, H) in [(33, 9), (1, 1)]:
        ref, ids_ref, _ = oracle_render(sc, W, H, 0, 2)
        for mode in (R.MODE_PARITY_REF, R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST):
            img, ids, _, _ = render(R, sc, W, H, 0, 2, mode)
            assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all()


@pytest.mark.parametrize("bounces", [1, 3, 4])
def test_max_bounce_override(R, bounces):
    sc = small_scene("is", "1k")
    W, H = 64, 40
    ref, _, _ = oracle_render(sc, W, H, 0, 2, max_bounce=bounces)
    img, _, _, _ = render(R, sc, W, H, 0, 2, R.MODE_PARITY_FAST, max_bounce=bounces)
    assert img.tobytes() == ref.tobytes()


def test_error_behaviour(R):
    L = R.lib()
    pt = R.PathTracer(32, 16, 2, mode=R.MODE_FAST)
    with pytest.raises(R.PtError, match="no scene"):
        pt.render_spp(0, 1)
    sc = small_scene("is", "1k")
    assert L.pt_calculate_hdr_cache(0, None, 4, 4, None) < 0 and b"null buffer" in L.pt_last_error()
    bad = sc.nodes.copy(); bad[1, 0] = 1.0                    # root's left child = itself
    with pytest.raises(R.PtError, match="child index"):
        pt.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache)
    bad = sc.nodes.copy(); leaf = np.where(bad[:, 3] > 0)[0][1]; bad[leaf, 4] = len(sc.tris)
    with pytest.raises(R.PtError, match="leaf range"):
        pt.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache)
    # what the production traversal relies on is checked at upload
    bad = sc.nodes.copy(); bad[leaf, 3] = -1.0
    with pytest.raises(R.PtError, match="negative triangle count"):
        pt.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache)
    leaves = np.where(sc.nodes[:, 3] > 0)[0]
    bad = sc.nodes.copy(); bad[leaves[2], 4] = bad[leaves[1], 4]          # two leaves claim the same triangles
    with pytest.raises(R.PtError, match="two leaves|no leaf"):
        pt.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache)
    inner = np.where(sc.nodes[2:, 3] == 0)[0][0] + 2
    bad = sc.nodes.copy(); bad[inner, 9:12] -= 0.25 * (bad[inner, 9:12] - bad[inner, 6:9])    # a parent box that no longer holds its children
    with pytest.raises(R.PtError, match="not inside its parent"):
        pt.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache)
    q = R.PathTracer(32, 16, 2, mode=R.MODE_PARITY_WIDE)               # the reference-topology modes take the array as it is
    q.upload_scene(sc.tris, bad, sc.hdr, sc.hdrCache); q.close()
    with pytest.raises(R.PtError):
        pt.set_region(0, 0, 33, 16)
    with pytest.raises(R.PtError):
        pt.set_band_interleave(2, 2)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
    small = np.empty(8, np.float32)
    assert L.pt_readback(pt._h, small.ctypes.data, small.nbytes) == -1
    with pytest.raises(R.PtError, match="max_bounce"):
        R.PathTracer(8, 8, 2, max_bounce=5)
    pt.close()


# ---- the three schedulings of the same arithmetic: megakernel, wavefront, persistent -------------------------------
PIPELINES = [0, 1, 2]


@pytest.mark.parametrize("kind", sorted(KINDS))
@pytest.mark.parametrize("pipeline", PIPELINES)
def test_every_pipeline_is_bit_identical_to_the_oracle(R, kind, pipeline):
    sc = small_scene(kind, "10k", hdr_res=(512, 256), camera=(10.0, 25.0, 2.5))
    W, H, NF = 210, 77, 3                                    # ragged against 8x4 warp tiles and 32x8 CTA tiles
    ref, ids_ref, _ = oracle_render(sc, W, H, 0, NF)
    for mode in (R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST):
        img, ids, st, _ = render(R, sc, W, H, 0, NF, mode, pipeline=pipeline)
        assert img.tobytes() == ref.tobytes(), "pipeline %d mode %d not bit-identical to the oracle" % (pipeline, mode)
        assert (ids == ids_ref).all()
    img, ids, st, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST, pipeline=pipeline)
    assert (ids == ids_ref).all() and rrmse(img, ref) <= RRMSE_TOL
    assert st["rays"] > W * H * NF and st["launches"] >= 1


@pytest.mark.parametrize("pipeline", PIPELINES)
def test_pipelines_continue_frames_split_regions_and_sum(R, pipeline):
    sc = small_scene("is", "1k")
    W, H, NF = 100, 52, 4
    full, _, _, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST, pipeline=0)
    # frame continuation in two calls
    img, _, _, pt = render(R, sc, W, H, 0, 1, R.MODE_PARITY_FAST, pipeline=pipeline, close=False)
    img, _, _, _ = render(R, sc, W, H, 1, NF - 1, R.MODE_PARITY_FAST, pt=pt)
    assert img.tobytes() == full.tobytes()
    # three rectangles tile the image
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_PARITY_FAST, pipeline=pipeline)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    for reg in [(0, 0, 37, H), (37, 0, W, 19), (37, 19, W, H)]:
        pt.set_region(*reg); pt.render_spp(0, NF)
    assert pt.readback().tobytes() == full.tobytes()
    pt.close()
    # interleaved bands over 3 ranks
    parts = []
    for rank in range(3):
        q = R.PathTracer(W, H, sc.variant, mode=R.MODE_PARITY_FAST, pipeline=pipeline)
        q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
        q.set_band_interleave(3, rank); q.render_spp(0, NF)
        parts.append(q.readback()); q.close()
    assert (parts[0] + parts[1] + parts[2]).tobytes() == full.tobytes()
    # interleaved 8x4 tiles over 3 ranks
    parts = []
    for rank in range(3):
        q = R.PathTracer(W, H, sc.variant, mode=R.MODE_PARITY_FAST, pipeline=pipeline)
        q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
        q.set_tile_interleave(3, rank); q.render_spp(0, NF)
        parts.append(q.readback()); q.close()
    assert (parts[0] + parts[1] + parts[2]).tobytes() == full.tobytes()
    # sample split: strided frames as plain sums
    acc = np.zeros_like(full)
    for rank in range(2):
        q = R.PathTracer(W, H, sc.variant, mode=R.MODE_PARITY_FAST, pipeline=pipeline)
        q.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); q.set_camera(sc.eye, sc.cam)
        q.render_frames(rank, NF // 2, 2, R.ACC_SUM)
        acc += q.readback(); q.close()
    acc[..., :3] /= NF
    assert rrmse(acc, full) < 1e-5


def test_metric_scene_every_pipeline(R):
    """is-1M (the metric scene) at a quarter of 1080p: all pipelines give the oracle's bits under the parity arithmetic."""
    sc = small_scene("is", "1M", hdr_res=(1024, 512), camera=(0.0, 0.0, 4.0))
    W, H = 480, 270
    ref, ids_ref, _ = oracle_render(sc, W, H, 0, 2)
    for pipeline in PIPELINES:
        img, ids, _, _ = render(R, sc, W, H, 0, 2, R.MODE_PARITY_FAST, pipeline=pipeline)
        assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all(), "pipeline %d" % pipeline
    # a close-up (camera inside the reference tree's degenerate region: the mesh fills the image)
    from opengl_ray_tracing_b200 import scene as S
    eye, cam = S.camera(20.0, 35.0, 1.2)
    ref, ids_ref, _ = oracle_render(type(sc)(sc.name, sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, eye, cam), 160, 90, 0, 1)
    pt = R.PathTracer(160, 90, sc.variant, mode=R.MODE_PARITY_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(eye, cam); pt.render_spp(0, 1)
    assert pt.readback().tobytes() == ref.tobytes() and (pt.readback_primary_ids() == ids_ref).all()
    assert (ids_ref >= 0).mean() > 0.5
    pt.close()


def test_wavefront_waves_and_lane_refill(R, monkeypatch):
    """The wavefront pipeline sizes its waves from the path-slot capacity: with 1 Mi slots a 512x512 image starts at 4 frames
    per wave, then grows to what the survivors of the first wave allow; 24 frames are several waves.  Bit-identical to the
    megakernel, also as two launches, also when an over-optimistic fill target forces waves to be repeated."""
    monkeypatch.setenv("PT_WF_PATHS_M", "1")
    for kind in ("is", "disney"):
        sc = small_scene(kind, "10k", hdr_res=(256, 128), camera=(10.0, 25.0, 2.5))
        W = H = 512
        NF = 24
        ref, ids_ref, st_ref, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST, pipeline=R.MEGAKERNEL)
        img, ids, st, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST, pipeline=R.WAVEFRONT)
        assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all() and st["rays"] == st_ref["rays"]
        assert st["launches"] >= 2 * (2 + 3 * sc.max_bounce)
        img, _, _, pt = render(R, sc, W, H, 0, 17, R.MODE_PARITY_FAST, pipeline=R.WAVEFRONT, close=False)
        img, ids, _, _ = render(R, sc, W, H, 17, NF - 17, R.MODE_PARITY_FAST, pt=pt)
        assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all()
        fast, ids_f, _, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST, pipeline=R.WAVEFRONT)
        assert (ids_f == ids_ref).all() and rrmse(fast, ref) <= RRMSE_TOL
        monkeypatch.setenv("PT_WF_FILL_PCT", "400")            # the host aims at 4x the slots there are: waves overflow and are repeated
        img, ids, st, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST, pipeline=R.WAVEFRONT)
        monkeypatch.delenv("PT_WF_FILL_PCT")
        assert img.tobytes() == ref.tobytes() and (ids == ids_ref).all() and st["rays"] == st_ref["rays"]


def test_readback_async_is_a_snapshot(R):
    """pt_readback_async copies the image AS OF THE CALL while later launches already overwrite the framebuffer."""
    import torch
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H = 256, 144
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(0, 4)
    want4 = pt.readback().copy()
    pt.clear()
    host = [torch.empty((H, W, 4), dtype=torch.float32).pin_memory() for _ in range(2)]
    pt.render_spp(0, 4)
    pt.readback_async(host[0].data_ptr(), host[0].numel() * 4)
    pt.render_spp(4, 60)                                       # overwrites the framebuffer while the copy may still run
    pt.readback_async(host[1].data_ptr(), host[1].numel() * 4)   # a second copy queues behind the first
    pt.render_spp(64, 8)
    pt.synchronize()
    assert host[0].numpy().tobytes() == want4.tobytes()
    pt2 = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    pt2.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt2.set_camera(sc.eye, sc.cam)
    pt2.render_spp(0, 64)
    assert host[1].numpy().tobytes() == pt2.readback().tobytes()
    pt2.close(); pt.close()


def test_readback_display_async_is_a_snapshot(R):
    """pt_readback_display_async: the display image AS OF THE CALL, byte for byte what pt_readback_display returns, while later
    launches (and a later synchronous display readback) go on."""
    import torch
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H = 256, 144
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    want = []
    for k in range(3):
        pt.render_spp(k, 1)
        want.append(pt.readback_display().copy())
    pt.clear()
    host = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(3)]
    for k in range(3):                                         # the display() loop, one sample per frame, copies overlapped
        pt.render_spp(k, 1)
        pt.readback_display_async(host[k].data_ptr(), host[k].numel())
    pt.render_spp(3, 29)                                       # overwrites the framebuffer while the last copy may still run
    later = pt.readback_display()                              # the synchronous form queues behind the copy in flight
    pt.synchronize()
    for k in range(3):
        assert host[k].numpy().tobytes() == want[k].tobytes(), "frame %d" % k
    assert not (later == want[2]).all()
    with pytest.raises(R.PtError, match="too small"):
        pt.readback_display_async(host[0].data_ptr(), 16)
    pt.close()


def test_counting_build_and_cost_map(R):
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H, NF = 128, 72, 2
    a, _, sa, _ = render(R, sc, W, H, 0, NF, R.MODE_FAST)
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST_COUNT)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    pt.enable_cost_map(True)
    pt.render_spp(0, NF)
    b, st = pt.readback(), pt.stats()
    assert a.tobytes() == b.tobytes(), "the counting build must not change the image"
    assert st["rays"] == sa["rays"] and st["inner_visits"] > 0 and st["tri_tests"] > 0
    assert 0 < st["shaded_hits"] <= st["rays"] and st["env_texels"] > 0 and st["leaf_candidates"] <= st["tri_tests"]
    assert sum(pt.ray_histogram()) == st["rays"]
    cost = pt.readback_cost_map()
    assert cost.shape == (H, W) and (cost > 0).all()
    pt.close()
    assert sa["inner_visits"] == 0                           # the production build carries no counters


def test_wavefront_rejects_what_it_cannot_run(R):
    with pytest.raises(R.PtError, match="megakernel only"):
        R.PathTracer(8, 8, 0, mode=R.MODE_PARITY_REF, pipeline=R.WAVEFRONT)
    with pytest.raises(R.PtError, match="max_bounce"):
        R.PathTracer(8, 8, 1, max_bounce=9, pipeline=R.WAVEFRONT)


def test_display_pass_matches_pass3(R):
    """pt_readback_display against the reference's OWN pass3.fsh (compiled as C++ by oracle/build_ref.py; its output on the
    golden framebuffers is committed as tests/golden/pass3.npz by make_golden_pass3.py): identical bytes — on the golden
    scenes rendered here (whose framebuffers equal the golden ones bit for bit), and on a synthetic framebuffer with zeros,
    denormals, values far above 1, inf and NaN written straight into the device framebuffer."""
    import os
    import torch
    from helpers import ROOT
    from opengl_ray_tracing_b200 import dist as D
    g = np.load(os.path.join(ROOT, "tests", "golden", "pass3.npz"))
    imgs = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    for kind in sorted(KINDS):
        sc = golden_scene(kind)
        pt = R.PathTracer(GOLDEN_W, GOLDEN_H, sc.variant, mode=R.MODE_PARITY_FAST)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
        pt.render_spp(0, 4)
        img, disp = pt.readback(), pt.readback_display()
        pt.close()
        assert img.tobytes() == imgs[kind + "_f4"].tobytes()
        assert disp.shape == (GOLDEN_H, GOLDEN_W, 4) and disp.tobytes() == g[kind + "_display"].tobytes(), kind
    fb = g["special_fb"]
    H, W = fb.shape[:2]
    pt = R.PathTracer(W, H, 0, mode=R.MODE_FAST)
    dev = D.DeviceFramebuffer(pt).tensor(torch.device("cuda", 0))
    dev.copy_(torch.from_numpy(fb))
    torch.cuda.synchronize()
    disp = pt.readback_display()
    del dev
    pt.close()
    assert disp.tobytes() == g["special_display"].tobytes()
    assert g["special_display"][..., :3].std() > 0


@pytest.mark.gpu
def test_hdr_cache_on_the_gpu_is_the_host_table_bit_for_bit(R):
    """calculateHdrCache (M:555-652) on the device vs libptscene's host restatement (itself pinned to the reference's code)."""
    from opengl_ray_tracing_b200 import scene as S
    rng = np.random.default_rng(5)
    maps = [S.synthHdr(512, 256, 1), S.synthHdr(2048, 1024, 3), S.synthHdr(96, 40, 2),
            (rng.random((33, 70, 3), dtype=np.float32) ** 8 * 50).astype(np.float32),           # ragged size, spiky
            np.ones((8, 8, 3), np.float32)]                                                     # flat: long runs of equal cdf steps
    for hdr in maps:
        want = S.calculateHdrCache(hdr, hdr.shape[1], hdr.shape[0])
        got = R.calculate_hdr_cache(hdr)
        assert got.shape == want.shape and got.tobytes() == np.ascontiguousarray(want, np.float32).tobytes()


@pytest.mark.gpu
def test_upload_without_a_table_builds_it_on_the_device(R):
    sc = small_scene("is", "1k")
    W, H, NF = 96, 64, 3
    a, ids_a, _, _ = render(R, sc, W, H, 0, NF, R.MODE_PARITY_FAST)
    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_PARITY_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, None)           # variant 2 with hdrCache = NULL
    pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(0, NF)
    b = pt.readback()
    pt.close()
    assert a.tobytes() == b.tobytes()
