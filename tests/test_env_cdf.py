"""PT_ENV_CDF_SEARCH (second half): environment light samples by binary search of the marginal /
conditional CDFs instead of the reference's precomputed hdrCache table.  Not the reference's image by definition; pinned by
  * the search returning the reference table's entries at the table's own grid points (CPU, vs libptscene's
    calculateHdrCache, itself bit-exact against the reference's code),
  * the sampled texel distribution following the table's pdf channel, which sums to 1 within 1e-3 (CPU),
  * GPU: PT_MODE_PARITY_FAST bit-identical to the oracle's CDF mode, PT_MODE_FAST within 1e-3 rRMSE, and the converged
    image mean within 0.5 % of variant 2 with the reference's table."""
import numpy as np
import pytest

from helpers import oracle_render, rrmse, small_scene


def test_search_reproduces_the_reference_table_at_its_grid_points():
    from opengl_ray_tracing_b200 import scene as S
    from oracle import pyoracle as O
    for (w, h, seed) in [(128, 64, 1), (96, 40, 2), (256, 128, 3)]:
        hdr = S.synthHdr(w, h, seed)
        cache = S.calculateHdrCache(hdr, w, h)
        ii, jj = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        # table entry [i][j] was computed from xi_1 = i / H, xi_2 = j / W (IS/main.cpp:616-634) and is read by the lookup
        # (xi_1', xi_2') = (column / W, row / H): the search takes the same pair
        xi = np.stack([(jj.astype(np.float32) / np.float32(w)).ravel(), (ii.astype(np.float32) / np.float32(h)).ravel()], 1).astype(np.float32)
        xy = O.env_cdf_sample(hdr, xi)
        assert ((xy[:, 0].astype(np.float32) / np.float32(w)).reshape(h, w) == cache[..., 0]).all()
        assert ((xy[:, 1].astype(np.float32) / np.float32(h)).reshape(h, w) == cache[..., 1]).all()
        assert (O.env_pdf(hdr) == cache[..., 2]).all()


def test_pdf_sums_to_one_and_samples_follow_it():
    from opengl_ray_tracing_b200 import scene as S
    from oracle import pyoracle as O
    w, h = 128, 64
    hdr = S.synthHdr(w, h, 1)
    pdf = O.env_pdf(hdr).astype(np.float64)
    assert abs(pdf.sum() - 1.0) <= 1e-3
    rng = np.random.default_rng(3)
    n = 2_000_000
    xy = O.env_cdf_sample(hdr, rng.random((n, 2), dtype=np.float32))
    assert (xy[:, 0] >= 0).all() and (xy[:, 0] < w).all() and (xy[:, 1] >= 0).all() and (xy[:, 1] <= h).all()
    # 8 x 4 blocks of 16 x 16 texels: observed mass vs the pdf's mass, within 5 sigma of the binomial + 1e-3 for the table's drift
    y = np.minimum(xy[:, 1], h - 1)
    obs = np.zeros((4, 8))
    np.add.at(obs, (y // 16, xy[:, 0] // 16), 1.0)
    obs /= n
    want = pdf.reshape(4, 16, 8, 16).sum(axis=(1, 3))
    sigma = np.sqrt(want * (1 - want) / n)
    assert (np.abs(obs - want) <= 5 * sigma + 1e-3 * want + 1e-5).all()


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from opengl_ray_tracing_b200 import renderer
    return renderer


def gpu(R, sc, W, H, n, mode, env):
    pt = R.PathTracer(W, H, sc.variant, mode=mode, env_sampler=env)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, None if env else sc.hdrCache)
    pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(0, n)
    img, ids, st = pt.readback(), pt.readback_primary_ids(), pt.stats()
    pt.close()
    return img, ids, st


@pytest.mark.gpu
def test_cdf_mode_matches_its_oracle_and_converges_to_the_table_mode(R):
    sc = small_scene("is", "10k", hdr_res=(512, 256))
    W, H = 160, 96
    ref, ids_ref, _ = oracle_render(sc, W, H, 0, 3)
    from oracle import pyoracle as O
    cdf_ref, cdf_ids, _ = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, 3, math=1, env_sampler=1)
    assert cdf_ref.tobytes() != ref.tobytes()                # another sampler: another image at fixed spp
    img, ids, st = gpu(R, sc, W, H, 3, R.MODE_PARITY_FAST, R.ENV_CDF_SEARCH)
    assert (ids == cdf_ids).all() and img.tobytes() == cdf_ref.tobytes(), "PARITY_FAST + CDF search is not bit-identical to the oracle's CDF mode"
    img, ids, _ = gpu(R, sc, W, H, 3, R.MODE_FAST, R.ENV_CDF_SEARCH)
    assert (ids == cdf_ids).all() and rrmse(img, cdf_ref) <= 1e-3
    # converged mean: 512 spp of each sampler, production mode
    a, _, _ = gpu(R, sc, W, H, 512, R.MODE_FAST, R.ENV_TABLE)
    b, _, _ = gpu(R, sc, W, H, 512, R.MODE_FAST, R.ENV_CDF_SEARCH)
    ma, mb = float(a[..., :3].astype(np.float64).mean()), float(b[..., :3].astype(np.float64).mean())
    assert abs(ma - mb) / ma <= 5e-3, "converged means differ by %.3f %%" % (100 * abs(ma - mb) / ma)


@pytest.mark.gpu
def test_cdf_mode_is_rejected_where_it_does_not_exist(R):
    with pytest.raises(R.PtError, match="PT_ENV_CDF_SEARCH"):
        R.PathTracer(8, 8, 0, env_sampler=R.ENV_CDF_SEARCH)
    with pytest.raises(R.PtError, match="PT_ENV_CDF_SEARCH"):
        R.PathTracer(8, 8, 2, pipeline=R.WAVEFRONT, env_sampler=R.ENV_CDF_SEARCH)
    with pytest.raises(R.PtError, match="env_sampler"):
        R.PathTracer(8, 8, 2, env_sampler=7)
