"""Multi-GPU host logic on CPU: world_size 2 over gloo.  The planning (interleaved bands / strided frames) and the
combine (torch.distributed reduce) are the product's code (opengl_ray_tracing_b200/dist.py); the renderer is a
stand-in that draws with the oracle, so the test needs no GPU."""
import os
import socket

import numpy as np
import pytest

from helpers import oracle_render, rrmse, small_scene
from opengl_ray_tracing_b200 import dist as D

W, H, NF = 40, 28, 4


class OracleRenderer:
    """Same surface as renderer.PathTracer, minus set_tile_interleave (so render_tiles issues one call per rectangle)."""

    def __init__(self, sc):
        self.sc, self.width, self.height = sc, W, H
        self.fb = np.zeros((H, W, 4), np.float32)
        self.region = (0, 0, W, H)

    def set_region(self, x0, y0, x1, y1):
        self.region = (x0, y0, x1, y1)

    def render_spp(self, first, n):
        self.fb, _, _ = oracle_render(self.sc, W, H, first, n, accum=self.fb, region=self.region)

    def render_frames(self, first, n, stride, accumulate):
        for k in range(n):                       # ACC_SUM: add each frame's colour
            one, _, _ = oracle_render(self.sc, W, H, first + k * stride, 1, region=self.region)
            one[..., :3] *= (first + k * stride + 1)          # undo mix(0, c, 1/(frame+1)) -> c
            self.fb[..., :3] += one[..., :3]
            self.fb[..., 3] = 1.0


def _worker(rank, world, port, split, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = small_scene("is", "1k")
    pt = OracleRenderer(sc)
    if split.startswith("tile"):
        D.render_tiles(pt, D.TilePlan(W, H, world, rank, band_rows=8, layout=split.split("-")[1]), 0, NF)
        out = D.combine(torch.from_numpy(pt.fb))
    else:
        D.render_samples(pt, D.SamplePlan(world, rank), 0, NF)
        out = D.combine(torch.from_numpy(pt.fb), scale=1.0 / NF)
    if rank == 0:
        np.save(out_path, out.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("split", ["tile-diagonal", "tile-bands", "sample"])
def test_two_ranks_match_one(tmp_path, split):
    import torch.multiprocessing as mp
    out_path = str(tmp_path / ("fb_%s.npy" % split))
    mp.spawn(_worker, args=(2, _free_port(), split, out_path), nprocs=2, join=True)
    got = np.load(out_path)
    sc = small_scene("is", "1k")
    ref, _, _ = oracle_render(sc, W, H, 0, NF)
    if split.startswith("tile"):
        assert got[..., :3].tobytes() == ref[..., :3].tobytes(), "tile split must be bit-identical to one renderer"
    else:
        assert rrmse(got, ref) < 1e-5          # same samples, different fp32 summation order
        assert (got[..., 3] == 1.0).all(), "the combined image carries the reference's alpha 1, not the sum of the ranks' alphas"


def test_plans_partition_the_work():
    for world in (1, 2, 3, 8):
        rows = np.zeros(1080, int)
        for r in range(world):
            for y0, y1 in D.TilePlan(1920, 1080, world, r).bands():
                rows[y0:y1] += 1
        assert (rows == 1).all()
        for (w, h) in ((1920, 1080), (854, 480), (33, 9)):          # diagonal 8x4 tiles cover every pixel exactly once
            cover = np.zeros((h, w), int)
            for r in range(world):
                plan = D.TilePlan(w, h, world, r)
                for x0, y0, x1, y1 in plan.tiles():
                    cover[y0:y1, x0:x1] += 1
                assert plan.pixels() == sum((x1 - x0) * (y1 - y0) for x0, y0, x1, y1 in plan.tiles())
            assert (cover == 1).all()
        frames = []
        for r in range(world):
            first, count, stride = D.SamplePlan(world, r).frames(5, 13)
            frames += [first + k * stride for k in range(count)]
        assert sorted(frames) == list(range(5, 18))
    px = [D.TilePlan(1920, 1080, 8, r, layout="bands").pixels() for r in range(8)]
    assert max(px) - min(px) <= 8 * 1920       # interleaving balances the bands
    px = [D.TilePlan(1920, 1080, 8, r).pixels() for r in range(8)]
    assert max(px) - min(px) <= 8 * 4 * 8      # ... and the diagonal tiles to within a few tiles
