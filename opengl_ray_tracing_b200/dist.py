"""Multi-GPU sharding of the path tracer: one process per GPU, scene replicated, pixels or samples split.

Every pixel and every frame index of the reference is independent (per-pixel seed from (px, py, frame),
pass1.fsh:73-76; Sobol index = frame + 1, pass1.fsh:795), so the path shards with no data-path collective;
the only exchange is combining the per-GPU fp32 framebuffers when an image is wanted:

  tile split    rank r renders the 8x4 pixel tiles (tx, s) with (tx + s) % world == r (or whole 8-row bands with
                b % world == r) of ALL frames (running mean, bit-identical to a single-GPU render); combine = sum of disjoint regions
                (x + 0 is exact) with one NCCL reduce, or nothing at all if only rank-local tiles are needed.
  sample split  rank r renders frames first + r, first + r + world, … of ALL pixels as a plain sum; combine =
                one NCCL reduce(sum) + a scale by 1/total_frames (fp32 sum order differs from the sequential
                running mean: covered by the 1e-3 rRMSE criterion, not bit-exact).

torch.distributed is plumbing only (rendezvous + the NCCL/gloo collective); the CPU tests drive the same
planning and combine code over gloo with a stand-in renderer.
"""
from __future__ import annotations

from dataclasses import dataclass

BAND_ROWS = 8      # one CTA tile row (PT_TILE_H): bands stay aligned to the kernel's 32x8 pixel tiles
TILE_COLS, TILE_ROWS = 8, 4     # one warp's pixel tile: the unit of the diagonal split


@dataclass(frozen=True)
class TilePlan:
    width: int
    height: int
    world: int
    rank: int
    band_rows: int = BAND_ROWS
    layout: str = "diagonal"       # "diagonal": 8x4 tiles, (tx + strip) % world == rank; "bands": whole bands, b % world == rank

    def tiles(self):
        """[(x0, y0, x1, y1)] owned by this rank under the diagonal layout (pt_set_tile_interleave): 8x4 pixel tiles
        (tx, strip) with (tx + strip) % world == rank — an even sample of the whole image for every rank."""
        out = []
        ns = (self.height + TILE_ROWS - 1) // TILE_ROWS
        ntx = (self.width + TILE_COLS - 1) // TILE_COLS
        for s in range(ns):
            for tx in range((self.rank - s) % self.world, ntx, self.world):
                out.append((tx * TILE_COLS, s * TILE_ROWS, min(self.width, (tx + 1) * TILE_COLS), min(self.height, (s + 1) * TILE_ROWS)))
        return out

    def rects(self):
        """The owned pixel rectangles of the chosen layout."""
        if self.layout == "bands":
            return [(0, y0, self.width, y1) for y0, y1 in self.bands()]
        return self.tiles()

    def bands(self):
        """[(y0, y1)] owned by this rank — interleaved so sky rows and geometry rows spread over all ranks."""
        out = []
        nb = (self.height + self.band_rows - 1) // self.band_rows
        for b in range(self.rank, nb, self.world):
            out.append((b * self.band_rows, min(self.height, (b + 1) * self.band_rows)))
        return out

    def pixels(self):
        return sum((x1 - x0) * (y1 - y0) for x0, y0, x1, y1 in self.rects())


@dataclass(frozen=True)
class SamplePlan:
    world: int
    rank: int

    def frames(self, first_frame: int, n_frames: int):
        """(first, count, stride) of the frames in [first_frame, first_frame + n_frames) owned by this rank."""
        count = (n_frames - self.rank + self.world - 1) // self.world if n_frames > self.rank else 0
        return first_frame + self.rank, count, self.world


def render_tiles(pt, plan: TilePlan, first_frame: int, n_frames: int):
    """Tile-split render of frames [first_frame, first_frame+n_frames) on this rank's tiles (running mean)."""
    if hasattr(pt, "set_tile_interleave"):          # the CUDA renderer does the interleave inside one launch
        pt.set_region(0, 0, plan.width, plan.height)
        if plan.layout == "bands":
            pt.set_band_interleave(plan.world, plan.rank)
        else:
            pt.set_tile_interleave(plan.world, plan.rank)
        pt.render_spp(first_frame, n_frames)
        return
    for x0, y0, x1, y1 in plan.rects():             # stand-in renderers (CPU tests): one call per rectangle
        pt.set_region(x0, y0, x1, y1)
        pt.render_spp(first_frame, n_frames)


def render_samples(pt, plan: SamplePlan, first_frame: int, n_frames: int):
    """Sample-split render: this rank's frames as a plain sum over the whole image."""
    from .renderer import ACC_SUM
    first, count, stride = plan.frames(first_frame, n_frames)
    if count:
        pt.render_frames(first, count, stride, ACC_SUM)


def combine(fb, out=None, dst: int = 0, scale: float | None = None, group=None):
    """Sums the per-rank framebuffers onto rank `dst` (torch tensor: CUDA -> NCCL, CPU -> gloo) and scales rgb by
    `scale` there (sample split: 1/total_frames).  The sum goes into `out` (a copy of fb), NOT into fb: each rank's
    framebuffer keeps only what that rank rendered, so rendering can continue and be combined again later.
    Returns `out` (meaningful on rank `dst`)."""
    import torch.distributed as dist
    if out is None:
        out = fb.clone()
    else:
        out.copy_(fb)
    multi = dist.is_initialized() and dist.get_world_size(group) > 1
    if multi:
        dist.reduce(out, dst=dst, op=dist.ReduceOp.SUM, group=group)
    if scale is not None and (not multi or dist.get_rank(group) == dst):
        import torch
        world = dist.get_world_size(group) if multi else 1
        # rgb * scale; alpha back to the reference's 1 (every rank contributed its own 1): one contiguous pass
        out *= torch.tensor([scale, scale, scale, 1.0 / world], dtype=out.dtype, device=out.device)
    return out


class DeviceFramebuffer:
    """Zero-copy torch view of a PathTracer's device framebuffer (W*H float4) via __cuda_array_interface__."""

    def __init__(self, pt):
        ptr, nbytes = pt.framebuffer_device_ptr()
        self.__cuda_array_interface__ = {"shape": (pt.height, pt.width, 4), "typestr": "<f4", "data": (ptr, False),
                                         "version": 3, "strides": None}
        self._keepalive = pt

    def tensor(self, device):
        import torch
        return torch.as_tensor(self, device=device)
