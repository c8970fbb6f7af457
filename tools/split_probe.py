"""One-GPU probe of the multi-GPU splits: what does rank 0 of an N-rank job cost, kernel only? (development aid)

    python tools/split_probe.py [workload] [spp]
For world in 1, 2, 4, 8: the tile split (band interleave) and the sample split (frame stride) of `spp` frames, with the
L2 flushed before each launch like bench.py does.  Efficiency = t(world 1) / (world * t(world))."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
name = sys.argv[1] if len(sys.argv) > 1 else "is-1M-1080p"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 256
sc, W, H, _ = bench.build_workload(name)
pt = R.PathTracer(W, H, sc.variant)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.ExternalStream(pt.stream())


def timed(fn, reps=5):
    best = []
    for _ in range(reps):
        with torch.cuda.stream(stream):
            flush.fill_(1)
        pt.synchronize(); pt.reset_stats()
        fn()
        pt.synchronize()
        best.append(pt.stats()["render_ms"])
    return sorted(best)[len(best) // 2]


pt.render_spp(0, 16); pt.synchronize()
base = None
for world in (1, 2, 4, 8):
    for rank in (range(world) if world == 8 else sorted({0, world - 1})):
        pt.set_band_interleave(world, rank)
        t_band = timed(lambda: pt.render_spp(0, spp))
        pt.set_tile_interleave(world, rank)
        t_tile = timed(lambda: pt.render_spp(0, spp))
        pt.set_band_interleave(1, 0)
        t_samp = timed(lambda: pt.render_frames(rank, spp // world, world, R.ACC_SUM))
        if base is None:
            base = t_tile
        print("world %d rank %d: bands %.3f ms (eff %.1f %%)   diagonal tiles %.3f ms (eff %.1f %%)   sample split %.3f ms (eff %.1f %%)"
              % (world, rank, t_band, 100 * base / (world * t_band), t_tile, 100 * base / (world * t_tile), t_samp, 100 * base / (world * t_samp)), flush=True)
pt.close()
