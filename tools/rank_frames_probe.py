"""Is the slowest rank of the 8-GPU sample split slow because of its FRAMES? (development aid, one GPU)

    python tools/rank_frames_probe.py
Kernel time of what rank r of 8 renders in a 256-spp step — frames r, r + 8, ... (round-robin) and frames 32 r .. 32 r + 31
(contiguous blocks) — all on ONE GPU, L2 flushed before each launch, median of 7."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
sc, W, H, _ = bench.build_workload("is-1M-1080p")
pt = R.PathTracer(W, H, sc.variant)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.ExternalStream(pt.stream())


def timed(first, n, stride, reps=7):
    ts = []
    for r in range(reps):
        with torch.cuda.stream(stream):
            flush.fill_(r)
        pt.synchronize(); pt.reset_stats()
        pt.render_frames(first, n, stride, R.ACC_SUM)
        pt.synchronize()
        st = pt.stats()
        ts.append((st["render_ms"], st["rays"]))
    ts.sort()
    return ts[len(ts) // 2]


pt.render_spp(0, 16); pt.synchronize()
for base in (0, 256, 1280):
    rr = [timed(base + r, 32, 8) for r in range(8)]
    cc = [timed(base + 32 * r, 32, 1) for r in range(8)]
    print("step at frame %d" % base)
    print("  round-robin ms  :", " ".join("%.3f" % t for t, _ in rr), "| Mrays:", " ".join("%.1f" % (n / 1e6) for _, n in rr), flush=True)
    print("  contiguous  ms  :", " ".join("%.3f" % t for t, _ in cc), "| Mrays:", " ".join("%.1f" % (n / 1e6) for _, n in cc), flush=True)
pt.close()
