"""Per-pixel cost heat map of one render launch (development aid): where do the SM cycles go?

    python tools/cost_map.py [workload] [frames]   ->  gpurun_out/cost_<workload>.npz + a text summary

Uses pt_enable_cost_map / pt_readback_cost_map (clock64 per pixel thread).  Prints the distribution per
32x8 CTA tile and per 8x4 warp tile, and the heaviest tiles with their primary-hit triangle ids.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from opengl_ray_tracing_b200 import renderer as R  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "is-1M-1080p"
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 8
sc, W, H, _ = bench.build_workload(name)
pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
pt.render_spp(0, nf)
pt.synchronize()
pt.enable_cost_map(True)
pt.reset_stats()
pt.render_spp(nf, nf)
cost = pt.readback_cost_map().astype(np.float64)
ids = pt.readback_primary_ids()
st = pt.stats()
print("%s: %d frames, %.2f ms, %.1f Mrays/s, rays/sample %.2f" % (name, nf, st["render_ms"], st["rays"] / st["render_ms"] / 1e3, st["rays"] / st["samples"]))
print("hit fraction %.3f" % float((ids >= 0).mean()))
Hc, Wc = H // 8 * 8, W // 32 * 32
c = cost[:Hc, :Wc]
# a warp's time = max over its 8x4 pixels; a CTA's time = max over its 8 warps
warp = c.reshape(Hc // 4, 4, Wc // 8, 8).max(axis=(1, 3))
cta = c.reshape(Hc // 8, 8, Wc // 32, 32).max(axis=(1, 3))
for nm, a in (("pixel", c), ("warp(max)", warp), ("cta(max)", cta)):
    q = np.percentile(a, [50, 90, 99, 99.9, 100])
    print("%-10s mean %10.0f  p50 %10.0f p90 %10.0f p99 %10.0f p99.9 %10.0f max %10.0f cycles" % (nm, a.mean(), *q))
print("sum of warp-max cycles / (148 SMs x 4 schedulers) = %.2f ms at 1.965 GHz if each scheduler ran one warp at a time" % (warp.sum() / 592 / 1.965e6))
print("kernel would need >= max CTA = %.2f ms" % (cta.max() / 1.965e6))
order = np.argsort(cta.ravel())[::-1][:12]
for o in order:
    ty, tx = divmod(int(o), cta.shape[1])
    blk = c[ty * 8:ty * 8 + 8, tx * 32:tx * 32 + 32]
    iy, ix = np.unravel_index(int(np.argmax(blk)), blk.shape)
    print("  CTA tile (%3d,%3d) rows %4d.. cols %4d..: max %.0f cycles (%.2f ms) at pixel (%d,%d) tri %d; tile mean %.0f"
          % (tx, ty, ty * 8, tx * 32, cta[ty, tx], cta[ty, tx] / 1.965e6, tx * 32 + ix, ty * 8 + iy, ids[ty * 8 + iy, tx * 32 + ix], blk.mean()))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "cost_%s.npz" % name), cost=cost.astype(np.uint32), ids=ids)
pt.close()
