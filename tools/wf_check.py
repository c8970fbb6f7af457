"""GPU development check: wavefront pipeline vs megakernel (bit-identical under the parity arithmetic) + timings."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_ray_tracing_b200 import scenes, renderer as R  # noqa: E402


def rrmse(a, b):
    return float(np.sqrt(((a[..., :3].astype(np.float64) - b[..., :3]) ** 2).sum() / max(1e-30, (b[..., :3].astype(np.float64) ** 2).sum())))


def render(sc, W, H, mode, pipeline, first, nf):
    pt = R.PathTracer(W, H, sc.variant, mode=mode, pipeline=pipeline)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
    pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(first, nf)
    img, ids, st = pt.readback(), pt.readback_primary_ids(), pt.stats()
    pt.close()
    return img, ids, st


print(R.build_info())
W, H, NF = 320, 200, 3
for kind in ("lambert", "disney", "is"):
    sc = scenes.build(kind, "10k", hdr_res=(512, 256), camera=(10.0, 25.0, 2.5))
    for mode in (R.MODE_PARITY_FAST, R.MODE_PARITY_WIDE, R.MODE_FAST):
        a, ia, sa = render(sc, W, H, mode, R.MEGAKERNEL, 0, NF)
        for pipe in (R.WAVEFRONT, R.PERSISTENT):
            b, ib, sb = render(sc, W, H, mode, pipe, 0, NF)
            print("%-8s mode %d pipeline %d: identical %s  ids equal %s  diff px %d  rRMSE %.2e  rays %d vs %d  launches %d vs %d"
                  % (kind, mode, pipe, a.tobytes() == b.tobytes(), bool((ia == ib).all()), int((a != b).any(axis=2).sum()), rrmse(b, a),
                     sa["rays"], sb["rays"], sa["launches"], sb["launches"]), flush=True)

if len(sys.argv) > 1 and sys.argv[1] == "quick":
    sys.exit(0)
for kind, size, hdr in (("is", "1M", (2048, 1024)), ("lambert", "100k", (1024, 512)), ("disney", "1M", (1024, 512))):
    sc = scenes.build(kind, size, hdr_res=hdr)
    for pipe in (R.MEGAKERNEL, R.WAVEFRONT, R.PERSISTENT):
        pt = R.PathTracer(1920, 1080, sc.variant, mode=R.MODE_FAST, pipeline=pipe)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
        pt.set_camera(sc.eye, sc.cam)
        pt.render_spp(0, 8); pt.synchronize(); pt.reset_stats()
        t0 = time.perf_counter()
        pt.render_spp(8, 16); st = pt.stats()
        wall = (time.perf_counter() - t0) * 1e3
        print("%s-%s 1080p pipeline %d: 16 frames %.2f ms (wall %.2f)  -> %.1f Mrays/s  %.1f spp/s  rays/sample %.2f  launches %d"
              % (kind, size, pipe, st["render_ms"], wall, st["rays"] / st["render_ms"] / 1e3, 16 / st["render_ms"] * 1e3,
                 st["rays"] / st["samples"], st["launches"]), flush=True)
        pt.close()
