"""What does a sample cost where? Renders sub-rectangles of a bench workload (sky / floor / mesh on is-1M) and prints the
kernel time per sample and per ray (development aid).   python tools/region_probe.py [workload] [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
name = sys.argv[1] if len(sys.argv) > 1 else "is-1M-1080p"
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 64
sc, W, H, _ = bench.build_workload(name)
pt = R.PathTracer(W, H, sc.variant)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
regions = {"whole": (0, 0, W, H), "top quarter": (0, 3 * H // 4, W, H), "bottom quarter": (0, 0, W, H // 4),
           "centre 1/5 x 1/4": (2 * W // 5, 3 * H // 8, 3 * W // 5, 5 * H // 8)}
for label, (x0, y0, x1, y1) in regions.items():
    pt.set_region(x0, y0, x1, y1)
    pt.render_spp(0, nf); pt.synchronize(); pt.reset_stats()
    pt.render_spp(nf, nf)
    st = pt.stats()
    print("%-18s %4dx%-4d: %8.3f ms  %.3f ns/sample  rays/sample %.3f  %.1f Mrays/s  %.2f Gsamples/s"
          % (label, x1 - x0, y1 - y0, st["render_ms"], st["render_ms"] * 1e6 / st["samples"], st["rays"] / st["samples"],
             st["rays"] / st["render_ms"] / 1e3, st["samples"] / st["render_ms"] / 1e6), flush=True)
pt.close()
