"""Renders a few frames of a bench workload with a chosen pipeline (for ncu launch lists)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
name = sys.argv[1]; pipe = int(sys.argv[2]); nf = int(sys.argv[3]); mode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
sc, W, H, _ = bench.build_workload(name)
pt = R.PathTracer(W, H, sc.variant, mode=mode, pipeline=pipe)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
pt.render_spp(0, nf); pt.synchronize(); pt.reset_stats()
pt.render_spp(nf, nf)
st = pt.stats()
print("%s pipeline %d mode %d: %d frames %.3f ms -> %.1f Mrays/s, rays/sample %.3f, launches %d" % (name, pipe, mode, nf, st["render_ms"], st["rays"] / st["render_ms"] / 1e3, st["rays"] / max(1, st["samples"]), st["launches"]))
if mode == 4:
    print("per ray: inner %.2f tris %.2f cand %.3f; tie re-traces %d of %d rays" % (st["inner_visits"] / st["rays"], st["tri_tests"] / st["rays"], st["leaf_candidates"] / st["rays"], st["tie_retraces"], st["rays"]))
if mode == 4:
    h = pt.ray_histogram()
    print("steps/ray histogram:", {("<%d" % (1 << k)): v for k, v in enumerate(h) if v})
pt.close()
