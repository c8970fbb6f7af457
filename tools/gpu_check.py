"""Quick GPU sanity run (development aid): parity of the three modes against the oracle + first timings."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_ray_tracing_b200 import scenes, renderer as R
from oracle import pyoracle as O

def rrmse(a, b):
    return float(np.sqrt(((a[..., :3].astype(np.float64) - b[..., :3]) ** 2).sum() / max(1e-30, (b[..., :3].astype(np.float64) ** 2).sum())))

print(R.build_info())
W, H, NF = 256, 192, 3
for kind in ("lambert", "disney", "is"):
    sc = scenes.build(kind, "10k", hdr_res=(512, 256), camera=(10.0, 25.0, 4.0))
    t0 = time.time()
    ref, ids_ref, _ = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, NF, math=1)
    t_or = time.time() - t0
    for mode in (R.MODE_PARITY_REF, R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST, R.MODE_FAST):
        pt = R.PathTracer(W, H, sc.variant, mode=mode)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
        pt.set_camera(sc.eye, sc.cam)
        pt.render_spp(0, NF)
        img = pt.readback(); ids = pt.readback_primary_ids(); st = pt.stats()
        nd = int((img != ref).any(axis=2).sum())
        print("%-8s mode %d: diff px %6d / %d  rRMSE %.3e  ids equal %s  mean %.6f vs %.6f  rays %d  %.2f ms (oracle %.2fs) depth %d inner %d mats %d"
              % (kind, mode, nd, W * H, rrmse(img, ref), bool((ids == ids_ref).all()), img[..., :3].mean(), ref[..., :3].mean(),
                 st["rays"], st["render_ms"], t_or, st["max_stack_depth"], st["n_inner_nodes"], st["n_materials"]), flush=True)
        pt.close()

# the metric scene (1M triangles; the reference builder's tree is 6781 levels deep here)
sc = scenes.build("is", "1M", hdr_res=(2048, 1024))
W2, H2 = 640, 360
ref, ids_ref, _ = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W2, H2, sc.eye, sc.cam, 0, 2, math=1)
for mode in (R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST, R.MODE_FAST):
    pt = R.PathTracer(W2, H2, sc.variant, mode=mode)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(0, 2)
    img = pt.readback(); ids = pt.readback_primary_ids(); st = pt.stats()
    print("is-1M    mode %d: diff px %6d / %d  rRMSE %.3e  ids equal %s  depth %d" % (mode, int((img != ref).any(axis=2).sum()), W2 * H2, rrmse(img, ref), bool((ids == ids_ref).all()), st["max_stack_depth"]), "fast depth", st["fast_tree_depth"], "build ms %.0f" % st["upload_build_ms"], flush=True)
    pt.close()

# first timings at 1080p
for kind, size in (("lambert", "100k"), ("is", "100k"), ("is", "1M"), ("disney", "1M")):
    sc = scenes.build(kind, size, hdr_res=(1024, 512))
    for mode in ((R.MODE_PARITY_REF, R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST, R.MODE_FAST) if size != "1M" else (R.MODE_PARITY_WIDE, R.MODE_PARITY_FAST, R.MODE_FAST)):
        pt = R.PathTracer(1920, 1080, sc.variant, mode=mode)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
        pt.set_camera(sc.eye, sc.cam)
        pt.render_spp(0, 2); pt.synchronize(); pt.reset_stats()
        nf = 4 if mode != R.MODE_PARITY_REF else 2
        pt.render_spp(2, nf); st = pt.stats()
        print("%s-%s 1080p mode %d: %d frames %.2f ms  -> %.1f Mrays/s  %.2f spp/s(frames/s)  rays/sample %.2f"
              % (kind, size, mode, nf, st["render_ms"], st["rays"] / st["render_ms"] / 1e3, nf / st["render_ms"] * 1e3, st["rays"] / st["samples"]), flush=True)
        pt.close()
