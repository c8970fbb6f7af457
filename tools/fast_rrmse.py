"""rRMSE of PT_MODE_FAST against the oracle on bench.py's centre crop of the metric scene, for several builds of libpt
(tools/build_variant.py): which floating-point freedoms of the production build cost how much parity.

    python tools/fast_rrmse.py "default nofmad precdiv" [frames]
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
from oracle import pyoracle as O

nf = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sc, W, H, _ = bench.build_workload("is-1M-1080p")
reg = bench.crop_region(W, H)
x0, y0, x1, y1 = reg
ref, ids_ref, _ = O.render(sc.variant, sc.max_bounce, sc.tris, sc.nodes, sc.hdr, sc.hdrCache, W, H, sc.eye, sc.cam, 0, nf, math=1, region=reg)
ref = ref[y0:y1, x0:x1].astype(np.float64)
here = os.path.join(ROOT, "opengl_ray_tracing_b200")
for v in sys.argv[1].split():
    R._lib, R._LIB_PATH = None, os.path.join(here, "libpt.so" if v == "default" else "libpt_%s.so" % v)
    for mode, name in ((R.MODE_FAST, "FAST"), (R.MODE_PARITY_FAST, "PARITY_FAST")):
        pt = R.PathTracer(W, H, sc.variant, mode=mode)
        pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam); pt.set_region(*reg)
        pt.render_spp(0, nf)
        img = pt.readback()[y0:y1, x0:x1].astype(np.float64); ids = pt.readback_primary_ids()[y0:y1, x0:x1]
        pt.close()
        rr = float(np.sqrt(((img[..., :3] - ref[..., :3]) ** 2).sum() / (ref[..., :3] ** 2).sum()))
        print("%-10s %-11s %d frames: rRMSE %.3e  ids equal %s  differing px %d  mean %.6f vs %.6f" % (
            v, name, nf, rr, bool((ids == ids_ref[y0:y1, x0:x1]).all()), int((img != ref).any(axis=2).sum()), img[..., :3].mean(), ref[..., :3].mean()), flush=True)
