"""A/B of libpt builds in ONE process: every variant library (tools/build_variant.py -> libpt_<name>.so) renders the same
workloads, scenes loaded once.

    python tools/ab.py "default smem16 tos" "is-1M-1080p:128 disney-1M-1080p:64" [--pipeline 2] [--reps 2]

Prints one line per (workload, variant): Grays/s of each repetition (device time of the launch, CUDA events inside libpt).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R

ap = argparse.ArgumentParser()
ap.add_argument("variants")
ap.add_argument("workloads")
ap.add_argument("--pipeline", type=int, default=2)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--mode", type=int, default=0)
ap.add_argument("--env", type=int, default=0, help="1: PT_ENV_CDF_SEARCH (variant 2 only)")
args = ap.parse_args()

here = os.path.join(ROOT, "opengl_ray_tracing_b200")
for wl in args.workloads.split():
    name, nf = wl.split(":")
    nf = int(nf)
    sc, W, H, _ = bench.build_workload(name)
    for v in args.variants.split():
        path = os.path.join(here, "libpt.so" if v == "default" else "libpt_%s.so" % v)
        if not os.path.exists(path):
            print("[%s] %s: missing %s" % (name, v, path), flush=True)
            continue
        R._lib, R._LIB_PATH = None, path
        try:
            pt = R.PathTracer(W, H, sc.variant, mode=args.mode, pipeline=args.pipeline, env_sampler=args.env)
            pt.upload_scene(sc.tris, sc.nodes, sc.hdr, None if args.env else sc.hdrCache)
            pt.set_camera(sc.eye, sc.cam)
            pt.render_spp(0, nf); pt.synchronize()
            out = []
            for r in range(args.reps):
                pt.reset_stats()
                pt.render_spp(nf * (r + 1), nf)
                st = pt.stats()
                out.append(st["rays"] / st["render_ms"] / 1e6)
            print("[%s] pipeline %d %-12s %s Grays/s  (%.3f ms, %d launches, %.3f rays/sample)" % (
                name, args.pipeline, v, " ".join("%.2f" % g for g in out), st["render_ms"], st["launches"], st["rays"] / max(1, st["samples"])), flush=True)
            pt.close()
        except Exception as e:                        # a broken variant must not stop the sweep
            print("[%s] pipeline %d %-12s FAILED: %s" % (name, args.pipeline, v, e), flush=True)
