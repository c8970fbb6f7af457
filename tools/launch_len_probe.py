"""What does a launch cost beyond its frames?  (development aid, one GPU)

    python tools/launch_len_probe.py [workload]
Kernel time of a sample-split launch (frame stride 8, PT_ACC_SUM, whole image) of n = 4 .. 256 frames, with the L2 flushed
before the launch (as bench.py does) and without; a least-squares line t = a + b n gives the fixed cost `a` of a launch."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from opengl_ray_tracing_b200 import renderer as R
name = sys.argv[1] if len(sys.argv) > 1 else "is-1M-1080p"
sc, W, H, _ = bench.build_workload(name)
pt = R.PathTracer(W, H, sc.variant)
pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
pt.set_camera(sc.eye, sc.cam)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.ExternalStream(pt.stream())


def timed(n, do_flush, reps=5):
    ts = []
    for r in range(reps):
        if do_flush:
            with torch.cuda.stream(stream):
                flush.fill_(r)
        pt.synchronize(); pt.reset_stats()
        pt.render_frames(3, n, 8, R.ACC_SUM)
        pt.synchronize()
        ts.append(pt.stats()["render_ms"])
    return sorted(ts)[len(ts) // 2]


pt.render_spp(0, 16); pt.synchronize()
ns = [1, 2, 4, 8, 16, 32, 64, 128, 256]
print('PT_TAIL_FRAMES =', os.environ.get('PT_TAIL_FRAMES', '(default)'))
for do_flush in (True,):
    t = [timed(n, do_flush) for n in ns]
    b, a = np.polyfit(ns[4:], t[4:], 1)
    print("flush" if do_flush else "no flush", " ".join("%d:%.3f" % (n, x) for n, x in zip(ns, t)), "| fit t = %.3f + %.4f n ms" % (a, b), flush=True)
    print("   per-frame ms:", " ".join("%d:%.4f" % (n, x / n) for n, x in zip(ns, t)), flush=True)
pt.close()
