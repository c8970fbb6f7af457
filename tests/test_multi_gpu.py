"""pt_comm_* / pt_combine on real GPUs (needs >= 2 devices; one context and one host thread per GPU in ONE process, as
ptdrive --gpus N does): the combined image of a tile split is the single-GPU image bit for bit, that of a sample split is
within 1e-5 rRMSE of it, and the display bytes are pass3 of the combined image."""
import threading

import numpy as np
import pytest

from helpers import rrmse, small_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from opengl_ray_tracing_b200 import renderer
    return renderer


def run_ranks(R, sc, W, H, NF, world, split, overlap=False):
    uid = R.comm_unique_id()
    out, errs = {}, []

    def worker(rank):
        try:
            pt = R.PathTracer(W, H, sc.variant, device=rank, mode=R.MODE_FAST)
            pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
            pt.set_camera(sc.eye, sc.cam)
            pt.comm_init(world, rank, uid)
            info = pt.comm_info()
            if split == "tile":
                pt.set_tile_interleave(world, rank)
                pt.render_spp(0, NF)
                if overlap and rank == 0:          # PT_COMBINE_ASYNC: the copy lands at synchronize(), rendering goes on meanwhile
                    import torch
                    hf = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
                    hd = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
                    pt.combine(R.COMBINE_DISJOINT | R.COMBINE_DISPLAY | R.COMBINE_ASYNC, NF, hf.data_ptr(), hd.data_ptr())
                    pt.render_spp(NF, 4)
                    pt.synchronize()
                    res = (hf.numpy().copy(), hd.numpy().copy())
                elif overlap:
                    pt.combine(R.COMBINE_DISJOINT | R.COMBINE_DISPLAY | R.COMBINE_ASYNC, NF)
                    pt.render_spp(NF, 4)
                    res = None
                else:
                    res = pt.combine_readback(R.COMBINE_DISJOINT, NF, display=True)
            else:
                n = (NF - rank + world - 1) // world
                pt.render_frames(rank, n, world, R.ACC_SUM)
                res = pt.combine_readback(R.COMBINE_SUM, NF, display=True)
            pt.synchronize()
            if rank == 0:
                out["img"], out["disp"], out["info"] = res[0], res[1], info
            pt.close()
        except Exception as e:                       # noqa: BLE001 — reported by the main thread
            errs.append((rank, e))

    th = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    return out


def test_combine_tile_and_sample_split(R):
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H, NF = 200, 120, 8
    solo = R.PathTracer(W, H, sc.variant, device=0, mode=R.MODE_FAST)
    solo.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); solo.set_camera(sc.eye, sc.cam)
    solo.render_spp(0, NF)
    ref, ref_disp = solo.readback(), solo.readback_display()
    solo.close()
    t = run_ranks(R, sc, W, H, NF, 2, "tile")
    assert t["info"]["world"] == 2
    assert t["img"].tobytes() == ref.tobytes(), "tile split + pt_combine must be bit-identical to one GPU"
    assert t["disp"].tobytes() == ref_disp.tobytes()
    t = run_ranks(R, sc, W, H, NF, 2, "tile", overlap=True)
    assert t["img"].tobytes() == ref.tobytes() and t["disp"].tobytes() == ref_disp.tobytes(), "PT_COMBINE_ASYNC"
    s = run_ranks(R, sc, W, H, NF, 2, "sample")
    assert rrmse(s["img"], ref) <= 1e-5 and (s["img"][..., 3] == 1.0).all()
    assert (np.abs(s["disp"].astype(int) - ref_disp.astype(int)) <= 1).all()


def run_steps(R, sc, W, H, world, steps, spp, how, split="sample"):
    """`steps` steps of `spp` frames, each followed by a combine into its own pinned host image on rank 0 (sample split:
    per-rank sums of all frames so far; tile split: the running mean of the rank's tiles).  how = extra pt_combine flags
    (one value, or one per step)."""
    import torch
    from opengl_ray_tracing_b200 import dist as D
    uid = R.comm_unique_id()
    out, errs = {}, []
    meet = threading.Barrier(world)
    hows = list(how) if isinstance(how, (list, tuple)) else [how] * steps

    def worker(rank):
        try:
            pt = R.PathTracer(W, H, sc.variant, device=rank, mode=R.MODE_FAST)
            pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
            pt.set_camera(sc.eye, sc.cam)
            pt.comm_init(world, rank, uid)
            assert pt.comm_info()["peer_memory"], "two GPUs of one box without peer memory"
            hosts = [torch.empty((H, W, 4), dtype=torch.float32).pin_memory() for _ in range(steps)] if rank == 0 else None
            disps = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(steps)] if rank == 0 else None
            plan, tiles = D.SamplePlan(world, rank), D.TilePlan(W, H, world, rank)
            for s in range(steps):
                if split == "sample":
                    D.render_samples(pt, plan, s * spp, spp)
                    kind = R.COMBINE_SUM
                else:
                    D.render_tiles(pt, tiles, s * spp, spp)
                    kind = R.COMBINE_DISJOINT
                pt.combine(kind | R.COMBINE_DISPLAY | hows[s], (s + 1) * spp, hosts[s].data_ptr() if rank == 0 else None,
                           disps[s].data_ptr() if rank == 0 else None)
            if hows[-1] & R.COMBINE_PIPELINED:
                pt.combine_fence()
            pt.synchronize()
            if rank == 0:
                out["imgs"] = [h.numpy().copy() for h in hosts]
                out["disps"] = [d.numpy().copy() for d in disps]
            meet.wait(timeout=120)                  # nobody unmaps a buffer a peer's kernel may still read
            pt.close()
        except Exception as e:                       # noqa: BLE001 — reported by the main thread
            errs.append((rank, e))
            meet.abort()

    th = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    return out


@pytest.mark.parametrize("split", ["sample", "tile"])
def test_pipelined_combine_equals_the_synchronous_one(R, split):
    """PT_COMBINE_PIPELINED (snapshot + in-kernel flag waits on a second stream, under the next step's render) must hand
    rank 0 the very images the synchronous combine does, step after step — seven steps, so both slots are reused thrice."""
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H, steps, spp = 200, 120, 7, 4
    a = run_steps(R, sc, W, H, 2, steps, spp, 0, split)
    b = run_steps(R, sc, W, H, 2, steps, spp, R.COMBINE_PIPELINED, split)
    for s in range(steps):
        assert a["imgs"][s].tobytes() == b["imgs"][s].tobytes(), "pipelined combine differs from the synchronous one at step %d" % s
        assert a["disps"][s].tobytes() == b["disps"][s].tobytes()
    assert not (a["imgs"][0] == a["imgs"][steps - 1]).all()      # the steps do differ (more frames)
    # a synchronous combine right after pipelined ones (the order bench.py's parity step uses)
    c = run_steps(R, sc, W, H, 2, 4, spp, [R.COMBINE_PIPELINED, R.COMBINE_PIPELINED, 0, R.COMBINE_PIPELINED], split)
    for s in range(4):
        assert c["imgs"][s].tobytes() == a["imgs"][s].tobytes(), "mixed pipelined / synchronous combines, step %d" % s
