"""pt_comm_* / pt_combine with a world of ONE (runs on the single-GPU box): the NCCL bootstrap (dlopen, unique id,
ncclCommInitRank), the fused consume kernel (scale, alpha, pass3 tone map) and the asynchronous host copy, against the plain
single-context path.  The N > 1 form of the same code is tests/test_multi_gpu.py (needs two GPUs)."""
import numpy as np
import pytest

from helpers import rrmse, small_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from opengl_ray_tracing_b200 import renderer
    return renderer


def test_combine_with_a_world_of_one(R):
    import torch
    sc = small_scene("is", "10k", hdr_res=(256, 128))
    W, H, NF = 200, 120, 8
    ref_pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    ref_pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); ref_pt.set_camera(sc.eye, sc.cam)
    ref_pt.render_spp(0, NF)
    ref, ref_disp = ref_pt.readback(), ref_pt.readback_display()
    ref_pt.close()

    pt = R.PathTracer(W, H, sc.variant, mode=R.MODE_FAST)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache); pt.set_camera(sc.eye, sc.cam)
    uid = R.comm_unique_id()
    assert len(uid) == R.COMM_ID_BYTES and any(uid)
    pt.comm_init(1, 0, uid)
    assert pt.comm_info()["world"] == 1 and pt.comm_info()["rank"] == 0
    with pytest.raises(R.PtError, match="already has a communicator"):
        pt.comm_init(1, 0, uid)
    # tile-split form: the running mean as it is, alpha 1, display bytes = pass3 of it
    pt.render_spp(0, NF)
    img, disp = pt.combine_readback(R.COMBINE_DISJOINT, NF, display=True)
    assert img.tobytes() == ref.tobytes() and disp.tobytes() == ref_disp.tobytes()
    # sample-split form: plain sums divided by the frame count
    pt.clear()
    pt.render_frames(0, NF, 1, R.ACC_SUM)
    img, disp = pt.combine_readback(R.COMBINE_SUM, NF, display=True)
    assert rrmse(img, ref) <= 1e-5 and (img[..., 3] == 1.0).all()
    assert (np.abs(disp.astype(int) - ref_disp.astype(int)) <= 1).all()
    # asynchronous host copy: lands at synchronize(), rendering goes on meanwhile and must not disturb it
    hf = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    pt.clear()
    pt.render_spp(0, NF)
    pt.combine(R.COMBINE_DISJOINT | R.COMBINE_ASYNC, NF, hf.data_ptr(), None)
    pt.render_spp(NF, 16)
    pt.synchronize()
    assert hf.numpy().tobytes() == ref.tobytes()
    # PT_COMBINE_PIPELINED needs peers: with a world of one the flag means PT_COMBINE_ASYNC, and the fence has nothing to wait for
    hf.zero_()
    pt.clear()
    pt.render_spp(0, NF)
    pt.combine(R.COMBINE_DISJOINT | R.COMBINE_PIPELINED, NF, hf.data_ptr(), None)
    pt.combine_fence()
    pt.render_spp(NF, 16)
    pt.synchronize()
    assert hf.numpy().tobytes() == ref.tobytes()
    # argument checks
    with pytest.raises(R.PtError, match="total_frames"):
        pt.combine(R.COMBINE_SUM, 0)
    with pytest.raises(R.PtError, match="flags"):
        pt.combine(2, 1)
    with pytest.raises(R.PtError, match="flags"):
        pt.combine(32, 1)
    pt.close()
    q = R.PathTracer(8, 8, 0)
    with pytest.raises(R.PtError, match="pt_comm_init first"):
        q.combine(R.COMBINE_SUM, 1)
    q.close()
