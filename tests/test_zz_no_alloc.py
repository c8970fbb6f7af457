"""pt_calculate_hdr_cache_noalloc (no counterpart in the reference): the
calculateHdrCache table from DEVICE buffers the caller owns, with the temporaries in a static device array inside libpt.so
and no allocation in the call.  Checked through the C ABI against the existing, allocating entry point
(pt_calculate_hdr_cache) — the same bar as tests/test_gpu_parity.py uses for that table: identical bits — on two calls in a
row that share the static scratch (the second on a smaller, different map, so stale contents of the first would show).

This file sorts last on purpose; it passes alone in a fresh process and after the others."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from opengl_ray_tracing_b200 import renderer
    return renderer


def test_no_alloc_entry_point_matches_the_allocating_one_twice_in_a_row(R):
    import torch
    from opengl_ray_tracing_b200 import scene as S
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(9)
    maps = [S.synthHdr(512, 256, 1),                                                     # call 1: fills most of the scratch
            (rng.random((40, 96, 3), dtype=np.float32) ** 6 * 40).astype(np.float32),    # call 2: smaller, ragged, spiky
            S.synthHdr(512, 512, 2)]                                                     # call 3: exactly the stated maximum
    # the caller's buffers (the TEST may allocate; the entry point may not)
    d_in = [torch.from_numpy(np.ascontiguousarray(m)).to(dev) for m in maps]
    d_out = [torch.full(m.shape, float("nan"), dtype=torch.float32, device=dev) for m in maps]
    want = [R.calculate_hdr_cache(m) for m in maps]                                      # the existing entry point (host arrays, allocates)
    torch.cuda.synchronize()
    # first call: also resolves the static array's device address (cudaGetSymbolAddress) for this device
    R.calculate_hdr_cache_noalloc(d_in[0].data_ptr(), maps[0].shape[1], maps[0].shape[0], d_out[0].data_ptr())
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    for k in (1, 2):                                                                     # back to back, same scratch, no sync in between
        R.calculate_hdr_cache_noalloc(d_in[k].data_ptr(), maps[k].shape[1], maps[k].shape[0], d_out[k].data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free_before, "the no-alloc entry point changed the device's free memory"
    for k in range(3):
        got = d_out[k].cpu().numpy()
        assert got.shape == want[k].shape and got.tobytes() == np.ascontiguousarray(want[k], np.float32).tobytes(), "map %d" % k
    # and once more on the first map: the scratch now holds another map's intermediates
    d_out[0].fill_(float("nan"))
    R.calculate_hdr_cache_noalloc(d_in[0].data_ptr(), maps[0].shape[1], maps[0].shape[0], d_out[0].data_ptr(),
                                  stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert d_out[0].cpu().numpy().tobytes() == np.ascontiguousarray(want[0], np.float32).tobytes()


def test_no_alloc_entry_point_rejects_what_does_not_fit(R):
    import torch
    L = R.lib()
    buf = torch.zeros(16, dtype=torch.float32, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    assert L.pt_calculate_hdr_cache_noalloc(0, p, 1024, 512, p, None) == -4 and b"exceeds the static scratch" in L.pt_last_error()
    assert L.pt_calculate_hdr_cache_noalloc(0, p, 8192, 1, p, None) == -4
    assert L.pt_calculate_hdr_cache_noalloc(0, None, 4, 4, p, None) == -1 and b"null buffer" in L.pt_last_error()
    assert L.pt_calculate_hdr_cache_noalloc(0, p, 0, 4, p, None) == -1
