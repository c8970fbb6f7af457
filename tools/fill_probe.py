"""How fast can the bench flush L2?  256 MiB written with fill_ on one-byte / eight-byte / float views and with cudaMemsetAsync (development aid)."""
import torch, ctypes, time
n = 256 << 20
b = torch.empty(n, dtype=torch.uint8, device="cuda")
b64 = b.view(torch.int64); b32 = b.view(torch.float32)
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemsetAsync.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p]
s = torch.cuda.current_stream()
def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for name, fn in (("uint8 fill_", lambda: b.fill_(3)), ("int64 view fill_", lambda: b64.fill_(3)), ("float32 view fill_", lambda: b32.fill_(1.0)),
                 ("cudaMemsetAsync", lambda: rt.cudaMemsetAsync(b.data_ptr(), 3, n, ctypes.c_void_p(s.cuda_stream))), ("zero_", lambda: b.zero_())):
    ms = t(fn); print("%-20s %.4f ms  %.2f TB/s" % (name, ms, n / ms / 1e9))
