"""How fast can the six boundary slabs of a host-resident 256^3 f64 field cross PCIe?  (design input for the e2e path)
Compares copy-engine 2-D copies of the strided x-slab (8-byte rows at 2 KiB pitch) with contiguous copies of the same
number of bytes, H2D and D2H, from/to pinned host memory."""
import ctypes
import time

import torch

rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                 ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
n = 256
host = torch.zeros(n * n * n, dtype=torch.float64).pin_memory()
dev = torch.zeros(n * n * n, dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream().cuda_stream
H2D, D2H = 1, 2


def timed(fn, reps=20):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


pitch = n * 8
for w in (8, 16):
    x_h2d = timed(lambda: rt.cudaMemcpy2DAsync(dev.data_ptr() + 8, pitch, host.data_ptr() + 8, pitch, w, n * n, H2D, s))
    x_d2h = timed(lambda: rt.cudaMemcpy2DAsync(host.data_ptr(), pitch, dev.data_ptr(), pitch, w, n * n, D2H, s))
    print(f"x-slab ({w}-byte rows x {n*n}, pitch {pitch}): H2D {x_h2d:.1f} us, D2H {x_d2h:.1f} us")
y_h2d = timed(lambda: rt.cudaMemcpy2DAsync(dev.data_ptr() + pitch, pitch * n, host.data_ptr() + pitch, pitch * n, pitch, n, H2D, s))
print(f"y-slab (2 KiB rows x {n}): H2D {y_h2d:.1f} us")
c_h2d = timed(lambda: dev[:n * n].copy_(host[:n * n], non_blocking=True))
c_d2h = timed(lambda: host[:n * n].copy_(dev[:n * n], non_blocking=True))
print(f"contiguous 512 KiB: H2D {c_h2d:.1f} us, D2H {c_d2h:.1f} us")
