"""Host cost of one stream-ordered update_halo call, layer by layer (1 GPU, cfg2): un-prepared Python call, prepared
call object, and the C side alone (igg_debug_issue_loop: n calls issued from C)."""
import ctypes as C
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import torch
import igg_b200 as igg
from igg_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
t = igg.zeros((n, n, n))
s = torch.cuda.Stream()
call = igg.prepare_update_halo(t, stream=s, blocking=False)
lib = L.load()
K = 3000
def timeit(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        f()
    dt = (time.perf_counter() - t0) / K * 1e6
    torch.cuda.synchronize()
    return dt
# the device is slower than the host here, so the launch queue fills up: issue cost includes back-pressure once it does;
# keep K small enough?  (the queue holds ~1000 launches) -- report both a short and a long burst
for burst in (500, 3000):
    K = burst
    print(f"burst {burst}: un-prepared igg.update_halo {timeit(lambda: igg.update_halo(t, stream=s, blocking=False)):.2f} us/call, "
          f"prepared call object {timeit(call):.2f} us/call")
ns = C.c_double(0)
for burst in (500, 3000):
    torch.cuda.synchronize()
    L.check(lib.igg_debug_issue_loop(call._h, C.c_void_p(s.cuda_stream), burst, C.byref(ns)))
    torch.cuda.synchronize()
    print(f"burst {burst}: C side alone {ns.value / 1e3:.2f} us/call")
call.release()
igg.finalize_global_grid()
