"""Per-block timeline of single-GPU update_halo launches by tile kind (igg_trace_enable / igg_trace_read):
when blocks start, have their tile record, have their first tile's loads back (stores issued) and end.

    python profiles/trace1.py [n] [dims]      e.g.  python profiles/trace1.py 256 1,2,3
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import numpy as np
import torch
import igg_b200 as igg
from igg_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dims = tuple(int(x) for x in sys.argv[2].split(",")) if len(sys.argv) > 2 else (3, 1, 2)
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
g = igg.global_grid()
lib = L.load()
ring = [igg.zeros((n, n, n)) for _ in range(24)]
for t in ring:
    t.copy_(torch.rand_like(t))
for i in range(50):
    igg.update_halo(ring[i % 24], dims=dims, blocking=False)
torch.cuda.synchronize()
L.check(lib.igg_trace_enable(g.ctx, 1))
KIND = {1: "local", 8: "wrapx", 0: "push", 2: "unpack", 6: "push_ll", 7: "unpack_ll"}
for rep in range(4):
    igg.update_halo(ring[(50 + rep) % 24], dims=dims)
    buf = (C.c_uint64 * (8 * 1024))()
    nb = C.c_int(0)
    L.check(lib.igg_trace_read(g.ctx, buf, 1024, C.byref(nb)))
    a = np.frombuffer(buf, dtype=np.uint64, count=8 * nb.value).reshape(nb.value, 8).astype(np.int64)
    if rep < 2:
        continue
    t0 = a[:, 0].min()
    start = a[:, 0] - t0
    kind = a[:, 7] & 0xff
    ncols = (a[:, 7] >> 40) & 0xffffff
    print(f"rep {rep}: {nb.value} blocks, last start {int(start.max())} ns, last end "
          f"{int((start + (a[:, 6] - a[:, 1]) / 1.965).max())} ns")
    for kd in sorted(set(kind.tolist())):
        for narrow in (True, False):
            m = (kind == kd) & ((ncols <= 4) == narrow)
            if not m.any():
                continue
            rel = lambda k: start[m] + (a[m, k] - a[m, 1]) / 1.965
            q = lambda v: f"{int(np.percentile(v, 10))}/{int(np.median(v))}/{int(v.max())}"
            print(f"   {KIND.get(kd, kd)}{' narrow' if narrow else ''}: n={int(m.sum())} start {q(start[m])}  tile_rec {q(rel(2))}  "
                  f"first_tile_done {q(rel(3))}  end {q(rel(6))}   (p10/p50/max ns)")
igg.finalize_global_grid()
