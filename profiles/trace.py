"""Per-block timeline of one update_halo launch (igg_trace_enable / igg_trace_read), 1..N GPUs.

    python profiles/trace.py [n]                       # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 profiles/trace.py

Prints, per rank, when (ns after the earliest block start of that launch) blocks started, had their first tile record,
issued their first copy, signalled their pushes, saw their first arrival flag and ended.
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import numpy as np
import torch
import torch.distributed as dist
import igg_b200 as igg
from igg_b200 import _lib as L

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
g = igg.global_grid()
lib = L.load()
ring = [igg.zeros((n, n, n)) for _ in range(24)]
for t in ring:
    t.copy_(torch.rand_like(t))
for i in range(50):
    igg.update_halo(ring[i % 24], blocking=False)
torch.cuda.synchronize()
L.check(lib.igg_trace_enable(g.ctx, 1))
sm_ghz = torch.cuda.clock_rate() / 1e6 if hasattr(torch.cuda, "clock_rate") else 1.965
rows = []
for rep in range(5):
    if world > 1:
        dist.barrier()
    igg.update_halo(ring[(50 + rep) % 24])
    buf = (C.c_uint64 * (8 * 1024))()
    nb = C.c_int(0)
    L.check(lib.igg_trace_read(g.ctx, buf, 1024, C.byref(nb)))
    a = np.frombuffer(buf, dtype=np.uint64, count=8 * nb.value).reshape(nb.value, 8).astype(np.int64)
    t0 = a[:, 0].min()
    start = a[:, 0] - t0
    def rel(k):
        v = a[:, k]
        ok = v > 0
        return (start[ok] + (v[ok] - a[ok, 1]) / 1.965), ok
    out = {"rep": rep, "blocks": int(nb.value), "start_max": int(start.max())}
    for k, name in ((2, "tile_loaded"), (3, "copy_issued"), (4, "signalled"), (5, "flag_seen"), (6, "end")):
        v, ok = rel(k)
        if ok.any():
            out[name] = (int(v.min()), int(np.median(v)), int(v.max()), int(ok.sum()))
    # per tile class: kind (0 push, 1 local, 2 unpack) x narrow (strided x-face rows) / wide
    kind = a[:, 7] & 0xff
    ncols = (a[:, 7] >> 40) & 0xffffff
    nrows = (a[:, 7] >> 16) & 0xffffff
    narrow = (ncols <= 4) & (nrows > 1)
    cyc = lambda k0, k1, m: ((a[m, k1] - a[m, k0]) / 1.965)
    for kd, kn in ((0, "push"), (1, "local"), (2, "unpack")):
        for nw, nn in ((True, "xface"), (False, "wide")):
            m = (kind == kd) & (narrow == nw) & (a[:, 6] > 0)
            if not m.any():
                continue
            d = {"n": int(m.sum()), "start": int(np.median(start[m])),
                 "rec->copy": int(np.median(cyc(2, 3, m))), "rec->copy_max": int(cyc(2, 3, m).max())}
            if kd == 0:
                d["copy->signalled"] = int(np.median(cyc(3, 4, m)))
                d["copy->signalled_max"] = int(cyc(3, 4, m).max())
            if kd == 2:
                d["start->flag"] = int(np.median(cyc(1, 5, m)))
                d["flag->end"] = int(np.median(cyc(5, 6, m)))
            out[f"{kn}_{nn}"] = d
    rows.append(out)
if world > 1:
    allrows = [None] * world
    dist.all_gather_object(allrows, rows)
else:
    allrows = [rows]
if rank == 0:
    for r, rr in enumerate(allrows):
        for o in rr[2:]:
            print(f"rank {r}", o)
igg.finalize_global_grid()
if world > 1:
    dist.destroy_process_group()
