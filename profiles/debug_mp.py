"""Debug helper (multi-rank): one periodic field, IPC mode only, DIFFERENT data in every call (so that a stale read
of an earlier call's staging shows up); after every update compare with the oracle and print, per rank, how many
cells are wrong, a sample, and the device exchange state.   usage: debug_mp.py [ncalls] [nx ny nz]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "implicitglobalgrid.jl_b200"), ROOT):
    sys.path.insert(0, p)
import numpy as np
import torch
import torch.distributed as dist
import igg_b200 as igg
from igg_b200 import _lib as L
from oracle import igg_oracle as O

torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("gloo")
me, n = dist.get_rank(), dist.get_world_size()
ncalls = int(sys.argv[1]) if len(sys.argv) > 1 else 4
gsz = tuple(int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (13, 9, 10)
kw = dict(periodx=1, periody=1, periodz=1)
grids = O.make_grids(n, *gsz, **kw)
igg.init_global_grid(*gsz, quiet=True, **kw)
g = igg.global_grid()
lib = L.load()
t = igg.zeros(gsz, dtype=torch.int64)
total_bad = 0
for call in range(ncalls):
    base = O.random_bits_field(gsz, 8, 1000 * call + 10 * me)
    want = [O.random_bits_field(gsz, 8, 1000 * call + 10 * r) for r in range(n)]
    O.update_halo(grids, [[w] for w in want])
    t.copy_(igg.from_numpy(base))
    torch.cuda.synchronize()
    dist.barrier()
    igg.update_halo(t, blocking=(call % 3 != 2))
    torch.cuda.synchronize()
    st = (C.c_uint64 * 32)()
    L.check(lib.igg_debug_state(g.ctx, st))
    got = igg.to_numpy(t).view(np.uint64)
    bad = np.argwhere(got != want[me])
    total_bad += len(bad)
    sample = [hex(int(got[tuple(b)])) for b in bad[:2]]
    msg = f"call {call} rank {me}: epoch {st[0]} bad {len(bad)} sample {sample}"
    out = [None] * n
    dist.all_gather_object(out, msg)
    if me == 0 and (call < 6 or any(" bad 0 " not in m for m in out)):
        print("\n".join(out), flush=True)
tb = torch.tensor([total_bad])
dist.all_reduce(tb)
if me == 0:
    print(f"TOTAL_BAD {int(tb)} over {ncalls} calls, fence_mode {os.environ.get('IGG_FENCE_MODE', '0')}", flush=True)
igg.finalize_global_grid()
dist.destroy_process_group()
