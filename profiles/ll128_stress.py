"""LL128 under stress on real hardware: many exchanges with data that changes every call, random per-rank delays in
front of every launch (rank skew), every cell of every field checked on the device after every call.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/ll128_stress.py [calls] [n]

The field holds a function of each cell's GLOBAL index and of the call number, so every rank knows in closed form
what its whole array (halo included: the grid is periodic) must hold after the update; before the update the halo
cells are overwritten with a poison value.  A torn or stale staging line shows up as a wrong halo cell.  Nothing is
read back to the host until the end.  Prints `LL128_STRESS_OK <calls> calls x <N> ranks` or the mismatch counts.
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import numpy as np
import torch
import torch.distributed as dist
import igg_b200 as igg

calls = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 48
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
g = igg.global_grid()
# global index of every local cell (ol = 2, hw = 1, periodic: local cell i is global cell coord*(n-2) + i - 1, wrapped)
ng = [g.dims[d] * (n - 2) for d in range(3)]
gi = [((g.coords[d] * (n - 2) + torch.arange(n, dtype=torch.int64, device="cuda") - 1) % ng[d]) for d in range(3)]
val = (gi[0][:, None, None] + ng[0] * (gi[1][None, :, None] + ng[1] * gi[2][None, None, :])) * 1000003
interior = torch.zeros((n, n, n), dtype=torch.bool, device="cuda")
interior[1:-1, 1:-1, 1:-1] = True
poison = torch.full((), -1, dtype=torch.int64, device="cuda")


def cm(t):          # a column-major array with the values of t
    a = igg.zeros((n, n, n), dtype=t.dtype)
    a.copy_(t)
    return a


val, interior = cm(val), cm(interior)
A, B = igg.zeros((n, n, n), dtype=torch.int64), igg.zeros((n, n + 1, n), dtype=torch.int64)      # B: staggered in y
valB = igg.zeros((n, n + 1, n), dtype=torch.int64)
giyB = (g.coords[1] * (n - 2) + torch.arange(n + 1, dtype=torch.int64, device="cuda") - 1) % (g.dims[1] * (n - 2))   # ol_y = 3
valB.copy_((gi[0][:, None, None] + ng[0] * (giyB[None, :, None] + ng[1] * gi[2][None, None, :])) * 1000003 + 17)
interiorB = igg.zeros((n, n + 1, n), dtype=torch.bool)
interiorB[1:-1, 1:-1, 1:-1] = True
E, EB = torch.empty_like(A), torch.empty_like(B)
bad = torch.zeros((), dtype=torch.int64, device="cuda")
rng = np.random.default_rng(99 + rank)
call = igg.prepare_update_halo(A, B, blocking=False)          # warm: grows the staging, collective
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.time()
for c in range(1, calls + 1):
    torch.add(val, c * 7919, out=E)
    torch.add(valB, c * 104729, out=EB)
    torch.where(interior, E, poison, out=A)
    torch.where(interiorB, EB, poison, out=B)
    if c % 3:
        torch.cuda._sleep(int(rng.integers(0, 30000)))          # up to ~15 us of skew against the neighbours
    call()
    bad += (A != E).sum() + (B != EB).sum()
    if c % 50000 == 0 and rank == 0:
        print(f"  {c} calls, {time.time() - t0:.0f} s", flush=True)
torch.cuda.synchronize()
nbad = int(bad.item())
allbad = [None] * world
if world > 1:
    dist.all_gather_object(allbad, nbad)
else:
    allbad = [nbad]
call.release()
igg.finalize_global_grid()
if rank == 0:
    mode = os.environ.get("IGG_MODE", "auto")
    if sum(allbad) == 0:
        print(f"LL128_STRESS_OK {calls} calls x {world} ranks, mode {mode}, {n}^3 + staggered field per rank, "
              f"{time.time() - t0:.0f} s, every cell checked on the device after every call", flush=True)
    else:
        print(f"LL128_STRESS_FAILED mismatching cells per rank: {allbad}", flush=True)
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if sum(allbad) == 0 else 1)
