"""CPU-side checks of the C-ABI boundary (no GPU): the library loads, exports every symbol include/igg_b200.h
declares, refuses to run the data path without a CUDA device, and its host-side pieces work across real processes
(shared-memory control plane; a world_size-2 gloo run that executes each rank's plan and ships the PUSH bytes with
torch.distributed send/recv, checked against the oracle)."""
import ctypes as C
import multiprocessing as mp
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from igg_b200 import _lib as L
from oracle import igg_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "igg_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(igg_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = L.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"libigg_b200.so does not export {name}"
    assert declared == set(L.DECLARED_SYMBOLS), declared ^ set(L.DECLARED_SYMBOLS)
    assert lib.igg_version() == 200
    want = b"igg_halo_exchange_kernel" if os.environ.get("IGG_MODE") in ("fence", "ipc") else b"igg_halo_exchange_ll128_kernel"
    assert lib.igg_kernel_name() == want
    assert lib.igg_bootstrap_size() == 256


def test_no_cpu_fallback():
    """Without a GPU the product path must fail loudly, not fall back to anything."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    (g,) = O.make_grids(1, 8, 8, 8, periodx=1)
    d = L.make_desc(g.nxyz, g.dims, g.coords, g.neighbors, g.periods, g.overlaps, g.halowidths, 0, 1, 0)
    ctx = C.c_void_p()
    rc = L.load().igg_create(C.byref(d), C.byref(ctx))
    assert rc == L.IGG_ERR_CUDA
    assert b"no CPU fallback" in L.load().igg_last_error()
    import igg_b200 as igg
    with pytest.raises(igg.IGGError, match="no CPU fallback"):
        igg.init_global_grid(8, 8, 8, quiet=True)
    assert not igg.grid_is_initialized()


def test_package_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "implicitglobalgrid.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".jl")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "igg_oracle" not in txt and "import oracle" not in txt and "from oracle" not in txt, f


def test_no_statement_hidden_in_a_trailing_comment():
    """A lost newline once glued `p.active_mask = ...;` onto the end of a // comment, which silently disabled the
    staging double buffering.  Flag any // comment that ends in what looks like a statement."""
    pat = re.compile(r"//.*\S\s{3,}[A-Za-z_][\w.\->\[\]]*\s*(=|\+=|-=)\s*[^=].*;\s*$")
    src = os.path.join(ROOT, "implicitglobalgrid.jl_b200", "csrc")
    bad = []
    for f in sorted(os.listdir(src)):
        if f.endswith((".cu", ".cpp", ".h")):
            for ln, line in enumerate(open(os.path.join(src, f), errors="replace"), 1):
                if pat.search(line):
                    bad.append(f"{f}:{ln}: {line.strip()[:100]}")
    assert not bad, "\n".join(bad)


def _ctrl_worker(me, nprocs, conn, rounds):
    sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
    from igg_b200 import _lib as LL
    lib = LL.load()
    name = C.create_string_buffer(64)
    h = C.c_void_p()
    rc = lib.igg_ctrl_selftest_create(me, nprocs, name, C.byref(h))
    conn.send((rc, name.raw))
    if rc:
        return
    names = conn.recv()
    rc = lib.igg_ctrl_selftest_run(h, me, nprocs, names, rounds)
    conn.send((rc, lib.igg_last_error()))


@pytest.mark.parametrize("nprocs", [2, 4])
def test_control_plane_across_processes(nprocs):
    ctx = mp.get_context("spawn")
    pipes, procs = [], []
    for r in range(nprocs):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ctrl_worker, args=(r, nprocs, b, 50))
        p.start()
        pipes.append(a)
        procs.append(p)
    names = b""
    for a in pipes:
        rc, raw = a.recv()
        assert rc == 0
        names += raw
    for a in pipes:
        a.send(names)
    for a in pipes:
        rc, msg = a.recv()
        assert rc == 0, msg
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    leftovers = [f for f in os.listdir("/dev/shm") if f.startswith("igg_b200_")]
    assert not leftovers, leftovers


GLOO_WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200")); sys.path.insert(0, ROOT)
from igg_b200 import _lib as L
from igg_b200 import topology as T
from oracle import igg_oracle as O

dist.init_process_group("gloo", init_method="env://")
me, n = dist.get_rank(), dist.get_world_size()
nx, ny, nz = 9, 7, 6
kw = dict(periodx=1, periody=1, periodz=0, overlaps=(4, 2, 2), halowidths=(2, 1, 1))
grids = O.make_grids(n, nx, ny, nz, **kw)                 # oracle view of all ranks (rank 0 checks with it)
# product-side topology of THIS rank
dims = T.dims_create(n, (0, 0, 0)); coords = T.cart_coords(me, dims)
nb = ([], [])
for d in range(3):
    l, r = T.cart_shift(coords, dims, (1, 1, 0), d, 1); nb[0].append(l); nb[1].append(r)
assert dims == grids[me].dims and coords == grids[me].coords and (tuple(nb[0]), tuple(nb[1])) == grids[me].neighbors
desc = L.make_desc((nx, ny, nz), dims, coords, nb, (1, 1, 0), (4, 2, 2), (2, 1, 1), me, n, 0)
shapes = [(nx, ny, nz), (nx + 1, ny, nz)]
fields = [O.random_bits_field(s, 8, 77 + 10 * me + i) for i, s in enumerate(shapes)]
mem = [f.reshape(-1, order="F").view(np.uint8) for f in fields]
cf = [L.make_field(0x10000 * (i + 1), s, (2, 1, 1), 8) for i, s in enumerate(shapes)]
ops, stg = L.plan_describe(desc, cf)
staging = np.full(max(stg, 1), 0xEE, np.uint8)

def rows(op):
    for z in range(op.nz):
        for y in range(op.ny):
            yield op.src_off + y * op.src_sy + z * op.src_sz, op.dst_off + y * op.dst_sy + z * op.dst_sz

# PUSH: pack into a message per (peer, offset-op), ship with gloo; the receiver places it at the staging offsets
sends = []
for k, op in enumerate(ops):
    if op.kind == L.IGG_OP_PUSH:
        payload = np.concatenate([mem[op.field][s:s + op.row_bytes] for s, _ in rows(op)])
        meta = torch.tensor([op.dst_off, op.dst_sy, op.dst_sz, op.row_bytes, op.ny, op.nz], dtype=torch.int64)
        sends.append((op.peer, meta, torch.from_numpy(payload.copy())))
    elif op.kind == L.IGG_OP_LOCAL:
        for s, d in rows(op):
            mem[op.field][d:d + op.row_bytes] = mem[op.field][s:s + op.row_bytes]
nrecv = sum(1 for op in ops if op.kind == L.IGG_OP_UNPACK)
reqs = [dist.isend(m, peer, tag=1) for peer, m, _ in sends]
peers_in = [op.peer for op in ops if op.kind == L.IGG_OP_UNPACK]
metas = []
for peer in peers_in:          # messages between a pair arrive in plan order on both sides
    m = torch.zeros(6, dtype=torch.int64); dist.recv(m, peer, tag=1); metas.append((peer, m))
for r in reqs: r.wait()
reqs = [dist.isend(p, peer, tag=2) for peer, _, p in sends]
for peer, m in metas:
    dst_off, dsy, dsz, rb, ny_, nz_ = [int(v) for v in m]
    buf = torch.zeros(rb * max(ny_, 1) * max(nz_, 1), dtype=torch.uint8); dist.recv(buf, peer, tag=2)
    b = buf.numpy(); k = 0
    for z in range(nz_):
        for y in range(ny_):
            o = dst_off + y * dsy + z * dsz
            staging[o:o + rb] = b[k:k + rb]; k += rb
for r in reqs: r.wait()
for op in ops:
    if op.kind == L.IGG_OP_UNPACK:
        for s, d in rows(op):
            mem[op.field][d:d + op.row_bytes] = staging[s:s + op.row_bytes]
# rank 0 collects and checks against the oracle
out = [None] * n
dist.all_gather_object(out, [f.copy() for f in fields])
if me == 0:
    want = [[O.random_bits_field(s, 8, 77 + 10 * r + i) for i, s in enumerate(shapes)] for r in range(n)]
    O.update_halo(grids, [[(a, (2, 1, 1)) for a in fr] for fr in want])
    for r in range(n):
        for i in range(len(shapes)):
            assert np.array_equal(out[r][i], want[r][i]), (r, i)
    print("GLOO_PLAN_OK")
dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [2])
def test_world_size_2_gloo_plan_execution(world, tmp_path):
    script = tmp_path / "gloo_worker.py"
    script.write_text(f"ROOT = {ROOT!r}\n" + GLOO_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400), str(script)],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "GLOO_PLAN_OK" in r.stdout
