"""Multi-rank GPU parity worker: run under torchrun with one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/mp_worker.py

Every rank drives libigg_b200 through the public Python API (-> ctypes -> C ABI -> CUDA kernel with CUDA-IPC peer
stores); rank 0 gathers all fields and compares them bitwise with the oracle simulating the same N ranks.
Prints `MP_WORKER_OK <N>` on success.  (Test infrastructure: imports the oracle as the checker.)
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "implicitglobalgrid.jl_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import igg_b200 as igg  # noqa: E402
from oracle import igg_oracle as O  # noqa: E402

TDT = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}
# The IPC mode under test: contexts start in the one IGG_MODE names ("fence", "ll128", else the default "auto", which
# picks one of the two per call by size -- run it with a small IGG_BULK_THRESHOLD so that both get used); after a
# detour through the NCCL comparison path the worker goes back to the same one.
IPC_MODE = os.environ.get("IGG_MODE") if os.environ.get("IGG_MODE") in ("fence", "ll128") else "auto"


def case_list(n):
    nx, ny, nz = 13, 9, 10
    cases = [
        # (grid size, kwargs, shapes, per-field hw or None, elsize, dims, repeats)
        ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz)], None, 8, (3, 1, 2), 3),
        ((nx, ny, nz), dict(), [(nx, ny, nz), (nx + 1, ny, nz), (nx, ny + 1, nz), (nx, ny, nz + 1)], None, 8, (3, 1, 2), 2),
        ((nx, ny, nz), dict(periodx=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
         [(nx, ny, nz), (nx + 1, ny, nz), (nx - 1, ny + 2, nz + 1)], None, 4, (3, 1, 2), 2),
        ((nx, ny, nz), dict(periody=1, overlaps=(6, 4, 4), halowidths=(3, 1, 2)), [(nx, ny, nz), (nx, ny, nz + 1)], None, 2, (1, 3), 2),
        ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2)),
         [(nx, ny, nz), (nx, ny, nz), (nx + 1, ny, nz)], [(2, 2, 2), (1, 1, 2), (2, 1, 1)], 8, (3, 1, 2), 2),
        ((40, 30, 1), dict(periodx=1, periody=1), [(40, 30), (41, 30)], None, 8, (3, 1, 2), 2),
        ((64, 1, 1), dict(periodx=1), [(64,), (65,)], None, 4, (3, 1, 2), 2),
        # staging growth: a much bigger field after the small ones, then small again
        ((96, 80, 72), dict(periodx=1, periody=1, periodz=1), [(96, 80, 72), (97, 80, 72)], None, 8, (3, 1, 2), 3),
        ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz)], None, 1, (2, 1), 2),
        # rows of 1 KiB and more: with fields from igg.alloc_field the y/z faces go straight into the neighbour's halo
        ((160, 12, 10), dict(periodx=1, periody=1, periodz=1), [(160, 12, 10), (161, 12, 10)], None, 8, (3, 1, 2), 3),
        ((260, 12, 10), dict(periody=1, overlaps=(2, 4, 2), halowidths=(1, 2, 1)), [(260, 12, 10), (260, 13, 10), (260, 12, 11)], None, 4, (3, 1, 2), 2),
        ((256, 12, 10), dict(periodx=1, periody=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(256, 12, 10), (256, 13, 10), (256, 12, 11)], None, 4, (3, 1, 2), 2),
    ]
    return cases


def main():
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    me, n = dist.get_rank(), dist.get_world_size()
    ok = True
    # Committed golden vectors for this world size (tests/golden/, closed form, built without the oracle): every
    # rank compares its own fields with the expected bits, twice in a row from fresh inputs.
    import golden_cases as G
    for gi, case in G.cases(n):
        igg.init_global_grid(*case["grid"], quiet=True, **G.kwargs(case))
        # The bootstrap's LL128 link check + self-test: passed (and really ran) unless the run forces it to fail
        # (IGG_LL128_SELFTEST_FAIL=1), in which case the context must have fallen back to the fence protocol.
        import ctypes as C
        from igg_b200 import _lib as L
        st = (C.c_uint64 * 32)()
        L.check(L.load().igg_debug_state(igg.global_grid().ctx, st))
        mode, ll_ok, checked = int(st[31]) & 0xff, (int(st[31]) >> 8) & 1, int(st[31]) >> 16
        if os.environ.get("IGG_LL128_SELFTEST_FAIL") == "1":
            if ll_ok or mode != L.IGG_MODE_IPC_1PHASE:
                ok = False
                print(f"FALLBACK NOT TAKEN rank {me}: mode {mode}, ll128_ok {ll_ok}", flush=True)
        elif not ll_ok or checked < 4096 or mode != {"fence": 0, "ll128": 2, "auto": 3}[IPC_MODE]:
            ok = False
            print(f"LL128 SELF-TEST rank {me}: mode {mode}, ll128_ok {ll_ok}, {checked} line images checked", flush=True)
        hws = G.halowidths(case, igg.global_grid().halowidths)
        for rep in range(2):
            fields = [igg.from_numpy(a) for a in G.arrays(gi, me, "in")]
            igg.update_halo(*[(t, hw) for t, hw in zip(fields, hws)], dims=tuple(case["dims"]), blocking=(rep == 0))
            torch.cuda.synchronize()
            for i, (t, want) in enumerate(zip(fields, G.arrays(gi, me, "out"))):
                if not np.array_equal(igg.to_numpy(t).view(want.dtype), want):
                    ok = False
                    print(f"GOLDEN MISMATCH {case['name']} rank {me} field {i} rep {rep}", flush=True)
        igg.finalize_global_grid()
        dist.barrier()
    # IGG_WORKER_QUICK=1: only the checks that need no oracle run on rank 0 (golden vectors above, full-size
    # properties below) -- for short confirmation runs on many GPUs.
    quick = os.environ.get("IGG_WORKER_QUICK") == "1"
    for ci, (gsz, kw, shapes, hws, elsize, dims, reps) in enumerate([] if quick else case_list(n)):
        try:
            grids = O.make_grids(n, *gsz, **kw)
        except O.IGGError:
            continue
        igg.init_global_grid(*gsz, quiet=True, **kw)
        g = igg.global_grid()
        assert g.dims == grids[me].dims and g.coords == grids[me].coords and g.neighbors == grids[me].neighbors
        hws_ = hws or [g.halowidths] * len(shapes)
        base = [O.random_bits_field(s, elsize, 1000 * ci + 10 * me + i) for i, s in enumerate(shapes)]

        def load(tensors, rep):
            """Fresh data before every call: rep >= 0 -> throw-away data that differs from call to call (a stale read
            of an earlier call's staging would show), rep < 0 -> the data the oracle is run on."""
            torch.cuda.synchronize()
            for i, t in enumerate(tensors):
                a = base[i] if rep < 0 else O.random_bits_field(shapes[i], elsize, 777 + 1000 * ci + 100 * rep + 10 * me + i)
                t.copy_(igg.from_numpy(a))
            torch.cuda.synchronize()

        # the NCCL comparison path (reference-style dependent phases) must give the same bits as the fused kernel
        if ci in (0, 1, 2, 4, 5):
            igg.set_exchange_mode("nccl")
            tn = [igg.from_numpy(a) for a in base]
            for rep in (0, -1):
                load(tn, rep)
                igg.update_halo(*[(t, hw) for t, hw in zip(tn, hws_)], dims=dims, blocking=(rep == 0))
            torch.cuda.synchronize()
            igg.set_exchange_mode(IPC_MODE)
        else:
            tn = None
        ts = [igg.from_numpy(a) for a in base]
        for rep in list(range(reps)) + [-1]:
            load(ts, rep)
            igg.update_halo(*[(t, hw) for t, hw in zip(ts, hws_)], dims=dims, blocking=(rep % 2 == 0))
        torch.cuda.synchronize()
        # The device-resident exchange state must have advanced once per call for every neighbour this rank
        # exchanges with (it selects the staging half; stuck at 0 = no double buffering = a fast neighbour can
        # overwrite staging that is still being unpacked).
        import ctypes as C
        from igg_b200 import _lib as L
        st = (C.c_uint64 * 32)()
        L.check(L.load().igg_debug_state(g.ctx, st))
        counts = sorted(set(int(v) for v in st[1:28]))
        if n > 1 and (int(st[0]) < reps + 1 or counts != [0, reps + 1]):
            ok = False
            print(f"EXCHANGE STATE case {ci} rank {me}: epoch {int(st[0])} per-neighbour counts {counts}, "
                  f"want [0, {reps + 1}]", flush=True)
        if ci in (0, 2):
            # the same update captured into a CUDA graph and replayed (exchange state lives on the device)
            tg = [igg.from_numpy(a) for a in base]
            s = torch.cuda.Stream()
            call = igg.prepare_update_halo(*[(t, hw) for t, hw in zip(tg, hws_)], dims=dims, stream=s, blocking=False)
            torch.cuda.synchronize()
            dist.barrier()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s):
                call()
                call()
            for rep in (0, 1, -1):
                load(tg, rep)
                graph.replay()
            torch.cuda.synchronize()
            for i, (a, b) in enumerate(zip(tg, ts)):
                if not torch.equal(a, b):
                    ok = False
                    print(f"GRAPH/STREAM MISMATCH case {ci} rank {me} field {i}", flush=True)
        # The same calls on fields in page-locked HOST memory (zero-copy over PCIe).  Rows that start on 64-byte
        # boundaries exchange their x-faces through cached whole lines (fetch kernel, exchange on the cache,
        # write-back kernel); the others go cell by cell.  Same bits as the device-resident result.
        if ci in (0, 2, 4) or ci >= len(case_list(n)) - 3:
            th = [igg.zeros(s_, dtype=TDT[elsize], pin_memory=True) for s_ in shapes]
            for rep in list(range(reps)) + [-1]:
                load(th, rep)
                igg.update_halo(*[(t, hw) for t, hw in zip(th, hws_)], dims=dims, blocking=(rep % 2 == 0))
            torch.cuda.synchronize()
            for i, (a, b) in enumerate(zip(th, ts)):
                if not torch.equal(a, b.cpu()):
                    ok = False
                    print(f"HOST-FIELD MISMATCH case {ci} rank {me} field {i}: {int((a != b.cpu()).sum())} cells", flush=True)
            del th
        # The same calls on fields inside symmetric allocations (igg.alloc_field): boxes that span x with rows of
        # 1 KiB or more are stored straight into the neighbour's field (direct-to-halo), the rest is staged as before.
        # Same bits as the staged result, which is checked against the oracle below.
        td = [igg.alloc_field(s_, dtype=TDT[elsize]) for s_ in shapes]
        for rep in list(range(reps)) + [-1]:
            load(td, rep)
            igg.update_halo(*[(t, hw) for t, hw in zip(td, hws_)], dims=dims, blocking=(rep % 2 == 0))
        torch.cuda.synchronize()
        for i, (a, b) in enumerate(zip(td, ts)):
            if not torch.equal(a, b):
                ok = False
                print(f"DIRECT/STAGED MISMATCH case {ci} rank {me} field {i}: {int((a != b).sum())} cells", flush=True)
        dist.barrier()
        for t in td:
            igg.free_field(t)
        del td
        if tn is not None:
            for i, (a, b) in enumerate(zip(tn, ts)):
                if not torch.equal(a, b):
                    ok = False
                    print(f"NCCL/IPC MISMATCH case {ci} rank {me} field {i}", flush=True)
        mine = [igg.to_numpy(t).view(base[i].dtype) for i, t in enumerate(ts)]
        out = [None] * n
        dist.all_gather_object(out, mine)
        if me == 0:
            want = [[O.random_bits_field(s, elsize, 1000 * ci + 10 * r + i) for i, s in enumerate(shapes)] for r in range(n)]
            O.update_halo(grids, [[(a, hw) for a, hw in zip(fr, hws_)] for fr in want], dims=dims)
            for r in range(n):
                for i in range(len(shapes)):
                    if not np.array_equal(out[r][i], want[r][i]):
                        ok = False
                        bad = np.argwhere(out[r][i] != want[r][i])
                        print(f"MISMATCH case {ci} rank {r} field {i}: {len(bad)} cells, first {bad[:4].tolist()}", flush=True)
        # gather! of field 0 onto root 0 and onto the last rank (test/test_gather.jl:36-69,126-137)
        for root in sorted({0, n - 1}):
            A = ts[0]
            dimsg = g.dims[:A.dim()]
            if any(d > 1 for d in g.dims[A.dim():]):
                continue
            G = igg.zeros(tuple(s * d for s, d in zip(A.shape, dimsg)), dtype=A.dtype) if me == root else None
            igg.gather(A, G, root=root)
            blocks = [None] * n
            dist.all_gather_object(blocks, igg.to_numpy(A))
            # the same into page-locked HOST memory on root: A_global cannot be exported over CUDA IPC, so the blocks go
            # through the context's device image of A_global and root copies it out (the path kept for such memory)
            GH = igg.zeros(tuple(s * d for s, d in zip(A.shape, dimsg)), dtype=A.dtype, pin_memory=True) if me == root else None
            igg.gather(A, GH, root=root)
            if me == root:
                wantG = np.zeros(tuple(G.shape), dtype=blocks[0].dtype, order="F")
                O.gather(grids, blocks, wantG, root=root)
                if not np.array_equal(igg.to_numpy(G), wantG):
                    ok = False
                    print(f"GATHER MISMATCH case {ci} root {root}", flush=True)
                if not np.array_equal(igg.to_numpy(GH), wantG):
                    ok = False
                    print(f"GATHER (host A_global) MISMATCH case {ci} root {root}", flush=True)
        igg.finalize_global_grid()
        dist.barrier()
    # Calls with different field sets and `dims=` alternating on ONE context: every neighbour offset keeps its own
    # staging / flag region across plans (a neighbour that is a call ahead must not write where another neighbour's
    # data is still being unpacked).  Fresh data before every call, every call checked against the oracle.
    if not quick:
        gsz, kw = (13, 9, 10), dict(periodx=1, periody=1, periodz=1)
        grids = O.make_grids(n, *gsz, **kw)
        igg.init_global_grid(*gsz, quiet=True, **kw)
        # D is big enough for the fence protocol when the worker runs in "auto" mode with IGG_BULK_THRESHOLD=16384: the
        # two protocols then alternate on one context (each has staging regions of its own).
        shp = {"A": (13, 9, 10), "B": (14, 9, 10), "C": (13, 9, 11), "D": (13, 64, 48)}
        dev = {k: igg.zeros(s, dtype=torch.int64) for k, s in shp.items()}
        seq = [("A", (1,)), ("AB", (3, 1, 2)), ("B", (2,)), ("A", (3, 1, 2)), ("ABC", (1, 3)), ("C", (3,)),
               ("BC", (2, 1)), ("A", (1,)), ("ABC", (3, 1, 2)), ("B", (3, 1, 2))] * 2
        seq += [("D", (3, 1, 2)), ("A", (3, 1, 2)), ("D", (1,)), ("D", (3, 1, 2)), ("A", (1,)), ("D", (2, 3))]
        for k, (names, dims) in enumerate(seq):
            host = {nm: O.random_bits_field(shp[nm], 8, 5000 + 100 * k + 10 * me + i) for i, nm in enumerate(names)}
            torch.cuda.synchronize()
            for nm in names:
                dev[nm].copy_(igg.from_numpy(host[nm]))
            torch.cuda.synchronize()
            igg.update_halo(*[dev[nm] for nm in names], dims=dims, blocking=(k % 3 != 1))
            torch.cuda.synchronize()
            out = [None] * n
            dist.all_gather_object(out, [igg.to_numpy(dev[nm]).view(np.uint64) for nm in names])
            if me == 0:
                want = [[O.random_bits_field(shp[nm], 8, 5000 + 100 * k + 10 * r + i) for i, nm in enumerate(names)] for r in range(n)]
                O.update_halo(grids, want, dims=dims)
                for r in range(n):
                    for i in range(len(names)):
                        if not np.array_equal(out[r][i], want[r][i]):
                            ok = False
                            print(f"ALTERNATING-PLANS MISMATCH call {k} ({names}, dims {dims}) rank {r} field {i}", flush=True)
        igg.finalize_global_grid()
        dist.barrier()

    # BASELINE config 2 at full size (256^3 f64, periodic): too big to ship to rank 0, so size-independent checks:
    # (a) the fused path equals the NCCL path (itself checked against the oracle above) bit for bit, on the device;
    # (b) after the update every halo plane equals the neighbour's send plane (checked through torch.distributed).
    N = 256
    igg.init_global_grid(N, N, N, periodx=1, periody=1, periodz=1, quiet=True)
    g = igg.global_grid()
    gen = torch.Generator(device="cuda")
    gen.manual_seed(4321 + me)
    a = igg.zeros((N, N, N), dtype=torch.int64)
    a.copy_(torch.randint(-2**62, 2**62, (N, N, N), dtype=torch.int64, device="cuda", generator=gen))
    b = a.clone(memory_format=torch.preserve_format)
    c = igg.alloc_field((N, N, N), dtype=torch.int64)       # symmetric allocation: y/z faces direct-to-halo
    c.copy_(a)
    igg.update_halo(a)
    igg.update_halo(c, blocking=False)
    igg.update_halo(c)
    igg.set_exchange_mode("nccl")
    igg.update_halo(b)
    igg.set_exchange_mode(IPC_MODE)
    torch.cuda.synchronize()
    if not torch.equal(a, b):
        ok = False
        print(f"FULL-SIZE fused/NCCL MISMATCH rank {me}", flush=True)
    if not torch.equal(a, c):
        ok = False
        print(f"FULL-SIZE direct/staged MISMATCH rank {me}: {int((a != c).sum())} cells", flush=True)
    dist.barrier()
    igg.free_field(c)
    del c
    for d in range(3):
        right = g.neighbors[1][d]
        left = g.neighbors[0][d]
        mine_send = a.narrow(d, N - 2, 1).contiguous().cpu()        # my high-side send plane (ol=2, hw=1)
        theirs = torch.empty_like(mine_send)
        if right == me:
            theirs = mine_send
        else:
            reqs = [dist.isend(mine_send, right, tag=10 + d), dist.irecv(theirs, left, tag=10 + d)]
            for r_ in reqs:
                r_.wait()
        if not torch.equal(a.narrow(d, 0, 1).contiguous().cpu(), theirs):   # my low halo == left neighbour's high send
            ok = False
            print(f"FULL-SIZE halo/neighbour MISMATCH rank {me} dim {d}", flush=True)
    del a, b
    igg.finalize_global_grid()
    dist.barrier()

    # All ranks in a row along z, four full-size fields (one staggered in x: its rows are not 16-byte aligned): the
    # contiguous z-faces cross NVLink -- in the fence protocol through the bulk-copy pipeline (tiles of 16 KiB and
    # more) -- while x and y wrap inside each rank.  Fused path vs NCCL path (checked against the oracle above).
    # (x halo of two cells = 16 bytes, so that the z-face boxes, which leave the x halo out, keep 16-byte alignment.)
    igg.init_global_grid(N, N, N, dimx=1, dimy=1, dimz=n, periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2),
                         halowidths=(2, 1, 1), quiet=True)
    shapes4 = [(N, N, N), (N, N, N), (N + 1, N, N), (N, N, N)]
    fa = [igg.zeros(s_, dtype=torch.int64) for s_ in shapes4]
    for t in fa:
        t.copy_(torch.randint(-2**62, 2**62, tuple(t.shape), dtype=torch.int64, device="cuda", generator=gen))
    fb = [t.clone(memory_format=torch.preserve_format) for t in fa]
    fc = [igg.alloc_field(s_, dtype=torch.int64) for s_ in shapes4]     # symmetric allocations: z-faces direct-to-halo
    for x, y in zip(fc, fa):
        x.copy_(y)
    igg.update_halo(*fa, blocking=False)
    igg.update_halo(*fa)                          # twice: both staging halves, halos must not change
    for rep in range(3):
        igg.update_halo(*fc, blocking=(rep == 2))
    torch.cuda.synchronize()
    for i, (x, y) in enumerate(zip(fc, fa)):
        if not torch.equal(x, y):
            ok = False
            print(f"FULL-SIZE z-row direct/staged MISMATCH rank {me} field {i}: {int((x != y).sum())} cells", flush=True)
    dist.barrier()
    for x in fc:
        igg.free_field(x)
    del fc
    igg.set_exchange_mode("nccl")
    igg.update_halo(*fb)
    igg.set_exchange_mode(IPC_MODE)
    torch.cuda.synchronize()
    for i, (x, y) in enumerate(zip(fa, fb)):
        if not torch.equal(x, y):
            ok = False
            print(f"FULL-SIZE z-row fused/NCCL MISMATCH rank {me} field {i}: {int((x != y).sum())} cells", flush=True)
    del fa, fb
    igg.finalize_global_grid()
    dist.barrier()

    # BASELINE config 1 on several ranks: the diffusion3D example, fused CUDA step + update_halo!, bit for bit
    from igg_b200.diffusion3d import Diffusion3D
    from oracle import diffusion_oracle as D
    nd, nt = (40, 36, 33), 40
    grids, Ts, Cps, dxyz, dt = D.setup(n, nd)
    T0, Cp0 = Ts[me].copy(order="F"), Cps[me].copy(order="F")
    want = None if quick else D.run(grids, Ts, Cps, dxyz, dt, nt)[me]
    for overlap in (() if quick else (False, True, "graph")):
        sim = Diffusion3D(*nd, T=T0, Cp=Cp0)
        got = igg.to_numpy(sim.run(nt, overlap=False, graph=True) if overlap == "graph" else sim.run(nt, overlap=overlap))
        sim.finalize()
        if not np.array_equal(got.view(np.uint64), want.view(np.uint64)):
            ok = False
            print(f"DIFFUSION MISMATCH rank {me} overlap {overlap}: max abs diff {np.abs(got - want).max()}", flush=True)
        dist.barrier()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if me == 0 and int(flag) == 1:
        print(f"MP_WORKER_OK {n}", flush=True)
    dist.destroy_process_group()
    return 0 if int(flag) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
