"""All BASELINE.json configs beyond the headline one (which bench.py covers), 1..8 GPUs under torchrun:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/bench_configs.py [--out file]

cfg2   256^3 f64 periodic, one field                      (fused IPC kernel vs NCCL 3-phase comparison path)
cfg3i  256^3 f64 non-periodic, update_halo!(T)            (the diffusion3D example's exchange)
cfg3ii T, Vx, Vy, Vz staggered 256/257 f64 in one call, non-periodic and periodic
cfg4   512^3 f32, overlaps 4, halowidths 2, periodic, F = 1, 2, 4 fields per call
cfg5a  gather! of 256^3 f64 per GPU onto root
cfg5b  2-D 8192^2 f64 periodic, hw = 1, 2, 4, 8 (overlaps = 2 hw), F = 1 and 4

Device time per call from one CUDA graph of K calls on a ring of distinct field sets (cold L2), max over ranks.
Prints one JSON object per line; rank 0 writes them to --out as well.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import torch
import torch.distributed as dist
import igg_b200 as igg

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
ap.add_argument("--steps", type=int, default=400)
ap.add_argument("--only", default="")
ap.add_argument("--modes", default="ipc,nccl", help="comma list of ipc (= the default, auto), auto, ll128, fence, nccl")
ap.add_argument("--dims", default="", help="dimx,dimy,dimz to force a process grid (e.g. 1,1,2: the z-faces cross NVLink)")
ap.add_argument("--updims", default="", help="dims= of the update_halo calls, e.g. 3 or 3,2 (default: all)")
ap.add_argument("--sym", action="store_true", help="fields in symmetric allocations (igg.alloc_field): direct-to-halo")
ap.add_argument("--nvl", type=float, default=770.0, help="NVLink GB/s per direction used for the roofline")
args = ap.parse_args()
MODES = tuple(m for m in args.modes.split(",") if m)
UPDIMS = tuple(int(x) for x in args.updims.split(",")) if args.updims else (3, 1, 2)
FORCED = dict(zip(("dimx", "dimy", "dimz"), (int(x) for x in args.dims.split(",")))) if args.dims else {}

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
NVL, HBM = args.nvl, 6555.5
results = []


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def maxr(x):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def halo_bytes(g, shapes, hws, el):
    """sum over fields and dims with a halo of (#sides with a neighbour) * hw * prod(other sizes) * el;
    returns (all bytes sent, bytes sent to other ranks)."""
    tot = nvl = 0
    for shp, hw in zip(shapes, hws):
        s = tuple(shp) + (1,) * (3 - len(shp))
        for d in range(3):
            ol = g.overlaps[d] + (s[d] - g.nxyz[d])
            if ol < 2 * hw[d]:
                continue
            face = hw[d] * el
            for k in range(3):
                if k != d:
                    face *= s[k]
            for n in range(2):
                nb = g.neighbors[n][d]
                if nb < 0:
                    continue
                tot += face
                if nb != g.me:
                    nvl += face
    return tot, nvl


def time_update(name, gsz, kw, shapes, dtype, hws=None, ring_bytes=3.3e9, modes=None):
    if args.only and not any(o in name for o in args.only.split(",")):
        return
    modes = MODES if modes is None else tuple(m for m in MODES if m != "nccl")
    kw = dict(kw, **{k: v for k, v in FORCED.items() if gsz[("dimx", "dimy", "dimz").index(k)] > 1})
    igg.init_global_grid(*gsz, quiet=True, **kw)
    g = igg.global_grid()
    el = torch.empty((), dtype=dtype).element_size()
    hws_ = hws or [g.halowidths] * len(shapes)
    if halo_bytes(g, shapes, hws_, el)[0] == 0:      # e.g. one rank, nothing periodic: no exchange at all
        igg.finalize_global_grid()
        return
    set_bytes = sum(el * int(torch.tensor(s).prod()) for s in shapes)
    nring = max(2, min(24, int(ring_bytes // set_bytes)))
    ring = []
    for r in range(nring):
        fs = []
        for s in shapes:
            t = igg.alloc_field(s, dtype=dtype) if args.sym else igg.zeros(s, dtype=dtype)
            t.copy_(torch.randn(tuple(reversed(s)), device="cuda").to(dtype).permute(*reversed(range(len(s)))))
            fs.append(t)
        ring.append(fs)
    B, Bn = halo_bytes(g, shapes, hws_, el)
    if args.updims:          # only the updated dimensions count
        B = Bn = 0
        for shp, hw in zip(shapes, hws_):
            s3 = tuple(shp) + (1,) * (3 - len(shp))
            for d in [x - 1 for x in UPDIMS]:
                face = hw[d] * el
                for k in range(3):
                    if k != d:
                        face *= s3[k]
                for nn in range(2):
                    nb = g.neighbors[nn][d]
                    if nb >= 0:
                        B += face
                        Bn += face if nb != g.me else 0
    stream = torch.cuda.Stream()
    for mode in modes:
        if mode == "nccl" and world == 1:
            continue
        igg.set_exchange_mode(mode)
        calls = [igg.prepare_update_halo(*[(t, hw) for t, hw in zip(fs, hws_)], dims=UPDIMS, stream=stream, blocking=False) for fs in ring]
        with torch.cuda.stream(stream):
            for i in range(20):
                calls[i % nring]()
        barrier()
        K = args.steps
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        graph = None
        if mode != "nccl":
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                for i in range(K):
                    calls[i % nring]()
            graph.replay()
            barrier()
        with torch.cuda.stream(stream):
            e0.record(stream)
            if graph is not None:
                graph.replay()
            else:
                for i in range(K):
                    calls[i % nring]()
            e1.record(stream)
        barrier()
        us = maxr(e0.elapsed_time(e1)) * 1e3 / K
        t_roof = max(Bn / (NVL * 1e3), 2 * B / (HBM * 1e3)) if B else 0.0      # us
        rec = {"config": name + (f" dims={UPDIMS}" if args.updims else "") + (" [symmetric fields]" if args.sym else ""), "n_gpus": world, "dims": list(g.dims), "mode": mode, "fields": len(shapes),
               "update_halo_us": round(us, 2), "halo_bytes_per_gpu": B, "nvlink_bytes_per_gpu": Bn,
               "halo_GBps_per_gpu": round(B / us / 1e3, 1) if us else None, "roofline_us": round(t_roof, 2),
               "roofline_frac": round(t_roof / us, 3) if us else None,
               "launch": "cuda graph" if graph is not None else "stream-ordered calls", "ring": nring}
        results.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        del calls, graph
    igg.set_exchange_mode("ipc")
    del ring
    igg.finalize_global_grid()
    torch.cuda.empty_cache()
    barrier()


def time_gather():
    if args.only and "cfg5a" not in args.only:
        return
    if FORCED:
        return
    n = 256
    igg.init_global_grid(n, n, n, quiet=True)
    g = igg.global_grid()
    A = igg.zeros((n, n, n))
    A.copy_(torch.rand_like(A))
    G = igg.zeros(tuple(n * d for d in g.dims)) if rank == 0 else None
    for _ in range(2):
        igg.gather(A, G)
    barrier()
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        igg.gather(A, G)
    torch.cuda.synchronize()
    sec = maxr((time.perf_counter() - t0) / reps)
    ingress = (world - 1) * n ** 3 * 8
    rec = {"config": "cfg5a gather! 256^3 f64 per GPU -> root (device A_global)", "n_gpus": world, "dims": list(g.dims),
           "gather_ms": round(sec * 1e3, 3), "root_ingress_bytes": ingress,
           "roofline_ms": round(max(ingress / (NVL * 1e9), 2 * world * n ** 3 * 8 / (HBM * 1e9)) * 1e3, 3),
           "timing": "host wall clock of the blocking call (collective, includes the host hand-shake), max over ranks"}
    rec["roofline_frac"] = round(rec["roofline_ms"] / rec["gather_ms"], 3)
    results.append(rec)
    if rank == 0:
        print(json.dumps(rec), flush=True)
    igg.finalize_global_grid()
    barrier()


P = dict(periodx=1, periody=1, periodz=1)
f64, f32 = torch.float64, torch.float32
time_update("cfg2 256^3 f64 periodic", (256, 256, 256), P, [(256, 256, 256)], f64)
time_update("cfg3i 256^3 f64 non-periodic T only", (256, 256, 256), {}, [(256, 256, 256)], f64)
stag = [(256, 256, 256), (257, 256, 256), (256, 257, 256), (256, 256, 257)]
time_update("cfg3ii staggered T,Vx,Vy,Vz f64 non-periodic", (256, 256, 256), {}, stag, f64)
time_update("cfg3ii staggered T,Vx,Vy,Vz f64 periodic", (256, 256, 256), P, stag, f64)
for F in (1, 2, 4):
    time_update(f"cfg4 512^3 f32 hw2 periodic F={F}", (512, 512, 512), dict(P, overlaps=(4, 4, 4), halowidths=(2, 2, 2)),
                [(512, 512, 512)] * F, f32, ring_bytes=4.5e9)
time_gather()
for hw in (1, 2, 4, 8):
    for F in (1, 4):
        time_update(f"cfg5b 2-D 8192^2 f64 periodic hw={hw} F={F}", (8192, 8192, 1),
                    dict(periodx=1, periody=1, overlaps=(2 * hw, 2 * hw, 2), halowidths=(hw, hw, 1)),
                    [(8192, 8192)] * F, f64, ring_bytes=4.5e9, modes=("ipc",))
if rank == 0 and args.out:
    with open(args.out, "w") as f:
        for r in results:
            f.write(json.dumps(r) + "\n")
if world > 1:
    dist.destroy_process_group()
