#!/usr/bin/env python
"""bench.py -- update_halo! time and effective halo bandwidth on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (C restatement) on host cores

A "step" is one `update_halo!(T)` call on BASELINE config 2: one 256^3 Float64 field, overlaps (2,2,2),
halowidths (1,1,1), periodic in x/y/z, process grid from Dims_create(N) (1 -> 1x1x1 all self-exchange,
2 -> 2x1x1, 4 -> 2x2x1, 8 -> 2x2x2).  Weak scaling: every GPU always owns one 256^3 block.

Prints ONE JSON line.  `value` = halo bytes sent by all ranks per second (GB/s) over K stream-ordered calls timed
with one pair of CUDA events on the launching stream, max over ranks (one kernel launch per call, no CUDA graph).  The
per-call median with an event pair around every call and the same K launches replayed from a CUDA graph
are reported as extra keys.  `e2e` = the same metric
through the blocking public API on fields that live in pinned HOST memory (the kernel reads the send boxes and writes
the halos over PCIe inside the timed region).  Before anything is timed every rank checks the committed golden vectors
of its world size and a gather! (parity object of the line); a mismatch ends the run without a number.
Halo bytes per call follow SURVEY.md section 8(d): sum over dims with a halo of (#sides with a neighbour) * hw *
prod(other sizes) * sizeof(T) = 6 * 256*256*8 = 3,145,728 B per GPU.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "implicitglobalgrid.jl_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

N_LOCAL = 256
ELSIZE = 8
METRIC = "update_halo_effective_halo_bandwidth"
UNIT = "GB/s"
NVLINK_MEASURED_GBS = 770.0      # /opt/skills/guides/B200_PROFILING.md: measured peer copy per direction per GPU
HBM_FALLBACK_GBS = 6650.0        # same guide, fallback when MEASURED_PEAKS.json is absent


def halo_bytes_per_gpu(n=N_LOCAL, el=ELSIZE, hw=1):
    return 6 * hw * n * n * el


def dims_for(nprocs):
    from igg_b200 import topology as T
    return T.dims_create(nprocs, (0, 0, 0))


def nvlink_bytes_per_gpu(nprocs, n=N_LOCAL, el=ELSIZE):
    """Bytes a rank sends to OTHER ranks (faces in dimensions with more than one rank)."""
    d = dims_for(nprocs)
    return sum(2 * n * n * el for k in range(3) if d[k] > 1)


def workload_config(nranks, dims):
    """The `config` object of BOTH arms (b200 and reference): the same workload, worded once."""
    return {"workload": "cfg2: update_halo!(T), T 256^3 Float64 per rank (one rank per GPU), overlaps (2,2,2), "
                        "halowidths (1,1,1), periodic xyz, dims " + "x".join(map(str, dims)),
            "halo_bytes_per_rank_per_step": halo_bytes_per_gpu(),
            "nvlink_bytes_per_rank_per_step": nvlink_bytes_per_gpu(nranks),
            "ranks": nranks,
            "l2": "inputs larger than L2: the GPU arm steps through a ring of 24 distinct 128 MiB fields (~12 MiB of "
                  "touched sectors per step vs 126 MB L2); the CPU arm updates one 128 MiB field per rank",
            "timing": "GPU arm: CUDA events on the launching stream around K stream-ordered calls (barrier + synchronize "
                      "on both sides), max over ranks; CPU arm: host clock around K calls"}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return json.load(open(path)), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return {"hbm_gbs": HBM_FALLBACK_GBS}, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the benchmark loops run."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.err = index, [], set(), None, None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            while not self._stop_evt.is_set():
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                try:
                    r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.02)
        except Exception as e:  # NVML missing: report, do not fake
            self.err = repr(e)

    def stop(self):
        self._stop_evt.set()
        self.join(2.0)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), **({"error": self.err} if self.err else {})}


def cpu_reference_run(nranks, steps, warmup, threads):
    """The reference's CPU path (oracle/cpu_ref.c restatement) on `nranks` simulated ranks; returns (seconds/step, cores)."""
    import numpy as np
    from oracle import cpu_ref
    from oracle import igg_oracle as O
    dims = dims_for(nranks)
    n = N_LOCAL
    rng = np.random.default_rng(1234)
    fields = []
    for r in range(nranks):
        a = np.asfortranarray(rng.random((n, n, n)))
        fields.append([a])
    args = (dims, (1, 1, 1), (n, n, n), (2, 2, 2), fields, [(1, 1, 1)])
    if warmup:
        cpu_ref.update_halo(*args, copy_threads=threads, iters=warmup)
    sec = cpu_ref.update_halo(*args, copy_threads=threads, iters=steps)
    return sec / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    ncores = len(os.sched_getaffinity(0))
    n = args.gpus
    threads = max(1, ncores // n)               # JULIA_NUM_THREADS per rank: all host cores shared by the ranks
    os.environ.setdefault("OMP_NUM_THREADS", str(threads))
    steps = max(1, min(args.steps, 200))        # bounded sample: one step moves 3 MiB per rank on the CPU
    warmup = max(1, min(args.warmup, 20))
    sec = cpu_reference_run(n, steps, warmup, threads)
    value = n * halo_bytes_per_gpu() / sec / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n, "steps": steps, "warmup": warmup,
        "ms_per_step": sec * 1e3, "update_halo_us": sec * 1e6, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(n, dims_for(n)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(ncores, n * threads), "kind": "port",
                         "sample": f"{steps} update_halo! calls after {warmup} warm-up on {n} rank thread(s) x {threads} "
                                   "copy thread(s); C restatement of the reference CPU path (oracle/cpu_ref.c), the Julia "
                                   "reference cannot run on this box"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def golden_parity(igg, torch, dist, world, rank):
    """Parity against the committed golden vectors of this world size (tests/golden/, closed-form known answers built
    without the oracle), through the public API, BEFORE anything is timed.  Every case runs twice with different data:
    first on the bit-inverted input (a pure copy maps ~in to ~out), then on the input itself -- a read of stale staging
    data from the first call would show in the second.  Then one gather! against a closed-form global array.
    Returns (cases, failed, gather_status)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import golden_cases as G
    failed, ncases = 0, 0
    for gi, case in G.cases(world):
        ncases += 1
        igg.init_global_grid(*case["grid"], quiet=True, **G.kwargs(case))
        hws = G.halowidths(case, igg.global_grid().halowidths)
        ins, outs = G.arrays(gi, rank, "in"), G.arrays(gi, rank, "out")
        bad = False
        for rep, inv in enumerate((True, False)):
            # bitwise NOT of every element (F-ordered arrays: go through the flat byte view of the transpose)
            flip = (lambda a: np.invert(np.ascontiguousarray(a.T).view(np.uint8)).view(a.dtype).T) if inv else (lambda a: a)
            fields = [igg.from_numpy(np.asfortranarray(flip(a))) for a in ins]
            igg.update_halo(*[(t, hw) for t, hw in zip(fields, hws)], dims=tuple(case["dims"]), blocking=(rep == 1))
            torch.cuda.synchronize()
            for t, want in zip(fields, outs):
                got = np.ascontiguousarray(igg.to_numpy(t).T).view(np.uint8)
                if not np.array_equal(got, np.ascontiguousarray(np.asfortranarray(flip(want)).T).view(np.uint8)):
                    bad = True
        igg.finalize_global_grid()
        if bad:
            failed += 1
            sys.stderr.write(f"bench.py: GOLDEN MISMATCH {case['name']} on rank {rank}\n")
        if world > 1:
            dist.barrier()
    # gather!: block (cx,cy,cz) of a global array that holds every cell's column-major linear index
    nl = (12, 10, 8)
    igg.init_global_grid(*nl, quiet=True)
    g = igg.global_grid()
    NX, NY, NZ = (nl[d] * g.dims[d] for d in range(3))
    ix = np.arange(nl[0])[:, None, None] + g.coords[0] * nl[0]
    iy = np.arange(nl[1])[None, :, None] + g.coords[1] * nl[1]
    iz = np.arange(nl[2])[None, None, :] + g.coords[2] * nl[2]
    block = np.asfortranarray((ix + NX * (iy + NY * iz)).astype(np.int64))
    A = igg.from_numpy(block)
    gather_ok = True
    for root in sorted({0, world - 1}):
        AG = igg.zeros((NX, NY, NZ), dtype=torch.int64) if rank == root else None
        igg.gather(A, AG, root=root)
        if rank == root:
            got = igg.to_numpy(AG).reshape(-1, order="F")
            gather_ok = gather_ok and bool(np.array_equal(got, np.arange(NX * NY * NZ, dtype=np.int64)))
    igg.finalize_global_grid()
    return ncases, failed, gather_ok


def peer_copy_gbs(igg, L, C, torch, world, rank):
    """Plain peer copy to the right neighbour rank (256 MiB, best of 5, CUDA events), all ranks at once, through the very
    CUDA-IPC mappings the exchange uses (a symmetric allocation): the practical NVLink bandwidth per direction on this
    box, the denominator of the NVLink roofline.  None if it cannot be measured."""
    try:
        if world < 2:
            return None
        nb = 256 << 20
        t = igg.alloc_field((nb,), dtype=torch.uint8)
        gbs = C.c_double(0)
        torch.cuda.synchronize()
        L.check(L.load().igg_debug_peer_copy(igg.global_grid().ctx, C.c_void_p(t.data_ptr()), (rank + 1) % world, nb, 5, C.byref(gbs)))
        torch.cuda.synchronize()
        igg.free_field(t)
        return gbs.value
    except Exception as e:
        sys.stderr.write(f"bench.py: peer-copy measurement failed: {e!r}\n")
        return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    import igg_b200 as igg
    from igg_b200 import _lib as L
    import ctypes as C

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.stderr.write(f"bench.py: --gpus {args.gpus} needs torchrun with {args.gpus} ranks; running 1 rank.\n")
        args.gpus = world
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- parity first: golden vectors of this world size + gather!, untimed -------------------------------------
    golden_n, golden_failed, gather_ok = golden_parity(igg, torch, dist, world, rank)
    golden_failed = int(max_over_ranks(float(golden_failed)))
    gather_ok = max_over_ranks(0.0 if gather_ok else 1.0) == 0.0
    barrier()

    n = N_LOCAL
    igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
    g = igg.global_grid()
    B_send = halo_bytes_per_gpu()
    B_nvl = nvlink_bytes_per_gpu(world)

    # Ring of fields so that consecutive steps touch different memory: 32-byte-sector footprint of one update is
    # ~12 MiB per field (x-faces touch a full sector per 8-byte cell), ring footprint must exceed the 126 MB L2.
    ring = args.ring
    fields = []
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1234 + rank)
    for i in range(ring):
        t = igg.zeros((n, n, n), dtype=torch.float64)
        t.copy_(torch.rand((n, n, n), dtype=torch.float64, device="cuda", generator=gen).permute(2, 1, 0))
        fields.append(t)
    ctxp = C.c_void_p(g.ctx)
    lib = L.load()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()

    parity_calls = [0]

    def parity_check(f, call=None):
        """Size-independent parity property at the full 256^3 size (checked before and after the timed loop)
        on FRESH data every time (a stale read of an earlier call's staging would show): after an update the low halo
        plane of every dimension equals the high send plane of the neighbour below, and the high halo plane equals
        the low send plane of the neighbour above (ol = 2, hw = 1: send planes are index 1 and n-2), whoever those
        neighbours are.  Planes travel through torch.distributed (gloo), not through the library under test."""
        parity_calls[0] += 1
        torch.cuda.synchronize()
        f.copy_(torch.rand((n, n, n), dtype=torch.float64, device="cuda", generator=gen).permute(2, 1, 0)
                if f.is_cuda else torch.rand((n, n, n), dtype=torch.float64).permute(2, 1, 0))
        torch.cuda.synchronize()
        barrier()
        (call or (lambda: igg.update_halo(f)))()
        torch.cuda.synchronize()
        barrier()
        bad = 0
        for d in range(3):
            left, right = g.neighbors[0][d], g.neighbors[1][d]
            send_hi = f.narrow(d, n - 2, 1).contiguous().cpu()
            send_lo = f.narrow(d, 1, 1).contiguous().cpu()
            if right == g.me:
                from_left, from_right = send_hi, send_lo
            else:
                from_left, from_right = torch.empty_like(send_hi), torch.empty_like(send_lo)
                reqs = [dist.isend(send_hi, right, tag=20 + d), dist.irecv(from_left, left, tag=20 + d),
                        dist.isend(send_lo, left, tag=40 + d), dist.irecv(from_right, right, tag=40 + d)]
                for r_ in reqs:
                    r_.wait()
            if not torch.equal(f.narrow(d, 0, 1).contiguous().cpu().view(torch.int64), from_left.view(torch.int64)):
                bad += 1
            if not torch.equal(f.narrow(d, n - 1, 1).contiguous().cpu().view(torch.int64), from_right.view(torch.int64)):
                bad += 1
        return int(max_over_ranks(float(bad)))

    # One prepared (argument-checked once) stream-ordered call per field of the ring: a time loop's steady state.
    stream = torch.cuda.Stream()
    calls = [igg.prepare_update_halo(f, stream=stream, blocking=False) for f in fields]
    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            calls[i % ring]()
    barrier()
    parity_bad = parity_check(fields[0], calls[0])
    barrier()
    # Host cost of issuing one stream-ordered call: from Python (ctypes) and from C (what a compiled host pays).
    t_host0 = time.perf_counter()
    for i in range(200):
        calls[i % ring]()
    host_issue_us = (time.perf_counter() - t_host0) / 200 * 1e6
    barrier()
    ns = C.c_double(0)
    sp = C.c_void_p(stream.cuda_stream)
    L.check(lib.igg_debug_issue_loop(calls[0]._h, sp, 200, C.byref(ns)))
    host_issue_c_us = ns.value * 1e-3
    barrier()

    # ---- the timed region: K stream-ordered calls, one CUDA-event pair around each (the metric is the
    # per-call median), bracketed by barrier + synchronize.  The loop is issued from C (igg_debug_timed_calls), the way
    # a compiled host language issues it: with CPython in the loop the host needs ~15 us per iteration (two
    # torch Event.record + one ctypes call), more than the kernel takes, and the number measures the interpreter. ---
    K = args.steps
    harr = (C.c_void_p * ring)(*[c._h for c in calls])
    us = (C.c_float * K)()
    loop = C.c_double(0)
    # (1) THE TIMED REGION: K stream-ordered calls between ONE pair of CUDA events, barrier + synchronize on both sides
    l0 = lib.igg_launch_count(ctxp)
    barrier()
    L.check(lib.igg_debug_timed_calls(harr, ring, sp, K, L.IGG_ASYNC | 2, us, C.byref(loop)))
    barrier()
    launches = lib.igg_launch_count(ctxp) - l0
    loop_noev_us = max_over_ranks(loop.value)
    sec_per_step = loop_noev_us * 1e-6
    # (2) the same K calls with a CUDA-event pair around every call (SURVEY 8d's per-call median).  On this system a
    # timing event costs the stream ~2.5 us, so these numbers carry ~4-5 us of event time per call.
    barrier()
    L.check(lib.igg_debug_timed_calls(harr, ring, sp, K, L.IGG_ASYNC, us, C.byref(loop)))
    barrier()
    per_call = sorted(us)
    pct = lambda v, q: v[min(len(v) - 1, int(q * len(v)))]
    per_call_us = {k: max_over_ranks(pct(per_call, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}
    loop_us = max_over_ranks(loop.value)
    # the same with CPython in the loop (torch events + ctypes), reported beside it
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
    with torch.cuda.stream(stream):
        for i, (a, b) in enumerate(evs):
            a.record(stream)
            calls[i % ring]()
            b.record(stream)
    barrier()
    py_calls = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    per_call_python_us = {k: max_over_ranks(pct(py_calls, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}

    # ---- the same K launches captured once into a CUDA graph and replayed (extra key, not the headline) ----------
    graph_us = None
    if not args.no_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            for i in range(K):
                calls[i % ring]()
        graph.replay()                          # untimed: instantiation / upload
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            graph.replay()
            e1.record(stream)
        barrier()
        graph_us = max_over_ranks(e0.elapsed_time(e1)) * 1e3 / K
        # keep the same launches going for a moment so that the clock sampler sees the load (untimed; the same
        # number of replays on every rank: the call is collective)
        for _ in range(max(1, min(200, int(1.0 / max(graph_us * 1e-6 * K, 1e-6))))):
            graph.replay()
        torch.cuda.synchronize()
        barrier()

    # ---- the host-synchronous call (the reference's semantics: returns when the halos are written), host clock ----
    n_dist = 200
    blocking_calls = [igg.prepare_update_halo(f, stream=stream, blocking=True) for f in fields]
    barr = (C.c_void_p * ring)(*[c._h for c in blocking_calls])
    bus = (C.c_float * n_dist)()
    barrier()
    L.check(lib.igg_debug_timed_calls(barr, ring, sp, n_dist, L.IGG_SYNC, bus, None))
    barrier()
    host_us = sorted(bus)
    blocking_us = {k: max_over_ranks(pct(host_us, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}
    py_block = []
    for i in range(n_dist):
        t_h = time.perf_counter()
        blocking_calls[i % ring]()
        py_block.append((time.perf_counter() - t_h) * 1e6)
    barrier()
    py_block.sort()
    blocking_python_us = {k: max_over_ranks(pct(py_block, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}
    parity_bad += parity_check(fields[1 % ring], blocking_calls[1 % ring])
    barrier()

    # ---- e2e: blocking public API, fields in pinned host memory (zero-copy over PCIe) ----
    hring = 2
    hfields = []
    for i in range(hring):
        h = igg.zeros((n, n, n), dtype=torch.float64, pin_memory=True)
        h.copy_(fields[i].cpu())
        hfields.append(h)
    e2e_steps = max(10, min(args.steps, 200))
    for i in range(3):
        igg.update_halo(hfields[i % hring])
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        igg.update_halo(hfields[i % hring])          # blocking: returns when the halos are in host memory
    torch.cuda.synchronize()
    e2e_sec = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
    barrier()

    clocks = sampler.stop() if sampler else None

    # the host-resident fields of the e2e loop must satisfy the same property (whole-array checks are in tests/)
    for h in hfields:
        parity_bad += parity_check(h)
    nvl_measured = peer_copy_gbs(igg, L, C, torch, world, rank)
    if nvl_measured is not None:
        nvl_measured = -max_over_ranks(-nvl_measured)        # the slowest rank's figure
    barrier()
    if parity_bad or golden_failed or not gather_ok:
        raise SystemExit(f"bench.py: parity check failed ({parity_bad} halo planes differ, {golden_failed} golden cases "
                         f"failed, gather {'ok' if gather_ok else 'MISMATCH'}); no number reported")

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncores = len(os.sched_getaffinity(0))
        cb_steps, cb_warm = 300, 10
        sec = cpu_reference_run(1, cb_steps, cb_warm, ncores)
        cpu_baseline = {"value": B_send / sec / 1e9, "unit": UNIT, "cores": ncores, "kind": "port",
                        "us_per_step": sec * 1e6,
                        "sample": f"{cb_steps} update_halo! calls (same 256^3 f64 periodic workload, 1 rank, {ncores} copy "
                                  "threads) of the C restatement of the reference CPU path (oracle/cpu_ref.c); the Julia "
                                  "reference itself cannot run on this box"}

    calls.clear()
    blocking_calls.clear()
    igg.finalize_global_grid()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0

    peaks, peak_src = measured_peaks()
    hbm = float(peaks.get("hbm_gbs", HBM_FALLBACK_GBS))
    # (a figure far below the profiling guide's 770 GB/s means the measurement was disturbed: use the guide's then)
    nvl_ok = bool(nvl_measured) and nvl_measured > 0.5 * NVLINK_MEASURED_GBS
    nvl = nvl_measured if nvl_ok else NVLINK_MEASURED_GBS
    nvl_src = "peer copy measured in this run" if nvl_ok else "B200_PROFILING.md peer-copy figure"
    t_hbm = 2.0 * B_send / (hbm * 1e9)
    t_nvl = B_nvl / (nvl * 1e9)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(f"n{world}")
        except Exception:
            traffic = None
    if t_nvl > t_hbm:
        roofline = {"bound": "nvlink", "achieved": B_nvl / sec_per_step / 1e9, "peak": nvl, "unit": "GB/s",
                    "frac": t_nvl / sec_per_step, "traffic": traffic,
                    "note": f"bytes sent to other GPUs per launch / launch time vs peer bandwidth per direction ({nvl_src})"}
    else:
        roofline = {"bound": "hbm", "achieved": 2.0 * B_send / sec_per_step / 1e9, "peak": hbm, "unit": "GB/s",
                    "frac": t_hbm / sec_per_step, "traffic": traffic,
                    "note": f"algorithmic bytes = 2*sizeof(T) per halo cell = {2 * B_send} B per launch; peak {peak_src}"}
    if graph_us:
        roofline["frac_graph_replay"] = max(t_nvl, t_hbm) / (graph_us * 1e-6)
    line = {
        "metric": METRIC, "value": world * B_send / sec_per_step / 1e9, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec_per_step * 1e3,
        "update_halo_us": sec_per_step * 1e6, "halo_GBps_per_gpu": B_send / sec_per_step / 1e9,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(world, g.dims),
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": world * B_send / e2e_sec / 1e9, "unit": UNIT, "us_per_step": e2e_sec * 1e6,
                "h2d_bytes_per_step": B_send, "d2h_bytes_per_step": B_send, "steps": e2e_steps,
                "path": "blocking igg_update_halo on fields in pinned host memory; the kernels read the send boxes and "
                        "write the halos over PCIe (zero-copy; x-faces, wrapped or exchanged, through whole 64-byte lines), no "
                        "device-resident copy of the field"},
        "per_call_us": dict(per_call_us, n=K, how="SURVEY 8d's per-call figure: one CUDA-event pair around every stream-ordered call (not graph-replayed), loop issued from C, max over ranks; includes the two timing events (~2.5 us each on the stream)"),
        "per_call_python_us": dict(per_call_python_us, n=200, how="the same with CPython in the loop (torch events + ctypes call)"),
        "loop_with_per_call_events_us_per_call": loop_us,
        "graph_replay_us_per_call": graph_us,
        "blocking_call_us": dict(blocking_us, n=n_dist, how="host clock around the host-synchronous call (reference semantics), issued from C, device-resident fields, max over ranks"),
        "blocking_call_python_us": dict(blocking_python_us, n=n_dist, how="the same called from CPython through ctypes"),
        "parity": {"golden_cases": golden_n, "failed": golden_failed, "gather": "ok" if gather_ok else "MISMATCH",
                   "halo_planes": "ok" if parity_bad == 0 else f"{parity_bad} differ",
                   "how": "committed golden vectors of this world size, each run twice (inverted bits, then as is) through "
                          "igg.update_halo; gather! of a closed-form global array onto the first and the last rank; "
                          f"both halo planes of every dimension vs the neighbours' send planes on fresh 256^3 data, {parity_calls[0]} times"},
        "gpu_launches": int(launches),
        "launch_mode": "K stream-ordered prepared calls (igg_update_halo_prepared) issued from C through the C ABI "
                       "(igg_debug_timed_calls), one launch each, one CUDA-event pair around the K calls; not graph-replayed",
        "host_issue_us_per_call": host_issue_c_us,
        "host_issue_python_us_per_call": host_issue_us,
        "nvlink_peer_copy_GBps": nvl_measured,
        "clocks": clocks,
        "kernel": lib.igg_kernel_name().decode(),
        "protocol": os.environ.get("IGG_MODE", "auto") + " (auto: LL128 lines below IGG_BULK_THRESHOLD staged bytes per call, "
                    "fence protocol with 16-byte accesses above)",
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ring", type=int, default=24)
    ap.add_argument("--no-graph", action="store_true", help="skip the extra CUDA-graph replay measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
