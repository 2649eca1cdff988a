#!/usr/bin/env python
"""bench.py -- update_halo! time and effective halo bandwidth on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (C restatement) on host cores

A "step" is one `update_halo!(T)` call on BASELINE config 2: one 256^3 Float64 field, overlaps (2,2,2),
halowidths (1,1,1), periodic in x/y/z, process grid from Dims_create(N) (1 -> 1x1x1 all self-exchange,
2 -> 2x1x1, 4 -> 2x2x1, 8 -> 2x2x2).  Weak scaling: every GPU always owns one 256^3 block.

Prints ONE JSON line.  `value` = halo bytes sent by all ranks per second (GB/s), device-timed with CUDA events on the
launching stream, max over ranks; `e2e` = the same through the blocking public API on fields that live in pinned HOST
memory (the kernel reads the send boxes and writes the halos over PCIe inside the timed region).
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
        "config": {"workload": "cfg2: update_halo!(T), T 256^3 Float64 per rank, overlaps (2,2,2), halowidths (1,1,1), "
                               "periodic xyz, dims " + "x".join(map(str, dims_for(n))),
                   "halo_bytes_per_rank_per_step": halo_bytes_per_gpu(), "ranks": n},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(ncores, n * threads), "kind": "port",
                         "sample": f"{steps} update_halo! calls after {warmup} warm-up on {n} rank thread(s) x {threads} "
                                   "copy thread(s); C restatement of the reference CPU path (oracle/cpu_ref.c), the Julia "
                                   "reference cannot run on this box"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


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

    n = N_LOCAL
    igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
    g = igg.global_grid()
    B_send = halo_bytes_per_gpu()
    B_nvl = nvlink_bytes_per_gpu(world)

    # Ring of fields so that consecutive steps touch different memory: 32-byte-sector footprint of one update is
    # ~12 MiB per field (x-faces touch a full sector per 8-byte cell), ring footprint must exceed the 126 MB L2.
    ring = args.ring
    fields = []
    torch.manual_seed(1234 + rank)
    for i in range(ring):
        t = igg.zeros((n, n, n), dtype=torch.float64)
        t.copy_(torch.rand((n, n, n), dtype=torch.float64, device="cuda").permute(2, 1, 0))
        fields.append(t)
    stream = torch.cuda.current_stream()
    ctxp = C.c_void_p(g.ctx)
    lib = L.load()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()

    def parity_check(f):
        """Size-independent parity property (checked once before and once after the timed loop): after an
        update every low halo plane equals the high send plane of the neighbour below (ol = 2, hw = 1), in every
        dimension, whoever that neighbour is.  Planes travel through torch.distributed (gloo), not through the
        library under test."""
        bad = 0
        for d in range(3):
            left, right = g.neighbors[0][d], g.neighbors[1][d]
            send_hi = f.narrow(d, n - 2, 1).contiguous().cpu()
            if right == g.me:
                got_hi = send_hi
            else:
                got_hi = torch.empty_like(send_hi)
                reqs = [dist.isend(send_hi, right, tag=20 + d), dist.irecv(got_hi, left, tag=20 + d)]
                for r_ in reqs:
                    r_.wait()
            if not torch.equal(f.narrow(d, 0, 1).contiguous().cpu().view(torch.int64), got_hi.view(torch.int64)):
                bad += 1
        return int(max_over_ranks(float(bad)))

    # One prepared (argument-checked once) stream-ordered call per field of the ring: a time loop's steady state.
    stream = torch.cuda.Stream()
    calls = [igg.prepare_update_halo(f, stream=stream, blocking=False) for f in fields]
    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            calls[i % ring]()
    barrier()
    parity_bad = parity_check(fields[0])
    barrier()
    # Host cost of issuing one stream-ordered call (reported, not part of `value` when the graph path is used).
    t_host0 = time.perf_counter()
    for i in range(200):
        calls[i % ring]()
    host_issue_us = (time.perf_counter() - t_host0) / 200 * 1e6
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    graph = None
    if not args.no_graph:
        # The K timed steps are captured once into a CUDA graph (the exchange state lives on the device, so the
        # launch parameters are call-invariant) and replayed: the time loop of a graph-captured application.
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            for i in range(args.steps):
                calls[i % ring]()
        graph.replay()                          # untimed: instantiation / upload
        barrier()
    l0 = lib.igg_launch_count(ctxp)
    with torch.cuda.stream(stream):
        e0.record(stream)
        if graph is not None:
            graph.replay()
        else:
            for i in range(args.steps):
                calls[i % ring]()
        e1.record(stream)
    barrier()
    launches = args.steps if graph is not None else lib.igg_launch_count(ctxp) - l0
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    sec_per_step = dev_ms * 1e-3 / args.steps

    # keep the same launches going for a moment so that the clock sampler sees the load (untimed)
    # (the same number of calls on every rank: the call is collective)
    extra = max(1, min(200, int(1.0 / max(sec_per_step * args.steps, 1e-6))))
    for _ in range(extra):
        if graph is not None:
            graph.replay()
        else:
            for i in range(args.steps):
                calls[i % ring]()
    torch.cuda.synchronize()
    barrier()

    # ---- per-call distribution: one CUDA-event pair around every stream-ordered call, and the
    # host-synchronous call (the reference's semantics: returns when the halos are written) on the host clock ----
    n_dist = 200
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_dist)]
    with torch.cuda.stream(stream):
        for i, (a, b) in enumerate(evs):
            a.record(stream)
            calls[i % ring]()
            b.record(stream)
    barrier()
    per_call = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    blocking_calls = [igg.prepare_update_halo(f, stream=stream, blocking=True) for f in fields]
    host_us = []
    for i in range(n_dist):
        t_h = time.perf_counter()
        blocking_calls[i % ring]()
        host_us.append((time.perf_counter() - t_h) * 1e6)
    barrier()
    host_us.sort()
    pct = lambda v, q: v[min(len(v) - 1, int(q * len(v)))]
    per_call_us = {k: max_over_ranks(pct(per_call, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}
    blocking_us = {k: max_over_ranks(pct(host_us, q)) for k, q in (("p10", 0.1), ("p50", 0.5), ("p90", 0.9))}
    parity_bad += parity_check(fields[1 % ring])
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
    if parity_bad:
        raise SystemExit(f"bench.py: parity check failed ({parity_bad} halo planes differ); no number reported")

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

    igg.finalize_global_grid()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0

    peaks, peak_src = measured_peaks()
    hbm = float(peaks.get("hbm_gbs", HBM_FALLBACK_GBS))
    t_hbm = 2.0 * B_send / (hbm * 1e9)
    t_nvl = B_nvl / (NVLINK_MEASURED_GBS * 1e9)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(f"n{world}")
        except Exception:
            traffic = None
    if t_nvl > t_hbm:
        roofline = {"bound": "nvlink", "achieved": B_nvl / sec_per_step / 1e9, "peak": NVLINK_MEASURED_GBS, "unit": "GB/s",
                    "frac": t_nvl / sec_per_step, "traffic": traffic,
                    "note": "bytes sent to other GPUs per launch / launch time vs measured peer-copy bandwidth per direction"}
    else:
        roofline = {"bound": "hbm", "achieved": 2.0 * B_send / sec_per_step / 1e9, "peak": hbm, "unit": "GB/s",
                    "frac": t_hbm / sec_per_step, "traffic": traffic,
                    "note": f"algorithmic bytes = 2*sizeof(T) per halo cell = {2 * B_send} B per launch; peak {peak_src}"}
    line = {
        "metric": METRIC, "value": world * B_send / sec_per_step / 1e9, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec_per_step * 1e3,
        "update_halo_us": sec_per_step * 1e6, "halo_GBps_per_gpu": B_send / sec_per_step / 1e9,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "cfg2: update_halo!(T), T 256^3 Float64 per GPU, overlaps (2,2,2), halowidths (1,1,1), "
                               "periodic xyz, dims " + "x".join(map(str, g.dims)),
                   "halo_bytes_per_gpu_per_step": B_send, "nvlink_bytes_per_gpu_per_step": B_nvl,
                   "l2": f"inputs larger than L2: ring of {ring} distinct 128 MiB fields, one per step "
                         f"(~{12 * ring} MiB of touched sectors between reuses vs 126 MB L2)",
                   "timing": "CUDA events on the launching stream around K stream-ordered calls, max over ranks",
                   "e2e_path": "blocking igg_update_halo on fields in pinned host memory; the kernels read the send "
                               "boxes and write the halos over PCIe (zero-copy; x self-wraps through whole 64-byte "
                               "lines), no device-resident copy of the field"},
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": world * B_send / e2e_sec / 1e9, "unit": UNIT, "us_per_step": e2e_sec * 1e6,
                "h2d_bytes_per_step": B_send, "d2h_bytes_per_step": B_send, "steps": e2e_steps},
        "per_call_us": dict(per_call_us, n=n_dist, how="one CUDA-event pair per stream-ordered call (not graph-replayed), max over ranks"),
        "blocking_call_us": dict(blocking_us, n=n_dist, how="host clock around the host-synchronous call (reference semantics), device-resident fields, max over ranks"),
        "parity": "ok" if parity_bad == 0 else f"{parity_bad} halo planes differ from the neighbour's send plane",
        "gpu_launches": int(launches),
        "launch_mode": "one CUDA graph of K update_halo launches, replayed" if graph is not None else "K stream-ordered calls",
        "host_issue_us_per_call": host_issue_us,
        "clocks": clocks,
        "kernel": lib.igg_kernel_name().decode(),
        "protocol": "fence (release fence + flag per 8-KiB tile)" if os.environ.get("IGG_MODE") in ("fence", "ipc")
                    else "ll128 (120 payload bytes + 8-byte epoch per 128-byte staging line; default)",
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
    ap.add_argument("--no-graph", action="store_true", help="issue the timed steps one by one instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
