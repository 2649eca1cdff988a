"""Weak scaling of the diffusion3D step (BASELINE config 3 (i)): 256^3 Float64 per GPU, non-periodic, dims from
Dims_create(N), as the reference example (examples/diffusion3D_multigpu_CuArrays_novis.jl).

    python profiles/bench_diffusion.py                                                    # 1 GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/bench_diffusion.py

Prints, for the serial schedule (stencil, then update_halo!) and the overlapped one (shell, exchange || interior), the
device time per step (CUDA events over `nt` steps after warm-up, max over ranks).  Weak-scaling efficiency =
t_step(1 GPU) / t_step(N GPUs); the 1-GPU run does no exchange at all.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import torch
import torch.distributed as dist
import igg_b200 as igg
from igg_b200.diffusion3d import Diffusion3D

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 300


def maxr(x):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


sim = Diffusion3D(n, n, n)
out = {"config": f"diffusion3D {n}^3 f64 per GPU, non-periodic", "n_gpus": world, "dims": list(igg.global_grid().dims), "nt": nt}
for fast, overlap, graph in ((False, False, False), (False, True, False), (True, False, False), (True, True, False),
                             (False, False, True), (True, False, True), (True, True, True)):
    sim.fast = fast
    sim._graphs.clear()                     # the captured kernels carry the arithmetic variant
    sim.run(20, overlap=overlap, graph=graph)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sim.main)
    sim.run(nt, overlap=overlap, graph=graph)
    e1.record(sim.main)
    torch.cuda.synchronize()
    us = maxr(e0.elapsed_time(e1)) * 1e3 / nt
    out[("fast_" if fast else "exact_") + ("overlap" if overlap else "serial") + ("_graph_us" if graph else "_us")] = round(us, 2)
# stencil alone (no exchange) for reference
for fast in (False, True):
    sim.fast = fast
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sim.main)
    for _ in range(nt):
        sim._stencil(sim.T, sim.T2, 0, sim.main)
    e1.record(sim.main)
    torch.cuda.synchronize()
    us = maxr(e0.elapsed_time(e1)) * 1e3 / nt
    out[("fast_" if fast else "exact_") + "stencil_only_us"] = round(us, 2)
    out[("fast_" if fast else "exact_") + "stencil_GBps"] = round(3 * n ** 3 * 8 / us / 1e3, 1)
sim.finalize()
if rank == 0:
    print(json.dumps(out), flush=True)
if world > 1:
    dist.destroy_process_group()
