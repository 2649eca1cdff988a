"""Measures host-side cost per update_halo call vs device time per kernel (1 GPU, cfg2)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import torch
import igg_b200 as igg

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
ring = [igg.zeros((n, n, n)) for _ in range(24)]
for t in ring:
    t.copy_(torch.rand_like(t))
torch.cuda.synchronize()
for i in range(200):
    igg.update_halo(ring[i % 24], blocking=False)
torch.cuda.synchronize()
K = 5000
t0 = time.perf_counter()
for i in range(K):
    igg.update_halo(ring[i % 24], blocking=False)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host issue {1e6*(t1-t0)/K:.2f} us/call, total incl. drain {1e6*(t2-t0)/K:.2f} us/call")
# per-kernel device time: events around single launches with an idle gap before
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
for i, (a, b) in enumerate(ev):
    torch.cuda.synchronize()
    a.record(); igg.update_halo(ring[i % 24], blocking=False); b.record()
torch.cuda.synchronize()
ts = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
print(f"single-launch event time us: min {ts[0]:.2f} p50 {ts[len(ts)//2]:.2f} p90 {ts[int(len(ts)*0.9)]:.2f}")
# CUDA graph of 100 calls (single rank: no epochs involved)
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for i in range(10):
        igg.update_halo(ring[i % 24], stream=s, blocking=False)
    s.synchronize()
    try:
        with torch.cuda.graph(g, stream=s):
            for i in range(96):
                igg.update_halo(ring[i % 24], stream=s, blocking=False)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"graph replay: {e0.elapsed_time(e1)*1e3/(20*96):.2f} us/update")
    except Exception as e:
        print("graph capture failed:", repr(e)[:300])
igg.finalize_global_grid()
