"""Device time per update_halo by dimension subset and L2 temperature (1 GPU, cfg2), replayed from a CUDA graph so
that host launch cost does not show: which faces cost what."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "implicitglobalgrid.jl_b200"))
import torch
import igg_b200 as igg

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dt = {"f64": torch.float64, "f32": torch.float32}[sys.argv[2] if len(sys.argv) > 2 else "f64"]
igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
nring = 24
ring = [igg.zeros((n, n, n), dtype=dt) for _ in range(nring)]
for t in ring:
    t.copy_(torch.rand_like(t))
s = torch.cuda.Stream()


def timed(dims, fields, reps=96):
    with torch.cuda.stream(s):
        for i in range(8):
            igg.update_halo(fields[i % len(fields)], dims=dims, stream=s, blocking=False)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(reps):
                igg.update_halo(fields[i % len(fields)], dims=dims, stream=s, blocking=False)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(10):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (10 * reps)


for dims in [(1,), (2,), (3,), (2, 3), (1, 2, 3)]:
    cold = timed(dims, ring)
    hot = timed(dims, ring[:1])
    print(f"dims={dims}: cold-L2 {cold:.2f} us/update, hot-L2 {hot:.2f} us/update", flush=True)
igg.finalize_global_grid()
