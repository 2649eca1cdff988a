"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): one process per GPU under torchrun, CUDA-IPC
peer stores between the ranks, results gathered on rank 0 and compared bitwise with the oracle simulating the same
ranks.  See tests/mp_worker.py for the cases (periodic / non-periodic, staggered, wide halos, per-field halowidths,
1-D/2-D grids, dims subsets, staging growth, gather! with default and non-default root).  Run once per IPC
protocol: release fence + per-tile flags (default) and LL128 (flags inside the data lines)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("proto", ["fence", "ll128"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_multirank_vs_oracle(nranks, proto):
    if not torch.cuda.is_available() or torch.cuda.device_count() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", IGG_DEVICE_TIMEOUT_S="20", IGG_HOST_TIMEOUT_S="60", IGG_MODE=proto)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + nranks + (20 if proto == "ll128" else 0)), os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-4000:] + "\n" + r.stderr[-4000:])
    assert f"MP_WORKER_OK {nranks}" in r.stdout
