"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): one process per GPU under torchrun, CUDA-IPC
peer stores between the ranks, results gathered on rank 0 and compared bitwise with the oracle simulating the same
ranks.  See tests/mp_worker.py for the cases (periodic / non-periodic, staggered, wide halos, per-field halowidths,
1-D/2-D grids, dims subsets, staging growth, gather! with default and non-default root).  Run once per IPC
mode: the fence protocol (release fence + per-tile flags), LL128 (flags inside the data lines) and the default "auto"
(per call by size; with a 16 KiB threshold here, so that both protocols alternate on one context)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("proto", ["auto", "fence", "ll128"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_multirank_vs_oracle(nranks, proto):
    if not torch.cuda.is_available() or torch.cuda.device_count() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", IGG_DEVICE_TIMEOUT_S="20", IGG_HOST_TIMEOUT_S="60", IGG_MODE=proto,
               IGG_DIRECT_THRESHOLD="0")        # direct-to-halo also for the small cases (default: from 32 MiB per call)
    if proto == "auto":
        env["IGG_BULK_THRESHOLD"] = "16384"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + nranks + {"fence": 0, "ll128": 20, "auto": 40}[proto]), os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-4000:] + "\n" + r.stderr[-4000:])
    assert f"MP_WORKER_OK {nranks}" in r.stdout


def test_ll128_selftest_failure_falls_back_to_the_fence_protocol():
    """IGG_LL128_SELFTEST_FAIL=1 makes the bootstrap self-test see torn lines: every rank must end up in the fence
    protocol (the worker checks the context's mode) and the golden vectors + full-size properties must still hold."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", IGG_DEVICE_TIMEOUT_S="20", IGG_HOST_TIMEOUT_S="60", IGG_MODE="auto",
               IGG_LL128_SELFTEST_FAIL="1", IGG_WORKER_QUICK="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29671", os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-4000:] + "\n" + r.stderr[-4000:])
    assert "MP_WORKER_OK 2" in r.stdout and "LL128 is not safe on this system" in r.stderr
