"""BASELINE config 1: the reference's diffusion3D example (examples/diffusion3D_multicpu_novis.jl) at 64^3 Float64,
100 steps, non-periodic, one rank -- the fused CUDA step + update_halo! against the numpy restatement of the example.
Tolerance: none.  The kernel evaluates the example's operations in the example's order in IEEE double precision
without FMA contraction, so the temperature field must be bit-identical after 100 steps.  (Two ranks: mp_worker.py.)"""
import numpy as np
import pytest
import torch

import igg_b200 as igg
from igg_b200.diffusion3d import Diffusion3D
from oracle import diffusion_oracle as D

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu_and_clean_grid():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    yield
    if igg.grid_is_initialized():
        igg.finalize_global_grid()


@pytest.mark.parametrize("overlap", [False, True])
@pytest.mark.parametrize("n,kw", [((64, 64, 64), {}), ((40, 36, 33), dict(periodx=1, periodz=1))])
def test_diffusion3d_matches_example_bit_for_bit(n, kw, overlap):
    nt = 100
    grids, Ts, Cps, dxyz, dt = D.setup(1, n, **kw)
    T0, Cp0 = Ts[0].copy(order="F"), Cps[0].copy(order="F")
    want = D.run(grids, Ts, Cps, dxyz, dt, nt)[0]
    sim = Diffusion3D(*n, T=T0, Cp=Cp0, **kw)
    assert (sim.dx, sim.dy, sim.dz) == dxyz and sim.dt == dt
    got = igg.to_numpy(sim.run(nt, overlap=overlap))
    sim.finalize()
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), \
        f"max abs diff {np.abs(got - want).max()} after {nt} steps"
    assert np.abs(want - T0).max() > 1e-3          # the field did evolve


def test_fast_arithmetic_variant_within_tolerance():
    """The bandwidth-bound variant (reciprocal multiplications, FMA) is not bit-identical; tolerance 1e-12 relative
    to the field's range after 100 steps (observed ~1e-15)."""
    n, nt = (48, 40, 36), 100
    grids, Ts, Cps, dxyz, dt = D.setup(1, n)
    T0, Cp0 = Ts[0].copy(order="F"), Cps[0].copy(order="F")
    want = D.run(grids, Ts, Cps, dxyz, dt, nt)[0]
    sim = Diffusion3D(*n, T=T0, Cp=Cp0)
    sim.fast = True
    got = igg.to_numpy(sim.run(nt, overlap=True))
    sim.finalize()
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()


def test_own_initial_conditions_match_example():
    """The driver's own initial conditions (x_g/y_g/z_g of the package) equal the example's."""
    n = (32, 32, 32)
    grids, Ts, Cps, dxyz, dt = D.setup(1, n)
    sim = Diffusion3D(*n)
    assert np.array_equal(igg.to_numpy(sim.T), Ts[0]) and np.array_equal(igg.to_numpy(sim.Cp), Cps[0])
    sim.finalize()
