"""numpy restatement of the reference's diffusion3D example on N simulated ranks  --  TEST INFRASTRUCTURE ONLY.

Follows examples/diffusion3D_multicpu_novis.jl (CPU twin of examples/diffusion3D_multigpu_CuArrays_novis.jl):
physics and numerics lines :13-25, initial conditions :35-38, time loop :41-48.  Every arithmetic operation is a
separate numpy ufunc in the example's order (no fused multiply-add), which is also what Julia's broadcast does, so
the GPU kernel can be compared bit for bit.  Halo updates go through the oracle's update_halo.
"""
import numpy as np

from . import igg_oracle as O


def setup(nprocs, n, lam=1.0, cp_min=1.0, l=(10.0, 10.0, 10.0), **grid_kw):
    nx, ny, nz = n
    grids = O.make_grids(nprocs, nx, ny, nz, **grid_kw)
    g0 = grids[0]
    dx, dy, dz = (l[d] / (g0.nxyz_g[d] - 1) for d in range(3))                  # :23-25
    lx, ly, lz = l
    Ts, Cps = [], []
    for g in grids:
        x = np.array([O.x_g(g, 1, i, dx, nx) for i in range(1, nx + 1)]).reshape(-1, 1, 1)
        y = np.array([O.x_g(g, 2, i, dy, ny) for i in range(1, ny + 1)]).reshape(1, -1, 1)
        z = np.array([O.x_g(g, 3, i, dz, nz) for i in range(1, nz + 1)]).reshape(1, 1, -1)
        Cp = cp_min + (5 * np.exp(-(x - lx / 1.5) ** 2 - (y - ly / 2) ** 2 - (z - lz / 1.5) ** 2) +
                       5 * np.exp(-(x - lx / 3.0) ** 2 - (y - ly / 2) ** 2 - (z - lz / 1.5) ** 2))     # :35-36
        T = (100 * np.exp(-((x - lx / 2) / 2) ** 2 - ((y - ly / 2) / 2) ** 2 - ((z - lz / 3.0) / 2) ** 2) +
             50 * np.exp(-((x - lx / 2) / 2) ** 2 - ((y - ly / 2) / 2) ** 2 - ((z - lz / 1.5) / 2) ** 2))  # :37-38
        Ts.append(np.asfortranarray(T))
        Cps.append(np.asfortranarray(Cp))
    dt = min(dx * dx, dy * dy, dz * dz) * cp_min / lam / 8.1                     # :41
    return grids, Ts, Cps, (dx, dy, dz), dt


def step_rank(T, Cp, lam, dt, dx, dy, dz):
    """One rank's local update, in place (:43-47)."""
    qx = -lam * (T[1:, 1:-1, 1:-1] - T[:-1, 1:-1, 1:-1]) / dx
    qy = -lam * (T[1:-1, 1:, 1:-1] - T[1:-1, :-1, 1:-1]) / dy
    qz = -lam * (T[1:-1, 1:-1, 1:] - T[1:-1, 1:-1, :-1]) / dz
    dTedt = 1.0 / Cp[1:-1, 1:-1, 1:-1] * (-(qx[1:] - qx[:-1]) / dx - (qy[:, 1:] - qy[:, :-1]) / dy
                                          - (qz[:, :, 1:] - qz[:, :, :-1]) / dz)
    T[1:-1, 1:-1, 1:-1] = T[1:-1, 1:-1, 1:-1] + dt * dTedt


def run(grids, Ts, Cps, dxyz, dt, nt, lam=1.0):
    dx, dy, dz = dxyz
    for _ in range(nt):
        for T, Cp in zip(Ts, Cps):
            step_rank(T, Cp, lam, dt, dx, dy, dz)
        O.update_halo(grids, [[T] for T in Ts])                                  # :48
    return Ts
