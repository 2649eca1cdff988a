"""The reference's 3-D heat diffusion example as a driver of the halo path
(ref: examples/diffusion3D_multigpu_CuArrays_novis.jl).

    import igg_b200 as igg
    from igg_b200 import diffusion3d
    sim = diffusion3d.Diffusion3D(256, 256, 256)         # calls init_global_grid like the example
    sim.run(nt=100, overlap=True)
    sim.finalize()

One step = the fused stencil kernel igg_diffusion3d_step (the example's five broadcasts, same arithmetic) followed by
update_halo!(T).  With `overlap=True` the two outermost cell layers are computed first on a high-priority stream, the
halo exchange is issued right behind them and runs while the interior is computed on the main stream -- the
communication/computation overlap the reference advertises (ref: README.md:10).
"""
from __future__ import annotations

import ctypes as C

from . import _lib as L
from . import api


class Diffusion3D:
    def __init__(self, nx, ny, nz, lam=1.0, cp_min=1.0, lxyz=(10.0, 10.0, 10.0), init_grid=True, T=None, Cp=None,
                 **grid_kw):
        import numpy as np
        import torch
        self.torch = torch
        self.lam, self.cp_min = float(lam), float(cp_min)
        if init_grid:
            api.init_global_grid(nx, ny, nz, quiet=grid_kw.pop("quiet", True), **grid_kw)    # :22
        self.owns_grid = init_grid
        g = api.global_grid()
        self.n = (nx, ny, nz)
        lx, ly, lz = lxyz
        self.dx, self.dy, self.dz = lx / (api.nx_g() - 1), ly / (api.ny_g() - 1), lz / (api.nz_g() - 1)   # :23-25
        if T is None or Cp is None:
            shp = (nx, ny, nz)
            x = np.array([api.x_g(i, self.dx, shp) for i in range(1, nx + 1)]).reshape(-1, 1, 1)
            y = np.array([api.y_g(i, self.dy, shp) for i in range(1, ny + 1)]).reshape(1, -1, 1)
            z = np.array([api.z_g(i, self.dz, shp) for i in range(1, nz + 1)]).reshape(1, 1, -1)
            Cp = cp_min + (5 * np.exp(-(x - lx / 1.5) ** 2 - (y - ly / 2) ** 2 - (z - lz / 1.5) ** 2) +
                           5 * np.exp(-(x - lx / 3.0) ** 2 - (y - ly / 2) ** 2 - (z - lz / 1.5) ** 2))   # :35-36
            T = (100 * np.exp(-((x - lx / 2) / 2) ** 2 - ((y - ly / 2) / 2) ** 2 - ((z - lz / 3.0) / 2) ** 2) +
                 50 * np.exp(-((x - lx / 2) / 2) ** 2 - ((y - ly / 2) / 2) ** 2 - ((z - lz / 1.5) / 2) ** 2))  # :37-38
        self.T = api.from_numpy(np.asarray(T, dtype=np.float64))
        self.Cp = api.from_numpy(np.asarray(Cp, dtype=np.float64))
        self.T2 = api.zeros(self.n, dtype=torch.float64)
        self.dt = min(self.dx ** 2, self.dy ** 2, self.dz ** 2) * self.cp_min / self.lam / 8.1   # :41
        self._n3 = (C.c_int64 * 3)(nx, ny, nz)
        self._lib = L.load()
        lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
        self.main = torch.cuda.Stream()
        self.comm = torch.cuda.Stream(priority=-1)
        self._ev_shell, self._ev_halo, self._ev_prev = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self._calls = {}
        self._graphs = {}

    fast = False    # True: bandwidth-bound arithmetic (reciprocals + FMA), agrees with the exact variant to ~1e-15 relative

    def _stencil(self, src, dst, part, stream):
        rc = self._lib.igg_diffusion3d_step(src.data_ptr(), self.Cp.data_ptr(), dst.data_ptr(), self._n3, self.lam,
                                            self.dt, self.dx, self.dy, self.dz, part | (4 if self.fast else 0),
                                            stream.cuda_stream)
        if rc:
            L.check(rc)

    def _halo(self, t, stream):
        key = (t.data_ptr(), stream.cuda_stream)
        call = self._calls.get(key)
        if call is None:
            stream.synchronize()
            call = self._calls[key] = api.prepare_update_halo(t, stream=stream, blocking=False)
            return          # prepare_update_halo already performed this update
        call()

    def step(self, overlap=True):
        """T <- T + dt*dTedt on the inner cells, then update_halo!(T)   (:43-48)."""
        src, dst = self.T, self.T2
        if not overlap:
            self._stencil(src, dst, 0, self.main)
            self._halo(dst, self.main)
        else:
            self._ev_prev.record(self.main)                 # everything of the previous step is done on `main`
            self.comm.wait_event(self._ev_prev)
            self._stencil(src, dst, 1, self.comm)           # shell first, high priority
            self._halo(dst, self.comm)                      # exchange right behind it ...
            self._ev_halo.record(self.comm)
            self._stencil(src, dst, 2, self.main)           # ... while the interior is computed
            self.main.wait_event(self._ev_halo)
        self.T, self.T2 = dst, src

    def run(self, nt, overlap=True, graph=False):
        """`nt` time steps.  graph=True: two steps (T -> T2 -> T) are captured once into a CUDA graph -- stencil launches
        and update_halo! launches alike, the exchange state lives on the device -- and replayed, which removes the
        launch gaps between the kernels of a step."""
        torch = api._torch()
        if graph and nt >= 4:
            g = self._graphs.get(overlap)
            if g is None:
                self.step(overlap)                       # warm: prepares the update_halo! calls of both arrays
                self.step(overlap)
                nt -= 2
                self.main.synchronize()
                self.comm.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.main):
                    self.step(overlap)
                    self.step(overlap)
                self._graphs[overlap] = g
            with torch.cuda.stream(self.main):
                for _ in range(nt // 2):
                    g.replay()
            nt = nt % 2
        for _ in range(nt):
            self.step(overlap)
        self.main.synchronize()
        self.comm.synchronize()
        return self.T

    def finalize(self):
        self.main.synchronize()
        self.comm.synchronize()
        self._graphs.clear()
        self._calls.clear()
        if self.owns_grid:
            api.finalize_global_grid()
