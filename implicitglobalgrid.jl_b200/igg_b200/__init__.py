"""igg_b200 -- Python host side of libigg_b200, mirroring ImplicitGlobalGrid.jl's API for the halo path.

    import igg_b200 as igg
    me, dims, nprocs, coords, comm = igg.init_global_grid(256, 256, 256, periodx=1, periody=1, periodz=1)
    T = igg.zeros((256, 256, 256))
    igg.update_halo(T)
    igg.finalize_global_grid()

Importing the package does not load the shared library; the first call that needs it does, and fails loudly if it is
missing (there is no Python/CPU fallback for the data path).
"""
from ._lib import IGGError, LIB_PATH
from .api import (ArgumentError, GlobalGrid, alloc_field, free_field, finalize_global_grid, from_numpy, gather, get_global_grid, global_grid,
                  grid_is_initialized, init_global_grid, ix_g, iy_g, iz_g, nx_g, ny_g, nz_g, prepare_update_halo, set_exchange_mode,
                  select_device, tic, to_numpy,
                  toc, update_halo, x_g, y_g, z_g, zeros)
from .cellarray import CellArray

# Julia spelling aliases
update_halo_ = update_halo
gather_ = gather

__all__ = ["IGGError", "ArgumentError", "GlobalGrid", "LIB_PATH", "init_global_grid", "finalize_global_grid",
           "update_halo", "update_halo_", "prepare_update_halo", "set_exchange_mode", "gather", "gather_", "select_device", "get_global_grid", "global_grid",
           "grid_is_initialized", "nx_g", "ny_g", "nz_g", "x_g", "y_g", "z_g", "ix_g", "iy_g", "iz_g", "tic", "toc",
           "zeros", "alloc_field", "free_field", "from_numpy", "to_numpy", "CellArray"]
