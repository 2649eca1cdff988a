"""CellArray fields for update_halo!.

The reference accepts CellArrays.jl arrays and turns them into plain bits arrays before the halo update
(`extract` / `bitsarrays`, ref: src/shared.jl:246-250,287-289):

  * B = 0 (component-major storage): one array per cell component, each of the grid's shape and contiguous
    -> `celllength` ordinary fields in the same call;
  * B = 1 (cell-major storage): ONE array of the grid's shape whose elements are whole cells
    (`reshape(reinterpret(T, view(A.data,:)), size(A))`) -> one field with element size `celllength * sizeof(eltype)`;
  * any other B is rejected with the reference's message.

This module mirrors that on torch storage.  `CellArray.data` is one tensor in the memory order CellArrays.jl uses for
these two layouts; `bitsarrays()` returns what `update_halo` (api.py) hands to the C ABI, which takes any element
size and any number of fields per call.
"""
from __future__ import annotations

from .topology import ArgumentError


class StructArrayView:
    """An array of the grid's shape whose elements are `cell_bytes`-byte cells (B = 1).  Quacks like the few tensor
    methods the host wrapper uses (data_ptr, shape, stride, dim, element_size, dtype, device)."""

    def __init__(self, storage, grid_shape, celllength):
        self.storage = storage                      # keeps the memory alive
        self.shape = tuple(int(s) for s in grid_shape)
        self._cell_bytes = int(celllength) * storage.element_size()
        self.dtype = ("cell", storage.dtype, int(celllength))
        self.device = storage.device

    def data_ptr(self):
        return self.storage.data_ptr()

    def dim(self):
        return len(self.shape)

    def element_size(self):
        return self._cell_bytes

    def stride(self):
        out, acc = [], 1
        for s in self.shape:
            out.append(acc)
            acc *= s
        return tuple(out)


class CellArray:
    """CellArray{Cell, N, B} with Cell = a `celldims`-shaped static array of `dtype` (e.g. SMatrix{2,2,Float64}).

    `data` holds everything: Julia shape (dims..., celllength) for B = 0 and (celllength, dims...) for B = 1,
    column-major like every array of this package (so `data[..., c]` / `data[c, ...]` is component c)."""

    def __init__(self, celldims, dims, dtype=None, B=0, device=None, pin_memory=False):
        from . import api
        if B not in (0, 1):
            raise ArgumentError("only CellArrays with B=0 or B=1 are supported")          # ref: src/shared.jl:289
        self.celldims = tuple(int(c) for c in (celldims if isinstance(celldims, (tuple, list)) else (celldims,)))
        self.dims = tuple(int(d) for d in (dims if isinstance(dims, (tuple, list)) else (dims,)))
        self.B = B
        self.celllength = 1
        for c in self.celldims:
            self.celllength *= c
        shape = self.dims + (self.celllength,) if B == 0 else (self.celllength,) + self.dims
        self.data = api.zeros(shape, dtype=dtype, device=device, pin_memory=pin_memory)

    @property
    def shape(self):
        return self.dims

    def field(self, c):
        """Component c (0-based) of every cell as an array of the grid's shape: contiguous for B = 0, strided for
        B = 1 (a view; not something update_halo takes -- use bitsarrays)."""
        return self.data[..., c] if self.B == 0 else self.data[c]

    def fill_components(self, values):
        """Every component of cell (ix, iy, iz) := values[ix, iy, iz] (what the reference's tests do with
        `P .= eltype(P).(...)`, test/test_update_halo.jl:1006)."""
        for c in range(self.celllength):
            self.field(c).copy_(values)
        return self

    def bitsarrays(self):
        """ref: src/shared.jl:287-288."""
        if self.B == 0:
            return [self.data[..., c] for c in range(self.celllength)]
        return [StructArrayView(self.data, self.dims, self.celllength)]
