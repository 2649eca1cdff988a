"""CellArray fields: the host mirror of `extract` / `bitsarrays` (ref: src/shared.jl:246-250,
287-289) and, on the GPU, the reference's full-update tests run on cell arrays (test/test_update_halo.jl:17-20 define
Cell = SMatrix{2,2,T}; the loop at :975 runs every full-update test on CellArrays with B = 0 and B = 1)."""
import numpy as np
import pytest
import torch

import igg_b200 as igg
from igg_b200 import api
from igg_b200.cellarray import StructArrayView
from oracle import igg_oracle as O

nx, ny, nz = 7, 5, 6


def test_layouts_and_bitsarrays_on_the_host():
    A0 = igg.CellArray((2, 2), (nx, ny, nz), dtype=torch.float64, B=0, device="cpu")
    assert A0.celllength == 4 and tuple(A0.data.shape) == (nx, ny, nz, 4) and A0.data.stride() == (1, nx, nx * ny, nx * ny * nz)
    b0 = A0.bitsarrays()
    assert len(b0) == 4
    for c, f in enumerate(b0):                       # B = 0: one contiguous column-major array per component
        assert tuple(f.shape) == (nx, ny, nz) and f.stride() == (1, nx, nx * ny)
        assert f.data_ptr() == A0.data.data_ptr() + c * nx * ny * nz * 8
    A1 = igg.CellArray((2, 2), (nx, ny, nz), dtype=torch.float32, B=1, device="cpu")
    assert tuple(A1.data.shape) == (4, nx, ny, nz) and A1.data.stride() == (1, 4, 4 * nx, 4 * nx * ny)
    (b1,) = A1.bitsarrays()                          # B = 1: one array of whole cells
    assert isinstance(b1, StructArrayView) and b1.shape == (nx, ny, nz) and b1.element_size() == 16
    assert b1.data_ptr() == A1.data.data_ptr() and b1.stride() == (1, nx, nx * ny) and b1.dim() == 3
    assert A1.field(2).stride() == (4, 4 * nx, 4 * nx * ny)
    with pytest.raises(igg.ArgumentError, match="only CellArrays with B=0 or B=1 are supported"):
        igg.CellArray((2, 2), (nx, ny, nz), B=2, device="cpu")


def test_extract_expands_cell_arguments():
    A0 = igg.CellArray((3,), (nx, ny), dtype=torch.float64, B=0, device="cpu")
    A1 = igg.CellArray((3,), (nx, ny), dtype=torch.float64, B=1, device="cpu")
    plain = igg.zeros((nx, ny), device="cpu")
    out = api._extract_all((plain, A0, (A1, (2, 1, 1)), {"A": A0, "halowidths": (1, 1, 1)}))
    assert out[0] is plain and len(out) == 1 + 3 + 1 + 3
    assert isinstance(out[4], tuple) and isinstance(out[4][0], StructArrayView) and out[4][1] == (2, 1, 1)
    assert isinstance(out[5], dict) and tuple(out[5]["A"].shape) == (nx, ny)
    assert api._extract_all((plain,)) == (plain,)


CELL_CASES = [
    # (grid args, kwargs, grid shape of the cell array, zeroed widths, reference test line)
    ((nx, ny, 1), dict(periodx=1, periody=1), (nx, ny), (1, 1), 990),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), (nx, ny, nz), (1, 1, 1), 1004),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), (nx, ny, nz), (2, 1, 1), 1019),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), (nx, ny, nz + 1), (1, 1, 1), 1084),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), (nx + 1, ny, nz), (2, 1, 1), 1099),
]


@pytest.mark.gpu
@pytest.mark.parametrize("B", [0, 1])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("gargs,kw,shape,widths,line", CELL_CASES, ids=[f"L{c[4]}" for c in CELL_CASES])
def test_reference_full_update_on_cell_arrays(gargs, kw, shape, widths, line, dtype, B):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    igg.init_global_grid(*gargs, quiet=True, **kw)
    try:
        (g,) = O.make_grids(1, *gargs, **kw)
        npdt = np.float64 if dtype == torch.float64 else np.float32
        ref = np.asfortranarray(O.coord_field(g, shape).astype(npdt))
        zeroed = ref.copy(order="F")
        for d in range(len(shape)):
            idx = [slice(None)] * len(shape)
            idx[d] = list(range(widths[d])) + list(range(shape[d] - widths[d], shape[d]))
            zeroed[tuple(idx)] = 0
        P = igg.CellArray((2, 2), shape, dtype=dtype, B=B)
        # a different value in every component (c + 1) * ref, so that a mixed-up component shows
        for c in range(P.celllength):
            P.field(c).copy_(igg.from_numpy(((c + 1) * zeroed).astype(npdt)))
        igg.update_halo(P)
        for c in range(P.celllength):
            got = igg.to_numpy(P.field(c))
            assert np.array_equal(got, ((c + 1) * ref).astype(npdt)), (B, c)
        # also with explicit halowidths, mixed with a plain array of the same element type only for B = 0
        if B == 0:
            Q = igg.zeros(shape, dtype=dtype)
            Q.copy_(igg.from_numpy(zeroed))
            for c in range(P.celllength):
                P.field(c).copy_(igg.from_numpy(((c + 1) * zeroed).astype(npdt)))
            igg.update_halo((P, g.halowidths), (Q, g.halowidths))
            assert np.array_equal(igg.to_numpy(Q), ref)
            assert np.array_equal(igg.to_numpy(P.field(3)), (4 * ref).astype(npdt))
    finally:
        igg.finalize_global_grid()
