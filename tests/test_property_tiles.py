"""Randomised check of the tile tables (hypothesis): for random process grids, periodicities, overlaps, halowidths,
field shapes (staggered up and down, narrow and wide in x), element sizes, pointer alignments and `dims=` orders,
fields in device memory, in symmetric allocations (direct-to-halo) and in page-locked host memory (whole-line
x-faces), the tile tables of both IPC protocols -- executed on the CPU as
the kernels address memory (tests/test_tiles_cpu.py) -- give exactly what the oracle's three dependent phases give.
No GPU needed."""
import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

from igg_b200 import _lib as L
from oracle import igg_oracle as O
from test_plan_vs_oracle import flat_bytes, rand_field
from test_tiles_cpu import run_tables, struct_field


@st.composite
def configs(draw):
    ndim = draw(st.sampled_from([1, 2, 3, 3, 3]))
    dims = [draw(st.sampled_from([1, 2, 3])) if d < ndim else 1 for d in range(3)]
    periods = [draw(st.integers(0, 1)) if d < ndim else 0 for d in range(3)]
    hw = [draw(st.integers(1, 3)) if d < ndim else 1 for d in range(3)]
    ol = [draw(st.integers(2 * hw[d], 2 * hw[d] + 2)) if d < ndim else 2 for d in range(3)]
    n = [draw(st.integers(ol[d] + 2 * hw[d] + 1, ol[d] + 2 * hw[d] + 7)) if d < ndim else 1 for d in range(3)]
    if draw(st.booleans()):
        n[0] += 128          # rows of 1 KiB and more (with 8-byte elements): wide enough for direct-to-halo
    if draw(st.booleans()):
        n[0] = (n[0] + 15) // 16 * 16      # rows on 64-byte boundaries: host fields take their x-faces through whole lines
    nf = draw(st.integers(1, 3))
    shapes = []
    for _ in range(nf):
        shapes.append(tuple(n[d] + draw(st.sampled_from([0, 0, 1, -1, 2])) for d in range(ndim)))
    elsize = draw(st.sampled_from([1, 2, 4, 8, 8, 12, 16]))
    order = draw(st.permutations([d + 1 for d in range(ndim)]))
    order = tuple(order[:draw(st.integers(1, ndim))])
    seed = draw(st.integers(0, 10_000))
    return dims, periods, hw, ol, n, shapes, elsize, order, seed


@settings(max_examples=90, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(configs(), st.sampled_from([L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128, L.IGG_MODE_IPC_AUTO]),
       st.sampled_from(["device", "symmetric", "host"]))
def test_random_tile_tables_match_oracle(cfg, mode, where):
    dims, periods, hw, ol, n, shapes, elsize, order, seed = cfg
    nprocs = dims[0] * dims[1] * dims[2]
    try:
        grids = O.make_grids(nprocs, *n, dimx=dims[0], dimy=dims[1], dimz=dims[2], periodx=periods[0], periody=periods[1],
                             periodz=periods[2], overlaps=tuple(ol), halowidths=tuple(hw))
    except O.IGGError:
        return                                           # argument combination the reference rejects
    hws = [tuple(hw)] * len(shapes)
    mk = struct_field if elsize == 12 else rand_field
    base = [[mk(s, elsize, seed + 1000 * r + 10 * i) for i, s in enumerate(shapes)] for r in range(nprocs)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    try:
        O.update_halo(grids, [[(a, hws[i]) for i, a in enumerate(fr)] for fr in want], dims=order)
    except O.IGGError:
        return                                           # e.g. a field without any halo: rejected by check_fields
    got = [[a.copy(order="F") for a in fr] for fr in base]
    auto = mode == L.IGG_MODE_IPC_AUTO                    # (small calls: auto = LL128 tiles)
    run_tables(grids, got, hws, order, L.IGG_MODE_IPC_LL128 if auto else mode, direct=where == "symmetric", host=where == "host")
    for r in range(nprocs):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), (r, i)
