"""Pins oracle/igg_oracle.py to the reference's own closed-form fixtures (CPU only, no GPU).

Every test cites the reference test it restates (paths relative to /root/reference/).  Base sizes
nx,ny,nz = 7,5,6 as in test/test_update_halo.jl:111-113.
"""
import numpy as np
import pytest

from oracle import igg_oracle as O

nx, ny, nz = 7, 5, 6


def R(*pairs):
    return [tuple(p) for p in pairs]


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:373-403  sendranges / recvranges, overlaps=(2,2,3)
# ---------------------------------------------------------------------------------------------------
def test_ranges_default_halowidth():
    (g,) = O.make_grids(1, nx, ny, nz, periodx=1, periody=1, periodz=1, overlaps=(2, 2, 3))
    P = (nx, ny, nz)
    A = (nx - 1, ny + 2, nz + 1)
    hw = g.halowidths
    assert hw == (1, 1, 1)
    sP = P
    assert O.sendranges(g, 1, 1, P, hw) == R((2, 2), (1, sP[1]), (1, sP[2]))
    assert O.sendranges(g, 2, 1, P, hw) == R((sP[0] - 1, sP[0] - 1), (1, sP[1]), (1, sP[2]))
    assert O.sendranges(g, 1, 2, P, hw) == R((1, sP[0]), (2, 2), (1, sP[2]))
    assert O.sendranges(g, 2, 2, P, hw) == R((1, sP[0]), (sP[1] - 1, sP[1] - 1), (1, sP[2]))
    assert O.sendranges(g, 1, 3, P, hw) == R((1, sP[0]), (1, sP[1]), (3, 3))
    assert O.sendranges(g, 2, 3, P, hw) == R((1, sP[0]), (1, sP[1]), (sP[2] - 2, sP[2] - 2))
    assert O.recvranges(g, 1, 1, P, hw) == R((1, 1), (1, sP[1]), (1, sP[2]))
    assert O.recvranges(g, 2, 1, P, hw) == R((sP[0], sP[0]), (1, sP[1]), (1, sP[2]))
    assert O.recvranges(g, 1, 2, P, hw) == R((1, sP[0]), (1, 1), (1, sP[2]))
    assert O.recvranges(g, 2, 2, P, hw) == R((1, sP[0]), (sP[1], sP[1]), (1, sP[2]))
    assert O.recvranges(g, 1, 3, P, hw) == R((1, sP[0]), (1, sP[1]), (1, 1))
    assert O.recvranges(g, 2, 3, P, hw) == R((1, sP[0]), (1, sP[1]), (sP[2], sP[2]))
    sA = A
    for n in (1, 2):
        with pytest.raises(O.IGGError):
            O.sendranges(g, n, 1, A, hw)
        with pytest.raises(O.IGGError):
            O.recvranges(g, n, 1, A, hw)
    assert O.sendranges(g, 1, 2, A, hw) == R((1, sA[0]), (4, 4), (1, sA[2]))
    assert O.sendranges(g, 2, 2, A, hw) == R((1, sA[0]), (sA[1] - 3, sA[1] - 3), (1, sA[2]))
    assert O.sendranges(g, 1, 3, A, hw) == R((1, sA[0]), (1, sA[1]), (4, 4))
    assert O.sendranges(g, 2, 3, A, hw) == R((1, sA[0]), (1, sA[1]), (sA[2] - 3, sA[2] - 3))
    assert O.recvranges(g, 1, 2, A, hw) == R((1, sA[0]), (1, 1), (1, sA[2]))
    assert O.recvranges(g, 2, 2, A, hw) == R((1, sA[0]), (sA[1], sA[1]), (1, sA[2]))
    assert O.recvranges(g, 1, 3, A, hw) == R((1, sA[0]), (1, sA[1]), (1, 1))
    assert O.recvranges(g, 2, 3, A, hw) == R((1, sA[0]), (1, sA[1]), (sA[2], sA[2]))


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:404-437  halowidth > 1: overlaps=(6,4,4), halowidths=(3,1,2), 13x9x9
# ---------------------------------------------------------------------------------------------------
def test_ranges_halowidth_gt1():
    NX, NY, NZ = 13, 9, 9
    (g,) = O.make_grids(1, NX, NY, NZ, periodx=1, periody=1, periodz=1, overlaps=(6, 4, 4), halowidths=(3, 1, 2))
    hw = g.halowidths
    P = (NX, NY, NZ)
    A = (NX - 1, NY + 2, NZ + 1)
    assert O.sendranges(g, 1, 1, P, hw) == R((4, 6), (1, NY), (1, NZ))
    assert O.sendranges(g, 2, 1, P, hw) == R((NX - 5, NX - 3), (1, NY), (1, NZ))
    assert O.sendranges(g, 1, 2, P, hw) == R((1, NX), (4, 4), (1, NZ))
    assert O.sendranges(g, 2, 2, P, hw) == R((1, NX), (NY - 3, NY - 3), (1, NZ))
    assert O.sendranges(g, 1, 3, P, hw) == R((1, NX), (1, NY), (3, 4))
    assert O.sendranges(g, 2, 3, P, hw) == R((1, NX), (1, NY), (NZ - 3, NZ - 2))
    assert O.recvranges(g, 1, 1, P, hw) == R((1, 3), (1, NY), (1, NZ))
    assert O.recvranges(g, 2, 1, P, hw) == R((NX - 2, NX), (1, NY), (1, NZ))
    assert O.recvranges(g, 1, 2, P, hw) == R((1, NX), (1, 1), (1, NZ))
    assert O.recvranges(g, 2, 2, P, hw) == R((1, NX), (NY, NY), (1, NZ))
    assert O.recvranges(g, 1, 3, P, hw) == R((1, NX), (1, NY), (1, 2))
    assert O.recvranges(g, 2, 3, P, hw) == R((1, NX), (1, NY), (NZ - 1, NZ))
    a = A
    for n in (1, 2):
        with pytest.raises(O.IGGError):
            O.sendranges(g, n, 1, A, hw)
        with pytest.raises(O.IGGError):
            O.recvranges(g, n, 1, A, hw)
    assert O.sendranges(g, 1, 2, A, hw) == R((1, a[0]), (6, 6), (1, a[2]))
    assert O.sendranges(g, 2, 2, A, hw) == R((1, a[0]), (a[1] - 5, a[1] - 5), (1, a[2]))
    assert O.sendranges(g, 1, 3, A, hw) == R((1, a[0]), (1, a[1]), (4, 5))
    assert O.sendranges(g, 2, 3, A, hw) == R((1, a[0]), (1, a[1]), (a[2] - 4, a[2] - 3))
    assert O.recvranges(g, 1, 2, A, hw) == R((1, a[0]), (1, 1), (1, a[2]))
    assert O.recvranges(g, 2, 2, A, hw) == R((1, a[0]), (a[1], a[1]), (1, a[2]))
    assert O.recvranges(g, 1, 3, A, hw) == R((1, a[0]), (1, a[1]), (1, 2))
    assert O.recvranges(g, 2, 3, A, hw) == R((1, a[0]), (1, a[1]), (a[2] - 1, a[2]))


def index_coded(shape):
    """[iz*1e2 + iy*1e1 + ix ...] of test/test_update_halo.jl:441,639."""
    s = O.size3(shape)
    ix = np.arange(1, s[0] + 1).reshape(-1, 1, 1)
    iy = np.arange(1, s[1] + 1).reshape(1, -1, 1)
    iz = np.arange(1, s[2] + 1).reshape(1, 1, -1)
    return np.asfortranarray((iz * 100.0 + iy * 10.0 + ix).reshape(tuple(shape), order="F"))


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:438-495  write_h2h! / read_h2h! (incl. halowidth > 1)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("hw,ranges_by_dim", [
    ((1, 1, 1), {1: R((2, 2), (1, ny), (1, nz)), 2: R((1, nx), (3, 3), (1, nz)), 3: R((1, nx), (1, ny), (4, 4))}),
    ((3, 1, 2), {1: R((4, 6), (1, ny), (1, nz)), 2: R((1, nx), (4, 4), (1, nz)), 3: R((1, nx), (1, ny), (3, 4))}),
])
def test_write_read_h2h(hw, ranges_by_dim):
    P = index_coded((nx, ny, nz))
    P2 = np.zeros_like(P)
    for dim, rng in ranges_by_dim.items():
        buf = np.zeros(O.halosize(dim, P, hw), order="F")
        O.write_h2h(buf, P, rng)
        sl = tuple(slice(lo - 1, hi) for lo, hi in rng)
        assert np.array_equal(buf.ravel(order="F"), P[sl].ravel(order="F"))
        O.read_h2h(buf, P2, rng)
        assert np.array_equal(buf.ravel(order="F"), P2[sl].ravel(order="F"))


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:635-723  iwrite_sendbufs!: buffer contents, overlaps=(4,2,3), halowidths=(2,1,1)
# ---------------------------------------------------------------------------------------------------
def test_sendbuf_contents():
    (g,) = O.make_grids(1, nx, ny, nz, periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1))
    hw = g.halowidths
    P = index_coded((nx, ny, nz))
    A = index_coded((nx - 1, ny + 2, nz + 1))

    def packed(F, n, dim):
        buf = np.zeros(O.halosize(dim, F, hw), order="F")
        O.write_h2h(buf, F, O.sendranges(g, n, dim, F, hw))
        return buf.ravel(order="F")

    f = lambda x: x.ravel(order="F")
    assert np.array_equal(packed(P, 1, 1), f(P[2:4, :, :]))            # P.A[3:4,:,:]
    assert np.array_equal(packed(P, 2, 1), f(P[-4:-2, :, :]))          # P.A[end-3:end-2,:,:]
    assert O.ol(g, 1, A) < 2 * hw[0]                                   # A has no halo in x (buffer stays 0)
    assert np.array_equal(packed(P, 1, 2), f(P[:, 1, :]))              # P.A[:,2,:]
    assert np.array_equal(packed(A, 1, 2), f(A[:, 3, :]))              # A.A[:,4,:]
    assert np.array_equal(packed(P, 2, 2), f(P[:, -2, :]))             # P.A[:,end-1,:]
    assert np.array_equal(packed(A, 2, 2), f(A[:, -4, :]))             # A.A[:,end-3,:]
    assert np.array_equal(packed(P, 1, 3), f(P[:, :, 2]))              # P.A[:,:,3]
    assert np.array_equal(packed(A, 1, 3), f(A[:, :, 3]))              # A.A[:,:,4]
    assert np.array_equal(packed(P, 2, 3), f(P[:, :, -3]))             # P.A[:,:,end-2]
    assert np.array_equal(packed(A, 2, 3), f(A[:, :, -4]))             # A.A[:,:,end-3]


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:724-837  iread_recvbufs!: where the receive buffers land
# ---------------------------------------------------------------------------------------------------
def test_recvbuf_landing():
    (g,) = O.make_grids(1, nx, ny, nz, periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1))
    hw = g.halowidths
    P = np.zeros((nx, ny, nz), order="F")
    A = np.zeros((nx - 1, ny + 2, nz + 1), order="F")
    expect = {
        (1, 1): (lambda X: X[0:2, :, :]), (2, 1): (lambda X: X[-2:, :, :]),
        (1, 2): (lambda X: X[:, 0, :]), (2, 2): (lambda X: X[:, -1, :]),
        (1, 3): (lambda X: X[:, :, 0]), (2, 3): (lambda X: X[:, :, -1]),
    }
    for dim in (1, 2, 3):
        for n in (1, 2):
            for i, F in enumerate((P, A), start=1):
                if O.ol(g, dim, F) < 2 * hw[dim - 1]:
                    assert np.all(expect[(n, dim)](F) == 0.0)      # no halo -> untouched (test:754,766)
                    continue
                val = dim * 100.0 + n * 10.0 + i
                buf = np.full(O.halosize(dim, F, hw), val, order="F")
                O.read_h2h(buf, F, O.recvranges(g, n, dim, F, hw))
                assert np.all(expect[(n, dim)](F) == val)


# ---------------------------------------------------------------------------------------------------
# test/test_update_halo.jl:975-1344  full halo updates, one rank, periodic: field == reference afterwards
# ---------------------------------------------------------------------------------------------------
def zero_boundaries(A, widths):
    B = A.copy(order="F")
    for d in range(A.ndim):
        w = widths[d]
        idx = [slice(None)] * A.ndim
        idx[d] = list(range(w)) + list(range(A.shape[d] - w, A.shape[d]))
        B[tuple(idx)] = 0
    return B


PERIODIC_CASES = [
    # (grid args, kwargs, field shapes, zeroed boundary widths, test line)
    ((nx, 1, 1), dict(periodx=1), [(nx,)], (1,), 977),
    ((nx, ny, 1), dict(periodx=1, periody=1), [(nx, ny)], (1, 1), 990),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz)], (1, 1, 1), 1004),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
     [(nx, ny, nz)], (2, 1, 1), 1019),
    ((nx, 1, 1), dict(periodx=1), [(nx + 1,)], (1,), 1057),
    ((nx, ny, 1), dict(periodx=1, periody=1), [(nx, ny + 1)], (1, 1), 1070),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz + 1)], (1, 1, 1), 1084),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
     [(nx + 1, ny, nz)], (2, 1, 1), 1099),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz + 1), (nx + 1, ny, nz)], (1, 1, 1), 1293),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
     [(nx, ny, nz + 1), (nx + 1, ny, nz)], (2, 1, 1), 1318),
]


@pytest.mark.parametrize("gargs,kw,shapes,widths,line", PERIODIC_CASES, ids=[f"L{c[4]}" for c in PERIODIC_CASES])
@pytest.mark.parametrize("dtype", [np.float64, np.float32, np.complex64])
def test_full_update_periodic_single_rank(gargs, kw, shapes, widths, line, dtype):
    (g,) = O.make_grids(1, *gargs, **kw)
    refs, fields = [], []
    for shp in shapes:
        ref = O.coord_field(g, shp, dtype=np.float64)
        if np.issubdtype(dtype, np.complexfloating):
            ref = np.asfortranarray((1 + 1j) * ref).astype(dtype)       # test:1169
        else:
            ref = ref.astype(dtype)
        A = zero_boundaries(ref, widths)
        assert not np.array_equal(A, ref)                                # @require
        refs.append(ref)
        fields.append(A)
    O.update_halo([g], [fields])
    for A, ref in zip(fields, refs):
        assert np.array_equal(A, ref)


# test/test_update_halo.jl:1135-1165  no halo in one dimension: that dimension's boundary stays zero
@pytest.mark.parametrize("gargs,kw,shape,nohalo_dim", [
    ((nx, ny, 1), dict(periodx=1, periody=1), (nx - 1, ny + 2), 0),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), (nx + 2, ny - 1, nz + 1), 1),
])
def test_full_update_no_halo_in_one_dim(gargs, kw, shape, nohalo_dim):
    (g,) = O.make_grids(1, *gargs, **kw)
    ref = O.coord_field(g, shape)
    A = zero_boundaries(ref, (1,) * len(shape))
    O.update_halo([g], [[A]])
    inner = [slice(None)] * len(shape)
    inner[nohalo_dim] = slice(1, -1)
    assert np.array_equal(A[tuple(inner)], ref[tuple(inner)])
    edge = [slice(None)] * len(shape)
    edge[nohalo_dim] = [0, -1]
    assert np.all(A[tuple(edge)] == 0.0)


# test/test_update_halo.jl:1034-1054, 1114-1133  not periodic (any nprocs): interior kept, outer boundary stays 0,
# inner boundaries equal the reference
@pytest.mark.parametrize("nprocs", [1, 2, 4, 8, 12])
@pytest.mark.parametrize("stag", [(0, 0, 0), (0, 0, 1)])
def test_full_update_not_periodic(nprocs, stag):
    grids = O.make_grids(nprocs, nx, ny, nz)
    shape = (nx + stag[0], ny + stag[1], nz + stag[2])
    refs = [O.coord_field(g, shape) for g in grids]
    fields = [[zero_boundaries(r, (1, 1, 1))] for r in refs]
    O.update_halo(grids, fields)
    for g, (P,), ref in zip(grids, fields, refs):
        assert np.array_equal(P[1:-1, 1:-1, 1:-1], ref[1:-1, 1:-1, 1:-1])
        planes = [(0, P[0], ref[0], 0), (0, P[-1], ref[-1], -1), (1, P[:, 0], ref[:, 0], 0),
                  (1, P[:, -1], ref[:, -1], -1), (2, P[:, :, 0], ref[:, :, 0], 0), (2, P[:, :, -1], ref[:, :, -1], -1)]
        for d, p, r, side in planes:
            at_edge = (g.coords[d] == 0) if side == 0 else (g.coords[d] == g.dims[d] - 1)
            if at_edge:
                assert np.all(p == 0.0)
            else:
                assert np.array_equal(p[1:-1, 1:-1], r[1:-1, 1:-1])


# periodic, several ranks: the coordinate-coded field is globally consistent, so the update must restore it
@pytest.mark.parametrize("nprocs", [2, 4, 8])
def test_full_update_periodic_multirank(nprocs):
    grids = O.make_grids(nprocs, nx, ny, nz, periodx=1, periody=1, periodz=1)
    refs = [O.coord_field(g, (nx, ny, nz)) for g in grids]
    fields = [[zero_boundaries(r, (1, 1, 1))] for r in refs]
    O.update_halo(grids, fields)
    for (P,), ref in zip(fields, refs):
        assert np.array_equal(P, ref)


# the result does not depend on the order of `dims`
def test_dimension_order_independence():
    grids = O.make_grids(8, 9, 8, 7, periodx=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1))
    base = [[O.random_bits_field((9, 8, 8), 8, 100 + g.me), O.random_bits_field((10, 8, 7), 8, 200 + g.me)]
            for g in grids]
    results = []
    import itertools
    for order in itertools.permutations((1, 2, 3)):
        f = [[a.copy(order="F") for a in fr] for fr in base]
        O.update_halo(grids, f, dims=order)
        results.append(f)
    for res in results[1:]:
        for fr, fr0 in zip(res, results[0]):
            for a, b in zip(fr, fr0):
                assert np.array_equal(a, b)


# ---------------------------------------------------------------------------------------------------
# check_fields errors (src/update_halo.jl:418-480)
# ---------------------------------------------------------------------------------------------------
def test_check_fields_errors():
    (g,) = O.make_grids(1, nx, ny, nz, periodx=1, periody=1, periodz=1)
    P = np.zeros((nx, ny, nz), order="F")
    with pytest.raises(O.IGGError, match="has a halowidth less than 1"):
        O.update_halo([g], [[(P, (0, 1, 1))]])
    Q = np.zeros((nx - 1, ny - 1, nz - 1), order="F")
    with pytest.raises(O.IGGError, match="has no halo; remove it from the call"):
        O.update_halo([g], [[Q]])
    with pytest.raises(O.IGGError, match="is a duplicate of the one at the position 1"):
        O.update_halo([g], [[P, P]])
    with pytest.raises(O.IGGError, match="different type than the first field"):
        O.update_halo([g], [[P, np.zeros((nx, ny, nz), dtype=np.float32, order="F")]])


# ---------------------------------------------------------------------------------------------------
# init: topology + nxyz_g (test/test_init_global_grid.jl:21-97) and argument errors (:99-133)
# ---------------------------------------------------------------------------------------------------
def test_topology_and_global_size():
    (g,) = O.make_grids(1, 4, 4, 4, dimx=1, dimy=1, dimz=1, periodz=1)
    assert g.nxyz_g == (4, 4, 2) and g.neighbors == ((-1, -1, 0), (-1, -1, 0))      # test_init:73
    for n, want in [(1, (1, 1, 1)), (2, (2, 1, 1)), (4, (2, 2, 1)), (8, (2, 2, 2)), (6, (3, 2, 1)), (12, (3, 2, 2))]:
        assert O.dims_create(n, (0, 0, 0)) == want
    assert O.dims_create(8, (0, 0, 1)) == (4, 2, 1)
    grids = O.make_grids(8, 5, 5, 5, periodx=1)
    assert grids[0].dims == (2, 2, 2) and grids[0].nxyz_g == (6, 8, 8)
    assert [g.coords for g in grids] == [(0, 0, 0), (0, 0, 1), (0, 1, 0), (0, 1, 1), (1, 0, 0), (1, 0, 1),
                                         (1, 1, 0), (1, 1, 1)]
    assert grids[0].neighbors == ((4, -1, -1), (4, 2, 1))
    assert grids[7].neighbors == ((3, 5, 6), (3, -1, -1))


@pytest.mark.parametrize("args,kw", [
    ((1, 4, 4), {}), ((4, 1, 4), {}), ((4, 4, 1), dict(dimz=3)), ((4, 4, 4), dict(periodx=2)),
    ((4, 4, 4), dict(halowidths=(0, 1, 1))), ((4, 4, 4), dict(overlaps=(2, 2, 2), halowidths=(2, 1, 1))),
    ((2, 4, 4), dict(periodx=1)), ((4, 4, 4), dict(dimx=-1)),
])
def test_init_argument_errors(args, kw):
    with pytest.raises(O.IGGError):
        O.make_grids(1, *args, **kw)


# ---------------------------------------------------------------------------------------------------
# tools: nx_g / ix_g / x_g known answers (test/test_tools.jl:15-106 and 344-429, simulated 3x3x3)
# ---------------------------------------------------------------------------------------------------
def test_tools_single_rank():
    (g,) = O.make_grids(1, 5, 5, 5, dimx=1, dimy=1, dimz=1, periodz=1)
    P, Vx, Vz, A, Sxz = (5, 5, 5), (6, 5, 5), (5, 5, 6), (7, 5, 7), (3, 4, 3)
    assert [O.n_g(g, d) for d in (1, 2, 3)] == [5, 5, 3]
    assert [O.n_g(g, d + 1, Vx[d]) for d in range(3)] == [6, 5, 3]
    assert [O.n_g(g, d + 1, A[d]) for d in range(3)] == [7, 5, 3]
    assert [O.n_g(g, d + 1, Sxz[d]) for d in range(3)] == [3, 4, 3]
    iz = lambda S, w=True: [O.i_g(g, 3, i, S[2], w) for i in range(1, S[2] + 1)]
    assert iz(P) == [3, 1, 2, 3, 1] and iz(P, False) == [0, 1, 2, 3, 4]
    assert iz(Vz) == [3, 1, 2, 3, 1, 2] and iz(Vz, False) == [0, 1, 2, 3, 4, 5]
    assert iz(A) == [2, 3, 1, 2, 3, 1, 2] and iz(A, False) == [-1, 0, 1, 2, 3, 4, 5]
    assert iz(Sxz) == [1, 2, 3]
    assert [O.i_g(g, 1, i, A[0]) for i in range(1, 8)] == [1, 2, 3, 4, 5, 6, 7]
    dx = 8 / (O.n_g(g, 1) - 1)
    dz = 8 / (O.n_g(g, 3) - 1)
    xs = lambda S: [O.x_g(g, 1, i, dx, S[0]) for i in range(1, S[0] + 1)]
    ys = lambda S: [O.x_g(g, 2, i, dx, S[1]) for i in range(1, S[1] + 1)]
    zs = lambda S, w=True: [O.x_g(g, 3, i, dz, S[2], w) for i in range(1, S[2] + 1)]
    assert xs(P) == [0.0, 2.0, 4.0, 6.0, 8.0]
    assert zs(P) == [8.0, 0.0, 4.0, 8.0, 0.0] and zs(P, False) == [-4.0, 0.0, 4.0, 8.0, 12.0]
    assert xs(Vx) == [-1.0, 1.0, 3.0, 5.0, 7.0, 9.0]
    assert zs(Vz) == [6.0, -2.0, 2.0, 6.0, -2.0, 2.0] and zs(Vz, False) == [-6.0, -2.0, 2.0, 6.0, 10.0, 14.0]
    assert xs(A) == [-2.0, 0.0, 2.0, 4.0, 6.0, 8.0, 10.0]
    assert zs(A) == [4.0, 8.0, 0.0, 4.0, 8.0, 0.0, 4.0]
    assert xs(Sxz) == [2.0, 4.0, 6.0] and ys(Sxz) == [1.0, 3.0, 5.0, 7.0] and zs(Sxz) == [0.0, 4.0, 8.0]


def test_tools_simulated_3x3x3():
    grids = O.make_grids(27, 5, 5, 5, dimx=3, dimy=3, dimz=3, periodz=1)
    by = {g.coords: g for g in grids}
    P, A = (5, 5, 5), (6, 3, 7)
    g0 = grids[0]
    assert g0.nxyz_g == (11, 11, 9)
    assert [O.n_g(g0, d + 1, A[d]) for d in range(3)] == [12, 9, 9]
    ix = lambda c, S: [O.i_g(by[(c, 0, 0)], 1, i, S[0]) for i in range(1, S[0] + 1)]
    iz = lambda c, S, w=True: [O.i_g(by[(0, 0, c)], 3, i, S[2], w) for i in range(1, S[2] + 1)]
    assert ix(0, P) == [1, 2, 3, 4, 5] and ix(1, P) == [4, 5, 6, 7, 8] and ix(2, P) == [7, 8, 9, 10, 11]
    assert iz(0, P) == [9, 1, 2, 3, 4] and iz(1, P) == [3, 4, 5, 6, 7] and iz(2, P) == [6, 7, 8, 9, 1]
    assert iz(2, P, False) == [6, 7, 8, 9, 10]
    assert ix(2, A) == [7, 8, 9, 10, 11, 12]
    assert iz(0, A) == [8, 9, 1, 2, 3, 4, 5] and iz(2, A) == [5, 6, 7, 8, 9, 1, 2]
    dx = 20 / (g0.nxyz_g[0] - 1)
    dz = 16 / (g0.nxyz_g[2] - 1)
    xs = lambda c, S: [O.x_g(by[(c, 0, 0)], 1, i, dx, S[0]) for i in range(1, S[0] + 1)]
    ys = lambda c, S: [O.x_g(by[(0, c, 0)], 2, i, dx, S[1]) for i in range(1, S[1] + 1)]
    zs = lambda c, S, w=True: [O.x_g(by[(0, 0, c)], 3, i, dz, S[2], w) for i in range(1, S[2] + 1)]
    assert xs(1, P) == [6.0, 8.0, 10.0, 12.0, 14.0] and xs(2, P) == [12.0, 14.0, 16.0, 18.0, 20.0]
    assert zs(0, P) == [16.0, 0.0, 2.0, 4.0, 6.0] and zs(2, P) == [10.0, 12.0, 14.0, 16.0, 0.0]
    assert xs(0, A) == [-1.0, 1.0, 3.0, 5.0, 7.0, 9.0] and xs(2, A) == [11.0, 13.0, 15.0, 17.0, 19.0, 21.0]
    assert ys(1, A) == [8.0, 10.0, 12.0]
    assert zs(0, A) == [14.0, 16.0, 0.0, 2.0, 4.0, 6.0, 8.0]
    assert zs(2, A, False) == [8.0, 10.0, 12.0, 14.0, 16.0, 18.0, 20.0]


# ---------------------------------------------------------------------------------------------------
# gather! (test/test_gather.jl:19-151)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("nprocs", [1, 2, 4, 8])
@pytest.mark.parametrize("nd", [1, 2, 3])
def test_gather(nprocs, nd):
    sizes = [(nx, 1, 1), (nx, ny, 1), (nx, ny, nz)][nd - 1]
    grids = O.make_grids(nprocs, *sizes, overlaps=(0, 0, 0))
    shp = sizes[:nd]
    blocks = [O.coord_field(g, shp) for g in grids]
    dims = grids[0].dims
    gshape = tuple(s * d for s, d in zip(shp, dims[:nd]))
    P_g = np.zeros(gshape, order="F")
    O.gather(grids, blocks, P_g)
    # P_g_ref = x_g-coded global array shifted to start at 0 (test_gather.jl:41-43); with overlaps 0 and dx=1 the
    # global coordinate of global index i is i-1.
    ref = np.zeros(O.size3(gshape), order="F")
    for d in range(nd):
        sh = [1, 1, 1]
        sh[d] = gshape[d]
        ref += (10.0 ** d) * np.arange(gshape[d]).reshape(sh)
    assert np.array_equal(P_g, ref.reshape(gshape, order="F"))


def test_gather_argument_errors_and_root():
    grids = O.make_grids(2, nx, ny, nz)
    d = grids[0].dims
    C = [np.zeros((nx, ny, nz), order="F") for _ in grids]
    with pytest.raises(O.IGGError, match="must be equal to the product"):
        O.gather(grids, C, np.zeros((nx * d[0], ny * d[1], nz * d[2] + 2), order="F"))
    with pytest.raises(O.IGGError, match="can't be `nothing` on the root"):
        O.gather(grids, C, None)
    A = [np.ones(nx) for _ in grids]
    A_g = np.zeros(nx * d[0])
    O.gather(grids, A, A_g, root=1)                                            # test_gather.jl:126-137
    assert np.all(A_g == 1.0)
    # 1-D block into a 3-D global array (test_gather.jl:73-79)
    P_g = np.zeros((nx * d[0], d[1], d[2]), order="F")
    O.gather(grids, [np.arange(nx, dtype=np.float64) + 10 * g.me for g in grids], P_g)
    assert np.array_equal(P_g[:, 0, 0], np.concatenate([np.arange(nx), np.arange(nx) + 10]))
