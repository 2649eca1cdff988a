"""GPU parity tests (run with `-m gpu` on a B200): the CUDA path, called through the Python host mirror ->
ctypes -> C ABI of libigg_b200.so, against the oracle on identical inputs.  Bit-exact (pure copying).

The single-rank cases restate the reference's own full-update tests (test/test_update_halo.jl:975-1344): a rank
that is its own neighbour in a periodic dimension exercises pack -> exchange -> unpack end to end.
Multi-rank cases live in test_gpu_multirank.py (they need more than one GPU).
"""
import numpy as np
import pytest
import torch

import igg_b200 as igg
from oracle import igg_oracle as O

pytestmark = pytest.mark.gpu

nx, ny, nz = 7, 5, 6


@pytest.fixture(autouse=True)
def _need_gpu_and_clean_grid():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    yield
    if igg.grid_is_initialized():
        igg.finalize_global_grid()


def zero_boundaries(A, widths):
    B = A.copy(order="F")
    for d in range(A.ndim):
        idx = [slice(None)] * A.ndim
        idx[d] = list(range(widths[d])) + list(range(A.shape[d] - widths[d], A.shape[d]))
        B[tuple(idx)] = 0
    return B


def bits_equal(t, ref):
    got = igg.to_numpy(t).reshape(-1, order="F").view(np.uint8)
    return np.array_equal(got, np.asfortranarray(ref).reshape(-1, order="F").view(np.uint8))


def rand_any(shape, elsize, seed):
    """F-ordered array of `elsize`-byte elements holding full-range random bits."""
    if elsize == 16:
        n = int(np.prod(shape))
        return O.random_bits_field((2 * n,), 8, seed).view(np.complex128).reshape(shape, order="F")
    return O.random_bits_field(shape, elsize, seed)


REFERENCE_CASES = [
    # (grid args, kwargs, shapes, zeroed widths, reference test line)
    ((nx, 1, 1), dict(periodx=1), [(nx,)], (1,), 977),
    ((nx, ny, 1), dict(periodx=1, periody=1), [(nx, ny)], (1, 1), 990),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz)], (1, 1, 1), 1004),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), [(nx, ny, nz)], (2, 1, 1), 1019),
    ((nx, 1, 1), dict(periodx=1), [(nx + 1,)], (1,), 1057),
    ((nx, ny, 1), dict(periodx=1, periody=1), [(nx, ny + 1)], (1, 1), 1070),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz + 1)], (1, 1, 1), 1084),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), [(nx + 1, ny, nz)], (2, 1, 1), 1099),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), [(nx, ny, nz + 1), (nx + 1, ny, nz)], (1, 1, 1), 1293),
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
     [(nx, ny, nz + 1), (nx + 1, ny, nz)], (2, 1, 1), 1318),
]


@pytest.mark.parametrize("gargs,kw,shapes,widths,line", REFERENCE_CASES, ids=[f"L{c[4]}" for c in REFERENCE_CASES])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32, torch.complex64, torch.int16])
def test_reference_full_update_fixtures(gargs, kw, shapes, widths, line, dtype):
    """Coordinate-coded field, boundaries zeroed, update_halo!, field == reference again."""
    igg.init_global_grid(*gargs, quiet=True, **kw)
    (g,) = O.make_grids(1, *gargs, **kw)
    npdt = {torch.float64: np.float64, torch.float32: np.float32, torch.complex64: np.complex64, torch.int16: np.int16}[dtype]
    refs, fields = [], []
    for shp in shapes:
        ref = O.coord_field(g, shp)
        ref = ((1 + 1j) * ref).astype(npdt) if dtype == torch.complex64 else ref.astype(npdt)
        ref = np.asfortranarray(ref)
        A = zero_boundaries(ref, widths)
        assert not np.array_equal(A, ref)
        refs.append(ref)
        fields.append(igg.from_numpy(A))
    igg.update_halo(*fields)
    for t, ref in zip(fields, refs):
        assert bits_equal(t, ref)
    # the x_g-coded values come out of the package's own tools as well
    assert igg.x_g(2, 1.0, fields[0]) == O.x_g(g, 1, 2, 1.0, shapes[0][0])


import golden_cases as G  # noqa: E402


@pytest.mark.parametrize("ci,case", G.cases(1), ids=[c["name"] for _, c in G.cases(1)])
def test_golden_single_rank(ci, case):
    """The committed closed-form vectors (tests/golden/, built without the oracle): CUDA path == expected bits."""
    igg.init_global_grid(*case["grid"], quiet=True, **G.kwargs(case))
    hws = G.halowidths(case, igg.global_grid().halowidths)
    fields = [igg.from_numpy(a) for a in G.arrays(ci, 0, "in")]
    igg.update_halo(*[(t, hw) for t, hw in zip(fields, hws)], dims=tuple(case["dims"]))
    for i, (t, want) in enumerate(zip(fields, G.arrays(ci, 0, "out"))):
        assert bits_equal(t, want), (case["name"], i)


@pytest.mark.parametrize("gargs,kw,shape,nohalo", [
    ((nx, ny, 1), dict(periodx=1, periody=1), (nx - 1, ny + 2), 0),                 # test/test_update_halo.jl:1135
    ((nx, ny, nz), dict(periodx=1, periody=1, periodz=1), (nx + 2, ny - 1, nz + 1), 1),  # :1150
])
def test_no_halo_in_one_dim(gargs, kw, shape, nohalo):
    igg.init_global_grid(*gargs, quiet=True, **kw)
    (g,) = O.make_grids(1, *gargs, **kw)
    ref = O.coord_field(g, shape)
    A = zero_boundaries(ref, (1,) * len(shape))
    t = igg.from_numpy(A)
    igg.update_halo(t)
    got = igg.to_numpy(t)
    inner = [slice(None)] * len(shape)
    inner[nohalo] = slice(1, -1)
    assert np.array_equal(got[tuple(inner)], ref[tuple(inner)])
    edge = [slice(None)] * len(shape)
    edge[nohalo] = [0, -1]
    assert np.all(got[tuple(edge)] == 0.0)


def test_not_periodic_single_rank_is_a_no_op():
    """test/test_update_halo.jl:1034-1054 with one rank: no neighbours, nothing may change."""
    igg.init_global_grid(nx, ny, nz, quiet=True)
    a = O.random_bits_field((nx, ny, nz), 8, 5)
    t = igg.from_numpy(a)
    igg.update_halo(t)
    assert bits_equal(t, a)


RANDOM_CASES = [
    # (grid size, kwargs, field shapes, per-field halowidths or None, elsize, dims)
    ((13, 9, 10), dict(periodx=1, periody=1, periodz=1, overlaps=(6, 4, 4), halowidths=(3, 1, 2)),
     [(13, 9, 10), (14, 9, 10), (13, 9, 11), (12, 11, 11)], None, 4, (3, 1, 2)),
    ((13, 9, 10), dict(periodx=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2)),
     [(13, 9, 10), (13, 10, 10)], [(2, 2, 2), (1, 1, 2)], 8, (3, 1, 2)),
    ((16, 12, 9), dict(periodx=1, periody=1, periodz=1), [(16, 12, 9), (17, 12, 9), (16, 13, 9), (16, 12, 10)], None, 8, (1, 2)),
    ((16, 12, 9), dict(periody=1), [(16, 12, 9)], None, 2, (2,)),
    ((33, 17, 20), dict(periodx=1, periody=1, periodz=1), [(33, 17, 20), (34, 17, 20)], None, 1, (3, 1, 2)),
    ((33, 17, 20), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2)),
     [(33, 17, 20), (33, 18, 21)], None, 16, (3, 1, 2)),
    ((40, 1, 1), dict(periodx=1, overlaps=(6, 2, 2), halowidths=(3, 1, 1)), [(40,), (41,)], None, 8, (3, 1, 2)),
    ((40, 30, 1), dict(periodx=1, periody=1), [(40, 30), (41, 30), (40, 31)], None, 4, (3, 1, 2)),
]


@pytest.mark.parametrize("case", range(len(RANDOM_CASES)))
def test_random_bits_vs_oracle(case):
    """Full-range random bit patterns (NaNs, denormals): any arithmetic or canonicalisation would show."""
    gsz, kw, shapes, hws, elsize, dims = RANDOM_CASES[case]
    igg.init_global_grid(*gsz, quiet=True, **kw)
    (g,) = O.make_grids(1, *gsz, **kw)
    hws = hws or [g.halowidths] * len(shapes)
    tdt = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64, 16: torch.complex128}[elsize]
    base = [rand_any(s, elsize, 1234 + i) for i, s in enumerate(shapes)]
    want = [a.copy(order="F") for a in base]
    O.update_halo([g], [[(a, hw) for a, hw in zip(want, hws)]], dims=dims)
    ts = [igg.from_numpy(a) for a in base]
    assert all(t.dtype == tdt for t in ts)
    for rep in range(3):          # idempotent: repeating the update changes nothing
        igg.update_halo(*[(t, hw) for t, hw in zip(ts, hws)], dims=dims)
        for t, w in zip(ts, want):
            assert bits_equal(t, w)


def test_misaligned_base_pointer():
    """A contiguous view that starts 8 bytes into an allocation: vector width must drop to 8 bytes."""
    igg.init_global_grid(16, 12, 9, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, 16, 12, 9, periodx=1, periody=1, periodz=1)
    a = O.random_bits_field((16, 12, 9), 8, 42)
    buf = torch.zeros(16 * 12 * 9 + 1, dtype=torch.int64, device="cuda")
    view = buf[1:].view(9, 12, 16).permute(2, 1, 0)
    view.copy_(igg.from_numpy(a))
    assert view.data_ptr() % 16 == 8
    want = a.copy(order="F")
    O.update_halo([g], [[want]])
    igg.update_halo(view)
    assert bits_equal(view, want)
    assert int(buf[0]) == 0


def test_full_size_256_f64_cell_id_pattern():
    """BASELINE config 2 on one GPU (periodic self-exchange), cell-id pattern with 0xA5-poisoned halos,
    whole 128 MiB array compared bitwise with the oracle; plus the halo-consistency property."""
    n = 256
    igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, n, n, n, periodx=1, periody=1, periodz=1)
    ix = np.arange(n, dtype=np.uint64).reshape(-1, 1, 1)
    iy = np.arange(n, dtype=np.uint64).reshape(1, -1, 1)
    iz = np.arange(n, dtype=np.uint64).reshape(1, 1, -1)
    a = np.asfortranarray((np.uint64(1) << np.uint64(48)) | (iz << np.uint64(32)) | (iy << np.uint64(16)) | ix)
    poison = np.uint64(0xA5A5A5A5A5A5A5A5)
    for d in range(3):
        idx = [slice(None)] * 3
        idx[d] = [0, n - 1]
        a[tuple(idx)] = poison
    want = a.copy(order="F")
    O.update_halo([g], [[want]])
    t = igg.from_numpy(a)
    igg.update_halo(t)
    got = igg.to_numpy(t).view(np.uint64)
    assert np.array_equal(got, want)
    assert not np.any(got == poison)
    # property: with overlap 2 and halowidth 1 the periodic wrap means plane 0 == plane n-2 and plane n-1 == plane 1
    for d in range(3):
        lo = [slice(None)] * 3; hi = [slice(None)] * 3
        lo[d], hi[d] = 0, n - 2
        assert np.array_equal(got[tuple(lo)], got[tuple(hi)])
        lo[d], hi[d] = n - 1, 1
        assert np.array_equal(got[tuple(lo)], got[tuple(hi)])


def test_512_f32_hw2_multifield_properties():
    """BASELINE config 4 shape on one GPU: 2 fields 512^3 Float32, overlaps 4, halowidths 2.  Too big for the
    numpy oracle to be quick on whole arrays, so: size-independent wrap properties on the GPU + oracle on planes."""
    n = 512
    igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2), quiet=True)
    torch.manual_seed(1234)
    fields = []
    for i in range(2):
        t = igg.zeros((n, n, n), dtype=torch.int32)
        t.copy_(torch.randint(-2**31, 2**31 - 1, (n, n, n), dtype=torch.int32, device="cuda"))
        fields.append(t)
    before = [t.clone() for t in fields]
    igg.update_halo(*fields)
    for t, b in zip(fields, before):
        # interior untouched
        assert torch.equal(t[2:-2, 2:-2, 2:-2], b[2:-2, 2:-2, 2:-2])
        # halo [0:2] == cells [n-4:n-2], halo [n-2:n] == cells [2:4], in every dimension (after the update)
        for d in range(3):
            assert torch.equal(t.narrow(d, 0, 2), t.narrow(d, n - 4, 2))
            assert torch.equal(t.narrow(d, n - 2, 2), t.narrow(d, 2, 2))
    # idempotence
    snap = [t.clone() for t in fields]
    igg.update_halo(*fields)
    for t, s in zip(fields, snap):
        assert torch.equal(t, s)


def test_pinned_host_fields_zero_copy():
    """Fields in page-locked host memory (the reference's CPU arrays): kernels read/write them over PCIe."""
    igg.init_global_grid(24, 16, 12, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, 24, 16, 12, periodx=1, periody=1, periodz=1)
    a = O.random_bits_field((24, 16, 12), 8, 7)
    t = igg.zeros((24, 16, 12), dtype=torch.int64, pin_memory=True)
    t.copy_(igg.from_numpy(a, device="cpu"))
    want = a.copy(order="F")
    O.update_halo([g], [[want]])
    igg.update_halo(t)
    assert bits_equal(t, want)
    pageable = igg.zeros((24, 16, 12), dtype=torch.int64, device="cpu")
    with pytest.raises(igg.IGGError, match="pageable host memory"):
        igg.update_halo(pageable)


HOST_LINE_CASES = [
    # (grid, kwargs, shapes, elsize, dims): page-locked host fields whose rows start on 64-byte boundaries take the
    # whole-line x self-wrap (igg_wrap_lines_kernel); the others in the same call take the cell path.
    ((32, 16, 12), dict(periodx=1, periody=1, periodz=1), [(32, 16, 12)], 8, (3, 1, 2)),
    ((32, 16, 12), dict(periodx=1, periody=1, periodz=1), [(32, 16, 12), (33, 16, 12), (32, 17, 12)], 8, (3, 1, 2)),
    ((32, 16, 12), dict(periodx=1), [(32, 16, 12)], 8, (3, 1, 2)),                      # x only periodic: all rows wrap
    ((32, 16, 12), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), [(32, 16, 12), (32, 16, 13)], 4, (3, 1, 2)),
    ((16, 10, 9), dict(periodx=1, periody=1, periodz=1, overlaps=(6, 2, 2), halowidths=(3, 1, 1)), [(16, 10, 9)], 8, (3, 1, 2)),
    ((8, 10, 9), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(8, 10, 9)], 16, (1, 2)),
    ((128, 9, 7), dict(periodx=1, periodz=1), [(128, 9, 7)], 1, (3, 1, 2)),
    ((64, 40, 1), dict(periodx=1, periody=1), [(64, 40)], 2, (3, 1, 2)),
    ((256, 1, 1), dict(periodx=1), [(256,)], 4, (3, 1, 2)),
    ((32, 16, 300), dict(periodx=1, periody=1, periodz=1), [(32, 16, 300)], 8, (3, 1, 2)),   # many wrap tiles
]


@pytest.mark.parametrize("gsz,kw,shapes,elsize,dims", HOST_LINE_CASES, ids=[f"h{i}" for i in range(len(HOST_LINE_CASES))])
def test_pinned_host_fields_whole_line_wrap(gsz, kw, shapes, elsize, dims):
    igg.init_global_grid(*gsz, quiet=True, **kw)
    (g,) = O.make_grids(1, *gsz, **kw)
    tdt = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64, 16: torch.complex128}[elsize]
    for rep in range(2):                                    # twice: the second call reuses the cached plan
        base = [rand_any(s, elsize, 50 + 10 * rep + i) for i, s in enumerate(shapes)]
        want = [a.copy(order="F") for a in base]
        O.update_halo([g], [want], dims=dims)
        ts = []
        for a in base:
            t = igg.zeros(a.shape, dtype=tdt, pin_memory=True)
            t.copy_(igg.from_numpy(a, device="cpu"))
            ts.append(t)
        from igg_b200 import _lib as L
        n0 = L.load().igg_launch_count(igg.global_grid().ctx)
        igg.update_halo(*ts, dims=dims)
        if kw.get("periody") and len(gsz) == 3 and gsz[1] > 1:
            # exchange kernel + wrap-lines kernel: proves the whole-line path ran (the cell path is one launch)
            assert L.load().igg_launch_count(igg.global_grid().ctx) - n0 == 2
        for t, w in zip(ts, want):
            assert bits_equal(t, w)
    # the same bits as device-resident fields give
    td = [igg.from_numpy(a) for a in base]
    igg.update_halo(*td, dims=dims)
    for t, w in zip(td, want):
        assert bits_equal(t, w)


def test_async_stream_ordered_call():
    igg.init_global_grid(32, 32, 32, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, 32, 32, 32, periodx=1, periody=1, periodz=1)
    a = O.random_bits_field((32, 32, 32), 8, 11)
    want = a.copy(order="F")
    O.update_halo([g], [[want]])
    s = torch.cuda.Stream()
    t = igg.from_numpy(a)
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        for _ in range(5):
            igg.update_halo(t, stream=s, blocking=False)
    s.synchronize()
    assert bits_equal(t, want)


@pytest.mark.parametrize("elsize", [12, 32, 36])
def test_struct_elements_through_the_c_abi(elsize):
    """Elements of any isbits size (CellArray B=1 cells of 2x2 Float64 = 32 B; 12- and 36-byte structs): torch has no
    such dtype, so the field is a byte tensor handed to igg_update_halo directly."""
    import ctypes as C
    from igg_b200 import _lib as L
    n = (17, 12, 9)
    kw = dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1))
    igg.init_global_grid(*n, quiet=True, **kw)
    (g,) = O.make_grids(1, *n, **kw)
    gg = igg.global_grid()
    rng = np.random.default_rng(elsize)
    dt = np.dtype((np.void, elsize))
    shape = (n[0] + 1, n[1], n[2])
    ncell = int(np.prod(shape))
    a = np.frombuffer(rng.bytes(ncell * elsize), dtype=dt).reshape(shape, order="F").copy(order="F")
    want = a.copy(order="F")
    O.update_halo([g], [[(want, gg.halowidths)]], check=False)
    raw = torch.from_numpy(a.reshape(-1, order="F").view(np.uint8).copy()).cuda()
    f = L.make_field(raw.data_ptr(), shape, gg.halowidths, elsize)
    arr = (L.Field * 1)(f)
    L.check(L.load().igg_update_halo(gg.ctx, 1, arr, None, 0, C.c_void_p(torch.cuda.current_stream().cuda_stream), 0))
    got = raw.cpu().numpy()
    assert np.array_equal(got, want.reshape(-1, order="F").view(np.uint8))


def test_cuda_graph_capture_and_replay():
    """The exchange state (epoch, staging parity) lives on the device, so a stream-ordered update_halo can be
    captured into a CUDA graph and replayed."""
    igg.init_global_grid(32, 24, 20, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, 32, 24, 20, periodx=1, periody=1, periodz=1)
    a = O.random_bits_field((32, 24, 20), 8, 21)
    b = O.random_bits_field((33, 24, 20), 8, 22)
    want = [a.copy(order="F"), b.copy(order="F")]
    O.update_halo([g], [want])
    s = torch.cuda.Stream()
    ta, tb = igg.from_numpy(a), igg.from_numpy(b)
    call = igg.prepare_update_halo(ta, tb, stream=s, blocking=False)     # warms plan + caches outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=s):
        call()
        call()
    for rep in range(3):
        ta.copy_(igg.from_numpy(a)); tb.copy_(igg.from_numpy(b))         # fresh, un-updated halos
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        assert bits_equal(ta, want[0]) and bits_equal(tb, want[1])


def test_argument_errors_mirror_reference():
    """check_fields messages (src/update_halo.jl:418-480) through the public API."""
    igg.init_global_grid(nx, ny, nz, periodx=1, periody=1, periodz=1, quiet=True)
    P = igg.zeros((nx, ny, nz))
    with pytest.raises(igg.IGGError, match="The field at position 1 has a halowidth less than 1."):
        igg.update_halo((P, (0, 1, 1)))
    Q = igg.zeros((nx - 1, ny - 1, nz - 1))
    with pytest.raises(igg.IGGError, match="The field at position 2 has no halo; remove it from the call."):
        igg.update_halo(P, Q)
    with pytest.raises(igg.IGGError, match="The field at position 2 is a duplicate of the one at the position 1"):
        igg.update_halo(P, P)
    R = igg.zeros((nx, ny, nz), dtype=torch.float32)
    with pytest.raises((igg.IGGError, igg.ArgumentError), match="different type than the first field"):
        igg.update_halo(P, R)
    with pytest.raises(igg.ArgumentError, match="non-contiguous"):
        igg.update_halo(torch.zeros(nx, ny, nz, dtype=torch.float64, device="cuda"))
    igg.finalize_global_grid()
    with pytest.raises(igg.ArgumentError, match="No function of the module can be called before init_global_grid"):
        igg.update_halo(P)
    with pytest.raises(igg.ArgumentError):
        igg.init_global_grid(4, 4, 4, periodx=1, overlaps=(4, 2, 2))   # nx < 2*ol-1 with periodx


def test_gather_single_rank_and_errors():
    """test/test_gather.jl:19-69 with one rank."""
    igg.init_global_grid(nx, ny, nz, overlaps=(0, 0, 0), quiet=True)
    (g,) = O.make_grids(1, nx, ny, nz, overlaps=(0, 0, 0))
    for shp in [(nx,), (nx, ny), (nx, ny, nz)]:
        a = O.coord_field(g, shp)
        P = igg.from_numpy(a)
        P_g = igg.zeros(shp)
        igg.gather(P, P_g)
        assert bits_equal(P_g, a)
    C_ = igg.zeros((nx, ny, nz))
    with pytest.raises(igg.IGGError, match="must be equal to the product of `size\\(A\\)` and `dims`"):
        igg.gather(C_, igg.zeros((nx, ny, nz + 2)))
    with pytest.raises(igg.IGGError, match="can't be `nothing` on the root"):
        igg.gather(C_, None)


def test_launch_counter_and_kernel_name():
    import ctypes as C
    from igg_b200 import _lib as L
    igg.init_global_grid(16, 16, 16, periodx=1, quiet=True)
    t = igg.zeros((16, 16, 16))
    ctx = igg.global_grid().ctx
    n0 = L.load().igg_launch_count(C.c_void_p(ctx))
    for _ in range(4):
        igg.update_halo(t)
    assert L.load().igg_launch_count(C.c_void_p(ctx)) == n0 + 4


def test_prepared_calls_through_the_c_abi():
    """igg_prepare / igg_update_halo_prepared: same bits as igg_update_halo, repeatedly, blocking and stream-ordered,
    with `dims=` subsets, and a clean error once the context is gone."""
    import ctypes as C
    from igg_b200 import _lib as L
    lib = L.load()
    kw = dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1))
    igg.init_global_grid(33, 17, 20, quiet=True, **kw)
    (g,) = O.make_grids(1, 33, 17, 20, **kw)
    gg = igg.global_grid()
    shapes = [(33, 17, 20), (34, 17, 20)]
    base = [O.random_bits_field(s, 8, 40 + i) for i, s in enumerate(shapes)]
    s = torch.cuda.Stream()
    sp = C.c_void_p(s.cuda_stream)
    handles = []
    for dims in [(3, 1, 2), (1,), (2, 3)]:
        want = [a.copy(order="F") for a in base]
        O.update_halo([g], [want], dims=dims)
        ts = [igg.from_numpy(a) for a in base]
        arr = (L.Field * 2)(*[L.make_field(t.data_ptr(), tuple(t.shape), gg.halowidths, 8) for t in ts])
        d = (C.c_int * len(dims))(*dims)
        h = C.c_void_p()
        L.check(lib.igg_prepare(gg.ctx, 2, arr, d, len(dims), C.byref(h)))
        handles.append(h)
        for rep in range(4):
            for t, a in zip(ts, base):
                t.copy_(igg.from_numpy(a))
            torch.cuda.synchronize()
            L.check(lib.igg_update_halo_prepared(h, sp, L.IGG_SYNC if rep % 2 == 0 else L.IGG_ASYNC))
            if rep % 2 == 0:          # host-synchronous: the halos are there on return, without any further wait
                got = [t.clone() for t in ts]
            else:
                s.synchronize()
                got = ts
            for t, w in zip(got, want):
                assert bits_equal(t, w), (dims, rep)
    bad = (L.Field * 2)(L.make_field(ts[0].data_ptr(), shapes[0], gg.halowidths, 8), L.make_field(ts[0].data_ptr(), shapes[0], gg.halowidths, 8))
    h2 = C.c_void_p()
    assert lib.igg_prepare(gg.ctx, 2, bad, None, 0, C.byref(h2)) == L.IGG_ERR_FIELDS      # duplicates are still refused
    L.check(lib.igg_release_prepared(handles.pop()))
    igg.finalize_global_grid()
    assert lib.igg_update_halo_prepared(handles[0], sp, L.IGG_SYNC) == L.IGG_ERR_STATE
    for h in handles:
        L.check(lib.igg_release_prepared(h))


@pytest.mark.parametrize("blocking_env", ["spin", "sync"])
def test_blocking_call_completion_paths(blocking_env, monkeypatch):
    """The host-synchronous call waits either on a host-mapped word the kernel's last block writes (default) or with
    cudaStreamSynchronize (IGG_BLOCKING=sync): on return the halos must be in memory either way, also for plans
    with many more tiles than blocks and right after stream-ordered calls."""
    monkeypatch.setenv("IGG_BLOCKING", blocking_env)
    n = (96, 80, 72)
    igg.init_global_grid(*n, periodx=1, periody=1, periodz=1, quiet=True)
    (g,) = O.make_grids(1, *n, periodx=1, periody=1, periodz=1)
    base = [O.random_bits_field(n, 8, 60 + i) for i in range(6)]
    want = [a.copy(order="F") for a in base]
    O.update_halo([g], [want])
    ts = [igg.from_numpy(a) for a in base]
    side = torch.cuda.Stream()
    for rep in range(3):
        for t, a in zip(ts, base):
            t.copy_(igg.from_numpy(a))
        torch.cuda.synchronize()
        igg.update_halo(*ts[:3], blocking=False)
        igg.update_halo(*ts)                      # blocking; returns when its launch is done
        with torch.cuda.stream(side):             # a stream that is NOT ordered behind the call's stream
            got = [t.clone() for t in ts]
        side.synchronize()
        for t, w in zip(got, want):
            assert bits_equal(t, w), rep
    igg.finalize_global_grid()


@pytest.mark.parametrize("mode", ["fence", "auto"])
def test_big_self_wraps_bulk_copy_pipeline(mode, monkeypatch):
    """More than 8 MiB of self-wraps in one call: the LOCAL tiles grow beyond 8 KiB and, in the fence kernel, tiles of
    long 16-byte-aligned rows go through the cp.async.bulk pipeline (global -> shared -> global); rows that are not
    16-byte aligned (the field staggered in x: odd row length) and the strided x-faces stay on the load/store path.
    (x halo of two 8-byte cells: the y/z face boxes, which leave the x halo out, start on 16-byte boundaries.)  Same
    bits as the oracle either way, twice in a row, blocking and stream-ordered."""
    monkeypatch.setenv("IGG_MODE", mode)
    n = (384, 160, 96)
    kw = dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1))
    igg.init_global_grid(*n, quiet=True, **kw)
    (g,) = O.make_grids(1, *n, **kw)
    shapes = [n, n, (n[0] + 1, n[1], n[2]), n, (n[0], n[1] + 1, n[2]), n, n]
    base = [O.random_bits_field(s, 8, 300 + i) for i, s in enumerate(shapes)]
    want = [a.copy(order="F") for a in base]
    O.update_halo([g], [want])
    ts = [igg.from_numpy(a) for a in base]
    for rep in range(2):
        for t, a in zip(ts, base):
            t.copy_(igg.from_numpy(a))
        torch.cuda.synchronize()
        igg.update_halo(*ts, blocking=(rep == 0))
        torch.cuda.synchronize()
        for i, (t, w) in enumerate(zip(ts, want)):
            assert bits_equal(t, w), (mode, rep, i)
    if mode == "fence":
        from igg_b200 import _lib as L
        gg = igg.global_grid()
        tiles, _ = L.plan_tiles(gg.desc, [L.make_field(t.data_ptr(), tuple(t.shape), gg.halowidths, 8) for t in ts], (3, 1, 2), L.IGG_MODE_IPC_1PHASE)
        assert any(t.bulk for t in tiles) and any(not t.bulk for t in tiles)
    igg.finalize_global_grid()


def test_symmetric_field_allocation_single_rank():
    """igg_alloc_field / igg_free_field through the host mirror: zero-initialised column-major arrays that update_halo!
    treats like any other device array (with one rank there is no neighbour to store into), slots are reused."""
    kw = dict(periodx=1, periody=1, periodz=1)
    igg.init_global_grid(160, 12, 10, quiet=True, **kw)
    (g,) = O.make_grids(1, 160, 12, 10, **kw)
    a = O.random_bits_field((160, 12, 10), 8, 77)
    want = a.copy(order="F")
    O.update_halo([g], [[want]])
    t = igg.alloc_field((160, 12, 10), dtype=torch.int64)
    assert int(t.abs().sum()) == 0 and tuple(t.shape) == (160, 12, 10) and t.stride() == (1, 160, 1920)
    t.copy_(igg.from_numpy(a))
    u = igg.alloc_field((161, 12, 10), dtype=torch.int64)
    igg.update_halo(t)
    assert bits_equal(t, want)
    p_t = t.data_ptr()
    igg.free_field(t)
    v = igg.alloc_field((160, 12, 10), dtype=torch.int64)       # the freed slot is handed out again
    v.copy_(igg.from_numpy(a))
    igg.update_halo(v, u)
    assert bits_equal(v, want)
    with pytest.raises(KeyError):
        igg.free_field(torch.zeros(4, device="cuda"))
    del p_t
    igg.finalize_global_grid()                                  # releases u and v
