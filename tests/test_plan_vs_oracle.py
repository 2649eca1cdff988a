"""Host logic of libigg_b200 vs the oracle, without a GPU.

`igg_plan_describe` returns the exact copy operations `igg_update_halo` would launch on a rank (one phase, up to
26 neighbours, byte offsets / strides / staging slots as the kernel uses them).  Here every simulated rank's plan is
executed with numpy on flat byte buffers -- PUSH into the peer's staging, LOCAL in place, then UNPACK -- and the
result must be bit-identical to the oracle's three sequential per-dimension phases (ref: src/update_halo.jl:51-88).
This is what ties the single-phase formulation to the reference semantics.
"""
import itertools

import numpy as np
import pytest

from igg_b200 import _lib as L
from oracle import igg_oracle as O

DT = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64, 16: np.complex128}


def rand_field(shape, elsize, seed):
    if elsize == 16:
        n = int(np.prod(shape))
        return O.random_bits_field((2 * n,), 8, seed).view(np.complex128).reshape(shape, order="F")
    return O.random_bits_field(shape, elsize, seed)


def desc_of(g, device=0):
    return L.make_desc(g.nxyz, g.dims, g.coords, g.neighbors, g.periods, g.overlaps, g.halowidths, g.me, g.nprocs,
                       device)


def flat_bytes(A):
    """Memory-order byte view of an F-contiguous array (no copy)."""
    assert A.flags.f_contiguous
    return A.reshape(-1, order="F").view(np.uint8) if A.ndim > 1 else A.view(np.uint8)


def run_lowlevel(op, src, dst):
    rb = op.row_bytes
    for z in range(op.nz):
        for y in range(op.ny):
            s = op.src_off + y * op.src_sy + z * op.src_sz
            d = op.dst_off + y * op.dst_sy + z * op.dst_sz
            dst[d:d + rb] = src[s:s + rb]


def execute_plans(grids, fields_per_rank, hws_per_field, dims=None, align=0):
    """fields_per_rank[r][i]: numpy F-order arrays (modified in place)."""
    n = len(grids)
    plans, stagings = [], []
    for r, g in enumerate(grids):
        cf = [L.make_field(0x10000 * (i + 1) + align, A.shape, hws_per_field[i], A.dtype.itemsize)
              for i, A in enumerate(fields_per_rank[r])]
        ops, stg = L.plan_describe(desc_of(g), cf, dims)
        plans.append(ops)
        stagings.append(np.full(max(stg, 1), 0xEE, dtype=np.uint8))
    mem = [[flat_bytes(A) for A in fr] for fr in fields_per_rank]
    # phase A: PUSH (into the peer's staging) and LOCAL (in place), in rank order -- order must not matter
    for r in range(n):
        for op in plans[r]:
            if op.kind == L.IGG_OP_PUSH:
                assert op.peer != grids[r].me
                run_lowlevel(op, mem[r][op.field], stagings[op.peer])
            elif op.kind == L.IGG_OP_LOCAL:
                assert op.peer == grids[r].me
                run_lowlevel(op, mem[r][op.field], mem[r][op.field])
    # phase B: UNPACK
    for r in range(n):
        for op in plans[r]:
            if op.kind == L.IGG_OP_UNPACK:
                run_lowlevel(op, stagings[r], mem[r][op.field])
    return plans


def compare(grids, shapes, hws, elsize, dims=None, align=0, seed=0):
    n = len(grids)
    base = [[rand_field(s, elsize, seed + 1000 * r + 10 * i) for i, s in enumerate(shapes)] for r in range(n)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    got = [[a.copy(order="F") for a in fr] for fr in base]
    fields_o = [[(a, hws[i]) for i, a in enumerate(fr)] for fr in want]
    O.update_halo(grids, fields_o, dims=dims if dims is not None else (3, 1, 2))
    plans = execute_plans(grids, got, hws, dims, align)
    for r in range(n):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), f"rank {r} field {i}"
    return plans


TOPOLOGIES = [(1, 1, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2), (3, 2, 1), (1, 1, 2), (1, 3, 1)]
PERIODS = list(itertools.product((0, 1), repeat=3))


@pytest.mark.parametrize("topo", TOPOLOGIES, ids=lambda t: "x".join(map(str, t)))
@pytest.mark.parametrize("periods", PERIODS, ids=lambda p: "p" + "".join(map(str, p)))
def test_default_overlap_all_topologies(topo, periods):
    nx, ny, nz = 7, 5, 6
    grids = O.make_grids(int(np.prod(topo)), nx, ny, nz, dimx=topo[0], dimy=topo[1], dimz=topo[2],
                         periodx=periods[0], periody=periods[1], periodz=periods[2])
    hw = grids[0].halowidths
    shapes = [(nx, ny, nz), (nx + 1, ny, nz), (nx, ny, nz + 1)]
    compare(grids, shapes, [hw] * 3, 8, seed=7)


@pytest.mark.parametrize("topo", [(1, 1, 1), (2, 2, 2), (3, 2, 1)], ids=lambda t: "x".join(map(str, t)))
@pytest.mark.parametrize("periods", [(1, 1, 1), (0, 1, 0), (0, 0, 0), (1, 0, 1)], ids=lambda p: "p" + "".join(map(str, p)))
@pytest.mark.parametrize("olhw", [((4, 2, 3), (2, 1, 1)), ((6, 4, 4), (3, 1, 2)), ((4, 4, 4), (2, 2, 2)), ((3, 3, 3), (1, 1, 1))],
                         ids=["ol423", "ol644", "ol444", "ol333"])
def test_wide_halos_and_staggering(topo, periods, olhw):
    ol, hw = olhw
    nx, ny, nz = 13, 9, 10
    grids = O.make_grids(int(np.prod(topo)), nx, ny, nz, dimx=topo[0], dimy=topo[1], dimz=topo[2],
                         periodx=periods[0], periody=periods[1], periodz=periods[2], overlaps=ol, halowidths=hw)
    # P, staggered Vx/Vz, and A (nx-1, ny+2, nz+1) which has no halo in x for ol_x - 1 < 2 hw_x
    shapes = [(nx, ny, nz), (nx + 1, ny, nz), (nx, ny, nz + 1), (nx - 1, ny + 2, nz + 1)]
    compare(grids, shapes, [hw] * 4, 4, seed=11)


@pytest.mark.parametrize("elsize", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("align", [0, 8, 4, 2, 1])
def test_element_sizes_and_alignment(elsize, align):
    """ComplexF16 (4 B), Int16, Float32/64, 16-byte elements; vector width must follow pointer alignment."""
    if align % elsize and elsize < 16:
        pytest.skip("an array of T is at least T-aligned")
    if elsize == 16 and align % 8:
        pytest.skip("complex128 arrays are 8-byte aligned")
    grids = O.make_grids(4, 8, 6, 5, dimx=2, dimy=2, dimz=1, periodx=1, periody=1, periodz=1)
    hw = grids[0].halowidths
    plans = compare(grids, [(8, 6, 5), (9, 6, 5)], [hw, hw], elsize, align=align, seed=3)
    for ops in plans:
        for op in ops:
            assert op.vec_bytes in (1, 2, 4, 8, 16)
            if op.kind != L.IGG_OP_UNPACK:
                assert (0x10000 * (op.field + 1) + align + op.src_off) % op.vec_bytes == 0
            if op.kind != L.IGG_OP_PUSH:
                assert (0x10000 * (op.field + 1) + align + op.dst_off) % op.vec_bytes == 0
            assert op.row_bytes % op.vec_bytes == 0


@pytest.mark.parametrize("dims", [(1,), (2,), (3,), (1, 2), (3, 1), (2, 3), (1, 2, 3), (2, 1, 3)])
def test_dims_subsets(dims):
    grids = O.make_grids(8, 7, 6, 5, periodx=1, periodz=1)
    hw = grids[0].halowidths
    compare(grids, [(7, 6, 5), (7, 7, 5)], [hw, hw], 8, dims=dims, seed=5)


def test_per_field_halowidths():
    """update_halo!(A, (A=C, halowidths=(2,2,2))) -- ref: src/update_halo.jl:16."""
    grids = O.make_grids(4, 12, 10, 9, dimx=2, dimy=1, dimz=2, periodx=1, periody=1, overlaps=(4, 4, 4),
                         halowidths=(2, 2, 2))
    compare(grids, [(12, 10, 9), (12, 10, 9), (13, 10, 9)], [(2, 2, 2), (1, 1, 2), (2, 1, 1)], 8, seed=9)


@pytest.mark.parametrize("shape,gargs,kw", [
    ((7,), (7, 1, 1), dict(periodx=1)),
    ((8,), (7, 1, 1), dict(periodx=1)),
    ((7, 5), (7, 5, 1), dict(periodx=1, periody=1)),
    ((7, 6), (7, 5, 1), dict(periody=1)),
    ((6, 7), (7, 5, 1), dict(periodx=1, periody=1)),       # no halo in x
])
@pytest.mark.parametrize("nprocs", [1, 2, 3])
def test_1d_2d_arrays(shape, gargs, kw, nprocs):
    grids = O.make_grids(nprocs, *gargs, **kw)
    hw = grids[0].halowidths
    compare(grids, [shape], [hw], 8, seed=13)


def test_reference_sizes_256_plan_shape():
    """cfg2: one 256^3 Float64 field, periodic, 2x2x2: 26 pushes + 26 unpacks, 6 faces of 254x254 cells etc."""
    grids = O.make_grids(8, 256, 256, 256, periodx=1, periody=1, periodz=1)
    f = [L.make_field(0x7f0000000000, (256, 256, 256), (1, 1, 1), 8)]
    ops, stg = L.plan_describe(desc_of(grids[0]), f)
    kinds = [op.kind for op in ops]
    assert kinds.count(L.IGG_OP_PUSH) == 26 and kinds.count(L.IGG_OP_UNPACK) == 26 and kinds.count(L.IGG_OP_LOCAL) == 0
    sent = sum(op.nbytes for op in ops if op.kind == L.IGG_OP_PUSH)
    # faces 6*254^2 + edges 12*254 + corners 8 cells == boundary shell of the 256^3 block minus ... == 256^3 - 254^3 cells
    assert sent == (256 ** 3 - 254 ** 3) * 8
    assert stg >= sent
    # z-faces are contiguous 254-cell rows with stride 256 (interior only) -> 16-byte vectors impossible (254*8 % 16 == 0 but offset 8)
    one = O.make_grids(1, 256, 256, 256, periodx=1, periody=1, periodz=1)
    ops1, stg1 = L.plan_describe(desc_of(one[0]), f)
    assert all(op.kind == L.IGG_OP_LOCAL for op in ops1) and len(ops1) == 26 and stg1 == 0 or stg1 >= 0


def test_error_paths():
    (g,) = O.make_grids(1, 7, 5, 6, periodx=1, periody=1, periodz=1)
    d = desc_of(g)
    f = L.make_field(0x1000, (6, 4, 5), (1, 1, 1), 8)
    with pytest.raises(L.IGGError, match="has no halo; remove it from the call"):
        L.plan_describe(d, [f])
    f0 = L.make_field(0x1000, (7, 5, 6), (0, 1, 1), 8)
    with pytest.raises(L.IGGError, match="has a halowidth less than 1"):
        L.plan_describe(d, [f0])
    fa, fb = L.make_field(0x1000, (7, 5, 6), (1, 1, 1), 8), L.make_field(0x1000, (7, 5, 6), (1, 1, 1), 8)
    with pytest.raises(L.IGGError, match="is a duplicate of the one at the position 1"):
        L.plan_describe(d, [fa, fb])
    fc = L.make_field(0x2000, (7, 5, 6), (1, 1, 1), 4)
    with pytest.raises(L.IGGError, match="different type than the first field"):
        L.plan_describe(d, [fa, fc])
    bad = desc_of(g)
    bad.neighbors[0][1] = L.IGG_PROC_NULL          # left = me, right = PROC_NULL
    with pytest.raises(L.IGGError, match="Incoherent neighbors in dimension 1"):
        L.plan_describe(bad, [fa])


THIN = [((4, 8, 8), (4, 2, 2), (2, 1, 1), [(4, 8, 8)]), ((4, 4, 8), (4, 4, 2), (2, 2, 1), [(4, 4, 8), (4, 5, 8)])]


@pytest.mark.parametrize("topo", [(2, 1, 1), (3, 1, 1), (3, 2, 1), (2, 2, 2), (3, 3, 1)], ids=lambda t: "x".join(map(str, t)))
@pytest.mark.parametrize("periods", [(0, 0, 0), (0, 0, 1)], ids=lambda p: "p" + "".join(map(str, p)))   # thin dims cannot be periodic
@pytest.mark.parametrize("thin", THIN, ids=["x", "xy"])
def test_arrays_without_interior_cells(topo, periods, thin):
    """Arrays made of halo cells only in an exchanged dimension (size == 2 * halowidth): the reference accepts them
    (has-halo predicate only, ref: src/update_halo.jl:239, src/shared.jl:220) and so does the single-phase plan -- a
    rank in the middle of such a dimension has empty face boxes in the other dimensions, its neighbours' data all
    arrives through the faces / edges of that dimension."""
    gsz, ol, hw, shapes = thin
    grids = O.make_grids(topo[0] * topo[1] * topo[2], *gsz, dimx=topo[0], dimy=topo[1], dimz=topo[2], periodx=periods[0],
                         periody=periods[1], periodz=periods[2], overlaps=ol, halowidths=hw)
    compare(grids, shapes, [hw] * len(shapes), 8)


def test_c_ranges_match_reference_known_answers():
    """igg_sendranges / igg_recvranges against test/test_update_halo.jl:404-437."""
    import ctypes as C
    lib = L.load()
    (g,) = O.make_grids(1, 13, 9, 9, periodx=1, periody=1, periodz=1, overlaps=(6, 4, 4), halowidths=(3, 1, 2))
    d = desc_of(g)
    out = (C.c_int64 * 6)()
    for shape in [(13, 9, 9), (12, 11, 10)]:
        f = L.make_field(0x1000, shape, g.halowidths, 8)
        for dim in (1, 2, 3):
            for n in (1, 2):
                for send, cfn, ofn in ((True, lib.igg_sendranges, O.sendranges), (False, lib.igg_recvranges, O.recvranges)):
                    rc = cfn(C.byref(d), C.byref(f), n, dim, out)
                    try:
                        want = ofn(g, n, dim, shape, g.halowidths)
                    except O.IGGError:
                        assert rc == L.IGG_ERR_FIELDS
                        assert b"Incoherent arguments: ol(A,dim)<2*halowidths[dim]." in lib.igg_last_error()
                        continue
                    assert rc == 0
                    assert [tuple(out[2 * k:2 * k + 2]) for k in range(3)] == want


@pytest.mark.parametrize("topo,periods", [((2, 1, 1), (1, 1, 1)), ((2, 2, 1), (1, 1, 1)), ((2, 2, 2), (1, 1, 1)),
                                          ((2, 2, 2), (0, 0, 0)), ((3, 2, 1), (1, 0, 1))])
def test_arrival_flags_pair_up_and_never_collide(topo, periods):
    """Every PUSH tile sets one arrival flag on the receiver and the matching UNPACK tile waits for exactly that
    flag: sender and receiver must derive the same flag numbers and tile counts, and no two tiles arriving at one
    rank may share a flag (a shared flag would let an unpack start before its data has landed)."""
    nx, ny, nz = 40, 36, 33
    grids = O.make_grids(int(np.prod(topo)), nx, ny, nz, dimx=topo[0], dimy=topo[1], dimz=topo[2],
                         periodx=periods[0], periody=periods[1], periodz=periods[2])
    hw = grids[0].halowidths
    shapes = [(nx, ny, nz), (nx + 1, ny, nz)]
    plans = []
    for g in grids:
        cf = [L.make_field(0x10000 * (i + 1), s, hw, 8) for i, s in enumerate(shapes)]
        plans.append(L.plan_describe(desc_of(g), cf)[0])
    for r, ops in enumerate(plans):
        used = set()
        for op in ops:
            if op.kind != L.IGG_OP_UNPACK:
                continue
            rng = set(range(op.flag_base, op.flag_base + op.ntiles))
            assert not (rng & used), "two arriving tiles share a flag"
            used |= rng
            # the sender's matching PUSH op: same field, opposite offset, aimed at me
            o = tuple(op.offset)
            match = [p for p in plans[op.peer] if p.kind == L.IGG_OP_PUSH and p.peer == r and p.field == op.field
                     and tuple(p.offset) == tuple(-x for x in o)]
            assert len(match) == 1
            assert match[0].flag_base == op.flag_base and match[0].ntiles == op.ntiles
            assert match[0].staging_off == op.staging_off and match[0].nbytes == op.nbytes


@pytest.mark.parametrize("elsize", [3, 12, 24, 32, 36])
def test_struct_elements_cellarray_b1(elsize):
    """Any isbits element size: a CellArray with B=1 is an array of cells of celllength*sizeof(T) bytes
    (ref: src/shared.jl:288, 2x2 Float64 cells = 32 B in test/test_update_halo.jl:17-18); 12- and 36-byte structs too."""
    grids = O.make_grids(4, 9, 7, 6, dimx=2, dimy=1, dimz=2, periodx=1, periody=1, periodz=1)
    hw = grids[0].halowidths
    shapes = [(9, 7, 6), (9, 8, 6)]
    rng = np.random.default_rng(elsize)
    dt = np.dtype((np.void, elsize))

    def rnd(shape):
        n = int(np.prod(shape))
        return np.frombuffer(rng.bytes(n * elsize), dtype=dt).reshape(shape, order="F").copy(order="F")

    base = [[rnd(s) for s in shapes] for _ in grids]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    got = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, [[(a, hw) for a in fr] for fr in want], check=False)
    execute_plans(grids, got, [hw, hw])
    for r in range(len(grids)):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), (r, i)
