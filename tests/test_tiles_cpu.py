"""Tile tables of the exchange kernels, executed on the CPU (no GPU needed).

`igg_plan_tiles` returns the very records the CUDA kernels walk (struct TileRec).  Here every simulated rank's table
is executed with numpy exactly as the kernels address memory -- fence mode (PUSH / LOCAL / UNPACK tiles of rows and
columns) and LL128 mode (PUSH_LL / UNPACK_LL tiles over line-formatted staging: 120 payload bytes + 8-byte epoch per
128-byte line) -- and the fields must come out bit-identical to the oracle, and to the committed golden vectors.
This pins the host side of IGG_MODE_IPC_LL128 (slot layout, line numbering, field-side geometry, word width)."""
import numpy as np
import pytest

import golden_cases as G
from igg_b200 import _lib as L
from oracle import igg_oracle as O
from test_plan_vs_oracle import desc_of, flat_bytes, rand_field

PUSH, LOCAL, UNPACK, PUSH_LL, UNPACK_LL, WRAPX, PUSH_DIRECT, WAIT = 0, 1, 2, 6, 7, 8, 9, 10
WRAP_LINES, HOST_LINES_X = 5, 11
LINE, PAYLOAD = 128, 120
OFFSETS = [(ox, oy, oz) for oz in (-1, 0, 1) for oy in (-1, 0, 1) for ox in (-1, 0, 1)]   # index = (ox+1)+3(oy+1)+9(oz+1)


def peer_of(g, nbr):
    o = OFFSETS[nbr]
    c = []
    for d in range(3):
        v = g.coords[d] + o[d]
        if g.periods[d]:
            v %= g.dims[d]
        assert 0 <= v < g.dims[d]
        c.append(v)
    return (c[0] * g.dims[1] + c[1]) * g.dims[2] + c[2]


def copy_rows(t, src, dst):
    """copy_tile of the kernel: nrows rows from row0, ncols words of (1 << vec_log2) bytes each."""
    nb = t.ncols << t.vec_log2
    for r in range(t.nrows):
        row = t.row0 + r
        y, z = row % t.ny, row // t.ny
        s = t.src_off + y * t.src_sy + z * t.src_sz
        d = t.dst_off + y * t.dst_sy + z * t.dst_sz
        dst[d:d + nb] = src[s:s + nb]


def wrap_x(t, mem):
    """wrapx_tile of the kernel: per (row, word) the low send word goes to the high halo and the high send word to the
    low halo; high send = high halo - delta, low halo = low send - delta (delta in dst_sy)."""
    W, delta = 1 << t.vec_log2, t.dst_sy
    for r in range(t.nrows):
        row = t.row0 + r
        y, z = row % t.ny, row // t.ny
        for c in range(t.ncols):
            off = y * t.src_sy + z * t.src_sz + c * W
            lo_send, hi_halo = t.src_off + off, t.dst_off + off
            a = mem[lo_send:lo_send + W].copy()
            b = mem[hi_halo - delta:hi_halo - delta + W].copy()
            mem[hi_halo:hi_halo + W] = a
            mem[lo_send - delta:lo_send - delta + W] = b


def wrap_lines(t, mem):
    """igg_wrap_lines_kernel: periodic x self-wrap of host rows through their first and last 64-byte line."""
    hwb, olb, pitch = t.dst_sy, t.dst_sz, t.dst_off
    for r in range(t.nrows):
        row = t.row0 + r
        y, z = row % t.ny, row // t.ny
        a = t.src_off + y * t.src_sy + z * t.src_sz
        L = np.concatenate([mem[a:a + 64], mem[a + pitch - 64:a + pitch]])
        L[0:hwb], L[128 - hwb:128] = L[128 - olb:128 - olb + hwb].copy(), L[olb - hwb:olb].copy()
        mem[a:a + 64], mem[a + pitch - 64:a + pitch] = L[:64], L[64:]


def host_lines(t, mem, cache, write_back):
    """igg_host_lines_kernel: first / last 64-byte line of every row <-> 128 bytes per row of the device cache."""
    pitch, sides = t.dst_off, t.dst_sz
    for r in range(t.nrows):
        row = t.row0 + r
        y, z = row % t.ny, row // t.ny
        a, c = t.src_off + y * t.src_sy + z * t.src_sz, t.dst_sy + row * 128
        for side, (h, d) in enumerate(((a, c), (a + pitch - 64, c + 64))):
            if (sides >> side) & 1:
                if write_back:
                    mem[h:h + 64] = cache[d:d + 64]
                else:
                    cache[d:d + 64] = mem[h:h + 64]


def box_offset(t, pb):
    """box_seek of the kernel: payload byte pb -> byte offset in the field."""
    r, c = divmod(pb, t.ncols)
    z, y = divmod(r, t.ny)
    return t.src_off + y * t.src_sy + z * t.src_sz + c


def run_tables(grids, fields_per_rank, hws, dims, mode, epoch=7, direct=False, host=False):
    n = len(grids)
    mem = [[flat_bytes(A) for A in fr] for fr in fields_per_rank]
    tables, stagings, caches = [], [], []
    for r, g in enumerate(grids):
        cf = [L.make_field(0x10000 * (i + 1), A.shape, hws[i], A.dtype.itemsize) for i, A in enumerate(fields_per_rank[r])]
        tiles, stg = L.plan_tiles(desc_of(g), cf, dims, mode | (0x100 if direct else 0) | (0x200 if host else 0))
        tables.append([t for t in tiles if t.kind not in (WRAP_LINES, HOST_LINES_X)])
        stagings.append(np.full(max(stg, 1), 0xEE, dtype=np.uint8))
        # host fields: whole-line tiles.  x self-wraps run on their own; exchanged x-faces are fetched into a cache
        # that the exchange tiles address as extra fields (one per host field with such faces, in field order).
        hx = [t for t in tiles if t.kind == HOST_LINES_X]
        cache = np.full(max([t.dst_sy + (t.row0 + t.nrows) * 128 for t in hx], default=0), 0xCC, dtype=np.uint8)
        for t in tiles:
            if t.kind == WRAP_LINES:
                wrap_lines(t, mem[r][t.field])
        for t in hx:
            host_lines(t, mem[r][t.field], cache, False)
        seen = []
        for t in hx:
            if t.field not in seen:
                seen.append(t.field)
                mem[r].append(cache[t.dst_sy:])
        caches.append((hx, cache))
    kinds = {t.kind for tb in tables for t in tb}
    assert kinds <= ({PUSH_LL, LOCAL, WRAPX, UNPACK_LL} if mode == L.IGG_MODE_IPC_LL128 else {PUSH, LOCAL, WRAPX, UNPACK}) | ({PUSH_DIRECT, WAIT} if direct else set())
    ep = np.frombuffer(np.uint64(epoch).tobytes(), dtype=np.uint8)
    for r, g in enumerate(grids):                       # everything that does not wait
        order = [t.kind for t in tables[r]]
        assert order == sorted(order, key=lambda k: {PUSH: 0, PUSH_LL: 0, PUSH_DIRECT: 0, WRAPX: 1, LOCAL: 1, UNPACK: 2, UNPACK_LL: 2, WAIT: 2}[k])
        for t in tables[r]:
            if t.kind == PUSH:
                copy_rows(t, mem[r][t.field], stagings[peer_of(g, t.nbr)])
            elif t.kind == LOCAL:
                copy_rows(t, mem[r][t.field], mem[r][t.field])
            elif t.kind == WRAPX:
                wrap_x(t, mem[r][t.field])
            elif t.kind == PUSH_DIRECT:     # straight into the neighbour's copy of the field; it must expect that tile
                q = peer_of(g, t.nbr)
                copy_rows(t, mem[r][t.field], mem[q][t.field])
                waits = [w for w in tables[q] if w.kind == WAIT and w.nbr == 26 - t.nbr and w.flag_idx == t.flag_idx]
                assert len(waits) == 1 and waits[0].field == t.field
            elif t.kind == PUSH_LL:
                W, nbytes, stg = 1 << t.vec_log2, t.dst_sy, stagings[peer_of(g, t.nbr)]
                assert t.ncols % W == 0 and nbytes % W == 0 and t.dst_off % LINE == 0
                for line in range(t.row0, t.row0 + t.nrows):
                    base = t.dst_off + line * LINE
                    for pb in range(line * PAYLOAD, min(line * PAYLOAD + PAYLOAD, nbytes), W):
                        s = box_offset(t, pb)
                        stg[base + pb - line * PAYLOAD: base + pb - line * PAYLOAD + W] = mem[r][t.field][s:s + W]
                    stg[base + PAYLOAD: base + LINE] = ep
    for r, g in enumerate(grids):                       # everything that waits for a neighbour
        for t in tables[r]:
            if t.kind == UNPACK:
                copy_rows(t, stagings[r], mem[r][t.field])
            elif t.kind == UNPACK_LL:
                W, nbytes, stg = 1 << t.vec_log2, t.dst_sy, stagings[r]
                for line in range(t.row0, t.row0 + t.nrows):
                    base = t.dst_off + line * LINE
                    assert np.array_equal(stg[base + PAYLOAD: base + LINE], ep), "line was never pushed"
                    for pb in range(line * PAYLOAD, min(line * PAYLOAD + PAYLOAD, nbytes), W):
                        d = box_offset(t, pb)
                        mem[r][t.field][d:d + W] = stg[base + pb - line * PAYLOAD: base + pb - line * PAYLOAD + W]
    for r in range(n):                                  # host fields: the cached lines go back whole
        hx, cache = caches[r]
        for t in hx:
            host_lines(t, mem[r][t.field], cache, True)
    # every line of every LL slot that is read was written exactly once per call: count lines
    if mode == L.IGG_MODE_IPC_LL128:
        for r, g in enumerate(grids):
            pushed = sum(t.nrows for q, gq in enumerate(grids) for t in tables[q] if t.kind == PUSH_LL and peer_of(gq, t.nbr) == r)
            assert pushed == sum(t.nrows for t in tables[r] if t.kind == UNPACK_LL)


CASES = [
    # nprocs, grid, kwargs, shapes, per-field halowidths or None, elsize, dims
    (2, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 8, (3, 1, 2)),
    (4, (13, 9, 10), dict(), [(13, 9, 10), (14, 9, 10), (13, 10, 10), (13, 9, 11)], None, 8, (3, 1, 2)),
    (8, (13, 9, 10), dict(periodx=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
     [(13, 9, 10), (14, 9, 10), (12, 11, 11)], None, 4, (3, 1, 2)),
    (8, (13, 9, 10), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2)),
     [(13, 9, 10), (13, 9, 10), (14, 9, 10)], [(2, 2, 2), (1, 1, 2), (2, 1, 1)], 8, (3, 1, 2)),
    (8, (40, 30, 1), dict(periodx=1, periody=1), [(40, 30), (41, 30)], None, 8, (3, 1, 2)),
    (4, (64, 1, 1), dict(periodx=1), [(64,), (65,)], None, 4, (3, 1, 2)),
    (8, (40, 36, 34), dict(periodx=1, periody=1, periodz=1), [(40, 36, 34)], None, 8, (3, 1, 2)),      # several tiles per slot
    (2, (13, 9, 10), dict(periody=1, overlaps=(6, 4, 4), halowidths=(3, 1, 2)), [(13, 9, 10), (13, 9, 11)], None, 2, (1, 3)),
    (8, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 1, (2, 1)),
    (8, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10), (14, 9, 10)], None, 16, (3, 1, 2)),
    (4, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 12, (3, 1, 2)),       # W = 4 < element
    # the rank is its own x neighbour: the two x-faces of a field wrap together (WRAPX tiles)
    (1, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10), (14, 9, 10)], None, 8, (3, 1, 2)),
    (1, (16, 9, 10), dict(periodx=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(16, 9, 10), (17, 10, 10)], None, 8, (3, 1, 2)),
    (1, (40, 24, 20), dict(periodx=1, periody=1, periodz=1), [(40, 24, 20)], None, 4, (3, 1, 2)),      # several WRAPX tiles
    (1, (13, 9, 10), dict(periodx=1, overlaps=(6, 2, 2), halowidths=(3, 1, 1)), [(13, 9, 10)], None, 2, (1,)),
    (1, (13, 9, 10), dict(periodx=1, periody=1), [(13, 9, 10)], None, 1, (1, 2)),
    (1, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 12, (3, 1, 2)),
    (1, (13, 9, 10), dict(periodx=1, periody=1, periodz=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(13, 9, 10)], None, 16, (3, 1, 2)),
    (1, (40, 30, 1), dict(periodx=1, periody=1), [(40, 30), (41, 30)], None, 8, (3, 1, 2)),
    (1, (64, 1, 1), dict(periodx=1), [(64,), (65,)], None, 4, (3, 1, 2)),
    (2, (13, 9, 10), dict(dimx=1, dimy=2, periodx=1, periody=1, periodz=1), [(13, 9, 10), (14, 9, 10)], None, 8, (3, 1, 2)),
    (4, (13, 9, 10), dict(dimx=1, dimy=2, dimz=2, periodx=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(13, 9, 10)], None, 4, (3, 1, 2)),
    # more than 2 MiB staged per call: the fence protocol's tiles grow beyond 8 KiB
    (2, (256, 256, 6), dict(dimx=1, dimy=1, dimz=2, periodx=1, periody=1, periodz=1), [(256, 256, 6)] * 3, None, 8, (3, 1, 2)),
]


def test_fence_tiles_grow_with_the_call():
    grids = O.make_grids(2, 256, 256, 6, dimx=1, dimy=1, dimz=2, periodx=1, periody=1, periodz=1)
    cf = [L.make_field(0x10000 * (i + 1), (256, 256, 6), grids[0].halowidths, 8) for i in range(3)]
    tiles, _ = L.plan_tiles(desc_of(grids[0]), cf, (3, 1, 2), L.IGG_MODE_IPC_1PHASE)
    assert 8192 < max((t.nrows * t.ncols) << t.vec_log2 for t in tiles if t.kind == PUSH) <= 16384
    few, _ = L.plan_tiles(desc_of(grids[0]), cf[:1], (3, 1, 2), L.IGG_MODE_IPC_1PHASE)
    assert max((t.nrows * t.ncols) << t.vec_log2 for t in few if t.kind == PUSH) <= 8192
    # the default mode picks LL128 for this call (3 MiB staged, below the 64 MiB threshold)
    auto, _ = L.plan_tiles(desc_of(grids[0]), cf, (3, 1, 2), L.IGG_MODE_IPC_AUTO)
    assert any(t.kind == PUSH_LL for t in auto) and not any(t.kind == PUSH for t in auto)


def test_wrapx_tiles_are_emitted():
    """The single-rank periodic cases above must actually go through WRAPX tiles (not plain LOCAL copies)."""
    (g,) = O.make_grids(1, 13, 9, 10, periodx=1, periody=1, periodz=1)
    tiles, _ = L.plan_tiles(desc_of(g), [L.make_field(0x10000, (13, 9, 10), g.halowidths, 8)], (3, 1, 2), L.IGG_MODE_IPC_AUTO)
    assert sum(t.kind == WRAPX for t in tiles) >= 1 and all(t.kind in (WRAPX, LOCAL) for t in tiles)


def struct_field(shape, elsize, seed):
    rng = np.random.default_rng(seed)
    return np.asfortranarray(rng.integers(0, 256, size=tuple(shape) + (elsize,), dtype=np.uint8).view(f"V{elsize}")[..., 0])


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("case", CASES, ids=[f"c{i}" for i in range(len(CASES))])
def test_tile_tables_match_oracle(case, mode):
    nprocs, gsz, kw, shapes, hws, elsize, dims = case
    grids = O.make_grids(nprocs, *gsz, **kw)
    hws = hws or [grids[0].halowidths] * len(shapes)
    mk = struct_field if elsize == 12 else rand_field
    base = [[mk(s, elsize, 1000 * r + 10 * i + 3) for i, s in enumerate(shapes)] for r in range(nprocs)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, [[(a, hws[i]) for i, a in enumerate(fr)] for fr in want], dims=dims)
    got = [[a.copy(order="F") for a in fr] for fr in base]
    run_tables(grids, got, hws, dims, mode)
    for r in range(nprocs):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), (r, i)


def staged_range(t):
    """Byte range of a staging half a PUSH / UNPACK tile touches (packed layout of its plan)."""
    if t.kind in (PUSH_LL, UNPACK_LL):
        return t.dst_off + t.row0 * LINE, t.dst_off + (t.row0 + t.nrows) * LINE
    off, sy, sz = (t.dst_off, t.dst_sy, t.dst_sz) if t.kind == PUSH else (t.src_off, t.src_sy, t.src_sz)
    lo, hi = None, None
    for row in (t.row0, t.row0 + t.nrows - 1):
        y, z = row % t.ny, row // t.ny
        a = off + y * sy + z * sz
        lo = a if lo is None else min(lo, a)
        hi = max(hi or 0, a + (t.ncols << t.vec_log2))
    return lo, hi


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("case", CASES, ids=[f"c{i}" for i in range(len(CASES))])
def test_every_tile_stays_inside_its_offsets_region(case, mode):
    """At run time every neighbour offset owns a region of the staging half (and of the flag array) whose position
    does not depend on the plan; the launch shifts an offset's base by (region - packed start).  That is only sound
    if everything a tile touches lies inside the packed range of ITS offset: the receiver's offset (26 - nbr) for
    PUSH, its own (nbr) for UNPACK."""
    nprocs, gsz, kw, shapes, hws, elsize, dims = case
    grids = O.make_grids(nprocs, *gsz, **kw)
    hws = hws or [grids[0].halowidths] * len(shapes)
    for g in grids:
        cf = [L.make_field(0x10000 * (i + 1), s, hws[i], elsize) for i, s in enumerate(shapes)]
        tiles, stg, reg = L.plan_tiles(desc_of(g), cf, dims, mode, regions=True)
        assert sum(reg["byte_need"]) == stg and all(n % 128 == 0 for n in reg["byte_need"])
        for o in range(27):
            assert reg["byte_pack"][o] == sum(reg["byte_need"][:o]) and reg["flag_pack"][o] == sum(reg["flag_need"][:o])
        for t in tiles:
            if t.kind in (LOCAL, WRAPX, PUSH_DIRECT, WAIT):
                continue
            o = 26 - t.nbr if t.kind in (PUSH, PUSH_LL) else t.nbr
            lo, hi = staged_range(t)
            assert reg["byte_pack"][o] <= lo and hi <= reg["byte_pack"][o] + reg["byte_need"][o], (t.kind, t.nbr)
            if t.kind in (PUSH, UNPACK):
                assert reg["flag_pack"][o] <= t.flag_idx < reg["flag_pack"][o] + reg["flag_need"][o]
            else:
                assert sum(reg["flag_need"]) == 0


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("ci,case", list(G.cases()), ids=[c["name"] for _, c in G.cases()])
def test_tile_tables_match_golden(ci, case, mode):
    grids = O.make_grids(case["nprocs"], *case["grid"], **G.kwargs(case))      # grid descriptions only
    hws = G.halowidths(case, grids[0].halowidths)
    fields = [G.arrays(ci, r, "in") for r in range(case["nprocs"])]
    run_tables(grids, fields, hws, tuple(case["dims"]), mode)
    for r in range(case["nprocs"]):
        for a, w in zip(fields[r], G.arrays(ci, r, "out")):
            assert np.array_equal(a, w), (case["name"], r)


DIRECT_CASES = [
    (2, (160, 9, 10), dict(periodx=1, periody=1, periodz=1), [(160, 9, 10), (161, 9, 10)], None, 8, (3, 1, 2)),
    (4, (160, 9, 10), dict(), [(160, 9, 10), (160, 10, 10), (160, 9, 11)], None, 8, (3, 1, 2)),
    (8, (130, 9, 10), dict(periodx=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)), [(130, 9, 10), (131, 9, 10)], None, 8, (3, 1, 2)),
    (8, (300, 12, 10), dict(periodx=1, periody=1, periodz=1), [(300, 12, 10)], None, 4, (3, 1, 2)),
    (4, (2100, 40, 1), dict(periodx=1, periody=1, dimx=1, dimy=4), [(2100, 40)], None, 8, (3, 1, 2)),     # rows longer than a tile
    (2, (160, 9, 10), dict(dimx=1, dimy=1, dimz=2, periodx=1, periody=1, periodz=1), [(160, 9, 10)], None, 8, (3,)),
    (8, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 8, (3, 1, 2)),           # too narrow: all staged
]


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("case", DIRECT_CASES, ids=[f"d{i}" for i in range(len(DIRECT_CASES))])
def test_direct_to_halo_tile_tables_match_oracle(case, mode):
    """Fields in symmetric allocations: boxes that span x (rows of at least 1 KiB) are stored straight into the
    neighbour's field (PUSH_DIRECT, paired with exactly one WAIT tile there), everything else stays staged."""
    nprocs, gsz, kw, shapes, hws, elsize, dims = case
    grids = O.make_grids(nprocs, *gsz, **kw)
    hws = hws or [grids[0].halowidths] * len(shapes)
    base = [[rand_field(s, elsize, 2000 * r + 10 * i + 5) for i, s in enumerate(shapes)] for r in range(nprocs)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, [[(a, hws[i]) for i, a in enumerate(fr)] for fr in want], dims=dims)
    got = [[a.copy(order="F") for a in fr] for fr in base]
    run_tables(grids, got, hws, dims, mode, direct=True)
    for r in range(nprocs):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), (r, i)
    cf = [L.make_field(0x10000 * (i + 1), s, hws[i], elsize) for i, s in enumerate(shapes)]
    tiles, _ = L.plan_tiles(desc_of(grids[0]), cf, dims, mode | 0x100)
    wide = shapes[0][0] * elsize >= 1024 and (grids[0].dims[1] > 1 or grids[0].dims[2] > 1)     # a y or z neighbour is another rank
    assert any(t.kind == PUSH_DIRECT for t in tiles) == wide


HOST_CASES = [
    # rows on 64-byte boundaries: x-faces through whole lines (wrapped: kind 5; exchanged: kind 11 + cache fields)
    (2, (16, 9, 10), dict(periodx=1, periody=1, periodz=1), [(16, 9, 10), (24, 9, 10)], None, 8, (3, 1, 2)),
    (2, (32, 9, 10), dict(overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(32, 9, 10), (32, 10, 10), (33, 9, 10)], None, 4, (3, 1, 2)),
    (8, (16, 9, 10), dict(periodx=1, periodz=1), [(16, 9, 10)], None, 8, (3, 1, 2)),
    (4, (64, 12, 1), dict(periodx=1, periody=1, overlaps=(4, 2, 2), halowidths=(2, 1, 1)), [(64, 12), (64, 13)], None, 2, (3, 1, 2)),
    (1, (16, 9, 10), dict(periodx=1, periody=1, periodz=1), [(16, 9, 10)], None, 8, (3, 1, 2)),
    (2, (16, 9, 10), dict(dimx=1, dimy=2, periodx=1, periody=1, periodz=1), [(16, 9, 10)], None, 8, (3, 1, 2)),
    (3, (16, 9, 10), dict(), [(16, 9, 10)], None, 8, (1,)),                     # non-periodic: edge ranks exchange one end only
    (2, (13, 9, 10), dict(periodx=1, periody=1, periodz=1), [(13, 9, 10)], None, 8, (3, 1, 2)),     # rows not aligned: per cell
]


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("case", HOST_CASES, ids=[f"h{i}" for i in range(len(HOST_CASES))])
def test_host_field_tile_tables_match_oracle(case, mode):
    """Fields in page-locked host memory: x self-wraps and exchanged x-faces through whole 64-byte lines."""
    nprocs, gsz, kw, shapes, hws, elsize, dims = case
    grids = O.make_grids(nprocs, *gsz, **kw)
    hws = hws or [grids[0].halowidths] * len(shapes)
    base = [[rand_field(s, elsize, 3000 * r + 10 * i + 7) for i, s in enumerate(shapes)] for r in range(nprocs)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, [[(a, hws[i]) for i, a in enumerate(fr)] for fr in want], dims=dims)
    got = [[a.copy(order="F") for a in fr] for fr in base]
    run_tables(grids, got, hws, dims, mode, host=True)
    for r in range(nprocs):
        for i in range(len(shapes)):
            assert np.array_equal(flat_bytes(got[r][i]), flat_bytes(want[r][i])), (r, i)
    cf = [L.make_field(0x10000 * (i + 1), s, hws[i], elsize) for i, s in enumerate(shapes)]
    tiles, _ = L.plan_tiles(desc_of(grids[0]), cf, dims, mode | 0x200)
    aligned = shapes[0][0] * elsize % 64 == 0
    x_remote, x_self = grids[0].dims[0] > 1, grids[0].dims[0] == 1 and grids[0].periods[0]
    assert any(t.kind == HOST_LINES_X for t in tiles) == (aligned and x_remote and 1 in dims)
    assert any(t.kind == WRAP_LINES for t in tiles) == (aligned and bool(x_self) and 1 in dims)
