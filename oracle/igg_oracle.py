"""CPU oracle for the halo-update path of ImplicitGlobalGrid.jl v0.17.1  --  TEST INFRASTRUCTURE ONLY.

This file is a numpy restatement of the reference's algorithm for `update_halo!` / `gather!` and the
host-side index helpers around them.  It simulates N MPI ranks inside one Python process.  It is the
checker for the CUDA path: only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` /
`--impl reference` legs may import it.  Nothing under `implicitglobalgrid.jl_b200/` imports it.

Parity pinning: the reference is Julia + MPI.jl and cannot be executed in this environment (no Julia, no MPI),
so the oracle is pinned by the reference's own *closed-form* test fixtures (known-answer ranges, buffer
contents, coordinate-coded full updates, `x_g`/`ix_g`/`nx_g` vectors), ported in `tests/test_oracle_*.py`.
Multi-rank message exchange is pinned by the model only (the reference tests it only under manual `mpiexec`).

Conventions: a Julia array of size (nx,ny,nz), column-major, is a numpy array of shape (nx,ny,nz) with
order='F'.  All index ranges below are 1-based inclusive like in the reference; helpers convert to slices.
`n` = 1 (left/low side) or 2 (right/high side); `dim` = 1..3.

All file:line citations are relative to /root/reference/.
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

NDIMS_MPI = 3            # src/shared.jl:29
NNEIGHBORS_PER_DIM = 2   # src/shared.jl:30
PROC_NULL = -1           # MPI.PROC_NULL stand-in (any negative value)


class IGGError(RuntimeError):
    """Stand-in for Julia's ErrorException raised by `error(...)` in the reference."""


# --------------------------------------------------------------------------------------------------
# MPI semantics that the reference relies on (MPI.jl 0.20, not vendored in /root/reference)
# --------------------------------------------------------------------------------------------------

def dims_create(nprocs: int, dims: Sequence[int]) -> Tuple[int, ...]:
    """MPI_Dims_create: fill the zero entries of `dims` so that prod(dims)==nprocs, as balanced as possible,
    free entries in non-increasing order.  Call site: src/create_global_grid.jl:94."""
    dims = list(int(d) for d in dims)
    fixed = 1
    for d in dims:
        if d < 0:
            raise IGGError("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
        if d > 0:
            fixed *= d
    free = [i for i, d in enumerate(dims) if d == 0]
    if nprocs % fixed != 0:
        raise IGGError("MPI_Dims_create: nprocs is not a multiple of the product of the fixed dims")
    rest = nprocs // fixed
    if not free:
        if rest != 1:
            raise IGGError("MPI_Dims_create: product of dims does not match nprocs")
        return tuple(dims)

    def factorizations(n, k, maxf):
        # non-increasing k-tuples of positive ints with product n and first factor <= maxf
        if k == 1:
            if n <= maxf:
                yield (n,)
            return
        for f in range(min(n, maxf), 0, -1):
            if n % f == 0:
                for tail in factorizations(n // f, k - 1, f):
                    yield (f,) + tail

    best = min(factorizations(rest, len(free), rest))  # lexicographically smallest == most balanced
    for i, f in zip(free, best):
        dims[i] = f
    return tuple(dims)


def cart_rank(coords: Sequence[int], dims: Sequence[int]) -> int:
    """MPI Cartesian communicators are row-major (last coordinate fastest); confirmed by src/gather.jl:40-46."""
    r = 0
    for c, d in zip(coords, dims):
        r = r * d + c
    return r


def cart_coords(rank: int, dims: Sequence[int]) -> Tuple[int, ...]:
    coords = []
    for d in reversed(dims):
        coords.append(rank % d)
        rank //= d
    return tuple(reversed(coords))


def cart_shift(coords, dims, periods, dim0: int, disp: int) -> Tuple[int, int]:
    """MPI_Cart_shift(direction=dim0, disp) -> (source, dest) = (left, right) neighbour ranks
    (src/create_global_grid.jl:98-100)."""
    out = []
    for s in (-disp, +disp):
        c = list(coords)
        c[dim0] += s
        if periods[dim0]:
            c[dim0] %= dims[dim0]
            out.append(cart_rank(c, dims))
        elif 0 <= c[dim0] < dims[dim0]:
            out.append(cart_rank(c, dims))
        else:
            out.append(PROC_NULL)
    return out[0], out[1]


# --------------------------------------------------------------------------------------------------
# Grid state (src/shared.jl:58-81, src/create_global_grid.jl:35-111, src/shared.jl:156-207)
# --------------------------------------------------------------------------------------------------

@dataclass
class GlobalGrid:
    nxyz_g: Tuple[int, int, int]
    nxyz: Tuple[int, int, int]
    dims: Tuple[int, int, int]
    overlaps: Tuple[int, int, int]
    halowidths: Tuple[int, int, int]
    origin: Tuple[float, float, float]
    origin_on_vertex: bool
    centerxyz: Tuple[bool, bool, bool]
    nprocs: int
    me: int
    coords: Tuple[int, int, int]
    neighbors: Tuple[Tuple[int, int, int], Tuple[int, int, int]]  # neighbors[n-1][dim-1]
    periods: Tuple[int, int, int]
    disp: int = 1

    def neighbor(self, n: int, dim: int) -> int:          # src/shared.jl:224
        return self.neighbors[n - 1][dim - 1]

    def has_neighbor(self, n: int, dim: int) -> bool:     # src/shared.jl:233
        return self.neighbor(n, dim) != PROC_NULL


def normalize_input(nx, ny, nz, dimx=0, dimy=0, dimz=0, periodx=0, periody=0, periodz=0,
                    overlaps=(2, 2, 2), halowidths=None, origin=(0.0, 0.0, 0.0), origin_on_vertex=False,
                    centerx=False, centery=False, centerz=False):
    """Value checks of src/shared.jl:156-207 (only those that concern the halo path)."""
    if halowidths is None:
        halowidths = tuple(max(1, o // 2) for o in overlaps)          # src/init_global_grid.jl:83
    dims = [int(dimx), int(dimy), int(dimz)]
    periods = [int(periodx), int(periody), int(periodz)]
    nxyz = [int(nx), int(ny), int(nz)]
    if any(d < 0 for d in dims):
        raise IGGError("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
    if any(p not in (0, 1) for p in periods):
        raise IGGError("Invalid arguments: periodx, periody, and periodz must be either 0 or 1.")
    if any(h < 1 for h in halowidths):
        raise IGGError("Invalid arguments: halowidths cannot be less than 1.")
    if any(o > 0 and h > o // 2 for o, h in zip(overlaps, halowidths)):
        raise IGGError("Incoherent arguments: if overlap is greater than 0, then halowidth cannot be greater "
                       "than overlap÷2, in each dimension.")
    if any(n < 1 for n in nxyz):
        raise IGGError("Invalid arguments: nx, ny, and nz cannot be less than 1.")
    for c, n, name in zip((centerx, centery, centerz), nxyz, "xyz"):
        if c and origin_on_vertex and n % 2 == 1:
            raise IGGError(f"Incoherent arguments: the grid cannot be centered on the origin with the constraint "
                           f"to have the origin on the cell vertex and n{name} being odd; set either "
                           f"`origin_on_vertex=false` or make n{name} even.")
        if c and (not origin_on_vertex) and n % 2 == 0:
            raise IGGError(f"Incoherent arguments: the grid cannot be centered on the origin with the constraint "
                           f"to have the origin the cell center and n{name} being even; set either "
                           f"`origin_on_vertex=true` or make n{name} odd.")
    if nx == 1:
        raise IGGError("Invalid arguments: nx can never be 1.")
    if ny == 1 and nz > 1:
        raise IGGError("Invalid arguments: ny cannot be 1 if nz is greater than 1.")
    if any(n == 1 and d > 1 for n, d in zip(nxyz, dims)):
        raise IGGError("Incoherent arguments: if nx, ny, or nz is 1, then the corresponding dimx, dimy or dimz "
                       "must not be set (or set 0 or 1).")
    if any(n < 2 * o - 1 and p > 0 for n, o, p in zip(nxyz, overlaps, periods)):
        raise IGGError("Incoherent arguments: if nx, ny, or nz is smaller than 2*overlaps[1]-1, 2*overlaps[2]-1 "
                       "or 2*overlaps[3]-1, respectively, then the corresponding periodx, periody or periodz must "
                       "not be set (or set 0).")
    for i in range(3):
        if nxyz[i] == 1 and dims[i] == 0:
            dims[i] = 1                                                # src/shared.jl:198
    return tuple(nxyz), tuple(dims), tuple(periods), tuple(overlaps), tuple(halowidths)


def make_grids(nprocs: int, nx: int, ny: int = 1, nz: int = 1, *, dimx=0, dimy=0, dimz=0,
               periodx=0, periody=0, periodz=0, overlaps=(2, 2, 2), halowidths=None,
               origin=(0.0, 0.0, 0.0), origin_on_vertex=False, centerx=False, centery=False, centerz=False,
               disp=1) -> List[GlobalGrid]:
    """`init_global_grid` on every one of `nprocs` simulated ranks (src/init_global_grid.jl:83-113,
    src/create_global_grid.jl:93-102).  Rank order is the identity (reorder ignored)."""
    nxyz, dims, periods, overlaps, halowidths = normalize_input(
        nx, ny, nz, dimx, dimy, dimz, periodx, periody, periodz, overlaps, halowidths,
        origin, origin_on_vertex, centerx, centery, centerz)
    dims = dims_create(nprocs, dims)
    nxyz_g = tuple(d * (n - o) + o * (p == 0) for d, n, o, p in zip(dims, nxyz, overlaps, periods))
    org = tuple(float(o) for o in (tuple(origin) + (0.0, 0.0, 0.0))[:3]) if isinstance(origin, (tuple, list)) \
        else (float(origin), 0.0, 0.0)
    grids = []
    for me in range(nprocs):
        coords = cart_coords(me, dims)
        left, right = [], []
        for i in range(NDIMS_MPI):
            l, r = cart_shift(coords, dims, periods, i, disp)
            left.append(l)
            right.append(r)
        grids.append(GlobalGrid(nxyz_g, nxyz, dims, overlaps, halowidths, org, origin_on_vertex,
                                (centerx, centery, centerz), nprocs, me, coords,
                                (tuple(left), tuple(right)), periods, disp))
    return grids


# --------------------------------------------------------------------------------------------------
# Fields, overlap, ranges (src/shared.jl:220,246-256; src/update_halo.jl:95,281-302)
# --------------------------------------------------------------------------------------------------

def size3(A_or_shape) -> Tuple[int, int, int]:
    shp = tuple(A_or_shape.shape) if hasattr(A_or_shape, "shape") else tuple(A_or_shape)
    return tuple(shp) + (1,) * (3 - len(shp))


def ol(gg: GlobalGrid, dim: int, A) -> int:
    """ol(dim, A) = overlaps[dim] + (size(A,dim) - nxyz[dim])   (src/shared.jl:220)."""
    return gg.overlaps[dim - 1] + (size3(A)[dim - 1] - gg.nxyz[dim - 1])


def wrap_field(gg: GlobalGrid, A, hw=None):
    """(A, halowidths) named tuple of the reference (src/shared.jl:252-256)."""
    if isinstance(A, tuple):
        return A[0], tuple(int(h) for h in A[1])
    return A, tuple(gg.halowidths) if hw is None else tuple(int(h) for h in hw)


def halosize(dim: int, A, hw) -> Tuple[int, int, int]:
    """src/update_halo.jl:95."""
    s = list(size3(A))
    s[dim - 1] = hw[dim - 1]
    return tuple(s)


def sendranges(gg: GlobalGrid, n: int, dim: int, A, hw) -> List[Tuple[int, int]]:
    """src/update_halo.jl:281-290.  Returns [(lo,hi)]*3, 1-based inclusive."""
    o = ol(gg, dim, A)
    if o < 2 * hw[dim - 1]:
        raise IGGError("Incoherent arguments: ol(A,dim)<2*halowidths[dim].")
    s = size3(A)
    if n == 2:
        i = s[dim - 1] - (o - 1)
    else:
        i = 1 + (o - hw[dim - 1])
    r = [(1, s[0]), (1, s[1]), (1, s[2])]
    r[dim - 1] = (i, i + hw[dim - 1] - 1)
    return r


def recvranges(gg: GlobalGrid, n: int, dim: int, A, hw) -> List[Tuple[int, int]]:
    """src/update_halo.jl:293-302."""
    o = ol(gg, dim, A)
    if o < 2 * hw[dim - 1]:
        raise IGGError("Incoherent arguments: ol(A,dim)<2*halowidths[dim].")
    s = size3(A)
    if n == 2:
        i = s[dim - 1] - (hw[dim - 1] - 1)
    else:
        i = 1
    r = [(1, s[0]), (1, s[1]), (1, s[2])]
    r[dim - 1] = (i, i + hw[dim - 1] - 1)
    return r


def _slices(ranges):
    return tuple(slice(lo - 1, hi) for lo, hi in ranges)


def _view3(A):
    """3-D view of a 1-3-D array without copying (trailing singleton dims, like size(A,3)==1 in Julia)."""
    return A.reshape(size3(A), order="F") if A.ndim < 3 else A


def write_h2h(sendbuf, A, ranges):
    """Pack: buffer (column-major box of shape halosize) <- A[ranges]   (src/update_halo.jl:308-321)."""
    sendbuf[...] = _view3(A)[_slices(ranges)]


def read_h2h(recvbuf, A, ranges):
    """Unpack: A[ranges] <- buffer   (src/update_halo.jl:324-337)."""
    _view3(A)[_slices(ranges)] = recvbuf


# --------------------------------------------------------------------------------------------------
# check_fields (src/update_halo.jl:418-480) -- the subset that does not need Julia type information
# --------------------------------------------------------------------------------------------------

def _join(idx):
    idx = [str(i) for i in idx]
    return ", ".join(idx[:-1]) + " and " + idx[-1] if len(idx) > 1 else idx[0]


def check_fields(gg: GlobalGrid, fields):
    inv = [i + 1 for i, (A, hw) in enumerate(fields) if any(hw[d] < 1 for d in range(A.ndim))]
    if len(inv) > 1:
        raise IGGError(f"The fields at positions {_join(inv)} have a halowidth less than 1.")
    if inv:
        raise IGGError(f"The field at position {inv[0]} has a halowidth less than 1.")
    no_halo = [i + 1 for i, (A, hw) in enumerate(fields)
               if all(ol(gg, d + 1, A) < 2 * hw[d] for d in range(A.ndim))]
    if len(no_halo) > 1:
        raise IGGError(f"The fields at positions {_join(no_halo)} have no halo; remove them from the call.")
    if no_halo:
        raise IGGError(f"The field at position {no_halo[0]} has no halo; remove it from the call.")
    dup = [(i + 1, j + 1) for i in range(len(fields)) for j in range(i + 1, len(fields))
           if fields[i][0] is fields[j][0]]
    if len(dup) > 2:
        raise IGGError("The pairs of fields with the positions " + _join([f"[{a}, {b}]" for a, b in dup]) +
                       " are the same; remove any duplicates from the call.")
    if dup:
        raise IGGError(f"The field at position {dup[0][1]} is a duplicate of the one at the position {dup[0][0]}; "
                       "remove the duplicate from the call.")
    diff = [i + 1 for i in range(1, len(fields))
            if fields[i][0].dtype != fields[0][0].dtype or fields[i][0].ndim != fields[0][0].ndim]
    if len(diff) > 1:
        raise IGGError(f"The fields at positions {_join(diff)} are of different type than the first field; make "
                       "sure that in a same call all fields are of the same type.")
    if diff:
        raise IGGError(f"The field at position {diff[0]} is of different type than the first field; make sure that "
                       "in a same call all fields are of the same type.")


# --------------------------------------------------------------------------------------------------
# update_halo! over N simulated ranks (src/update_halo.jl:30-89)
# --------------------------------------------------------------------------------------------------

def update_halo(grids: List[GlobalGrid], fields_per_rank, dims=(3, 1, 2), check=True):
    """`update_halo!(A...)` executed on every simulated rank, in place.

    fields_per_rank[r] = list of arrays or (array, halowidths) tuples of rank r (same order on every rank).
    Follows the reference's three sequential phases (default order z, x, y; src/update_halo.jl:30,51): in each
    phase every rank packs its send boxes of both sides for all fields (:52-54), the buffers travel to the
    neighbour (:57-66 MPI, or :68-74 local copy when the rank is its own neighbour), and are unpacked into the
    receive boxes (:78-81).  My send buffer of side n lands in the receive buffer of side 3-n on
    neighbor(n,dim) -- for the MPI branch this follows from the posting order (recv n=2,1; send n=1,2) and
    in-order matching; for the local branch it is explicit (:373-383).
    """
    nprocs = len(grids)
    F = [[wrap_field(g, f) for f in fields_per_rank[r]] for r, g in enumerate(grids)]
    if check:
        for r, g in enumerate(grids):
            check_fields(g, F[r])
    for dim in dims:
        # pack (iwrite_sendbufs!, src/update_halo.jl:236-243)
        sendbufs = [dict() for _ in range(nprocs)]
        for r, g in enumerate(grids):
            for n in (1, 2):
                if not g.has_neighbor(n, dim):
                    continue
                for i, (A, hw) in enumerate(F[r]):
                    if ol(g, dim, A) >= 2 * hw[dim - 1]:
                        buf = np.empty(halosize(dim, A, hw), dtype=A.dtype, order="F")
                        write_h2h(buf, A, sendranges(g, n, dim, A, hw))
                        sendbufs[r][(n, i)] = buf
        # coherence check (src/update_halo.jl:57,68,76)
        for r, g in enumerate(grids):
            nb = (g.neighbor(1, dim), g.neighbor(2, dim))
            if not (all(x != g.me for x in nb) or all(x == g.me for x in nb)):
                raise IGGError(f"Incoherent neighbors in dimension {dim}: either all neighbors must equal to me, "
                               "or none.")
        # exchange + unpack (irecv/isend or sendrecv_halo_local, then iread_recvbufs!)
        for r, g in enumerate(grids):
            for n in (1, 2):
                if not g.has_neighbor(n, dim):
                    continue
                dst = g.neighbor(n, dim)
                for i, (A, hw) in enumerate(F[r]):
                    if (n, i) not in sendbufs[r]:
                        continue
                    B, hwB = F[dst][i]
                    read_h2h(sendbufs[r][(n, i)], B, recvranges(grids[dst], 3 - n, dim, B, hwB))
    return None


# --------------------------------------------------------------------------------------------------
# gather! (src/gather.jl:25-54)
# --------------------------------------------------------------------------------------------------

def gather(grids: List[GlobalGrid], blocks, A_global, root: int = 0):
    """Every rank r contributes blocks[r]; root's A_global (numpy, order='F') receives block (cx,cy,cz) at
    [cx*nxA : (cx+1)*nxA, ...] (src/gather.jl:36-49; rank enumeration is row-major :40-47)."""
    g0 = grids[root]
    if A_global is None:
        raise IGGError("The input argument `A_global` can't be `nothing` on the root.")
    A = blocks[root]
    N2, N = A.ndim, A_global.ndim
    if N2 > N:
        raise IGGError("The number of dimension of `A` must be less than or equal to the number of dimensions "
                       "of `A_global`.")
    if N > 3:
        raise IGGError("The number of dimensions of `A_global` must be less than or equal to the number of "
                       "dimensions of the Cartesian grid of MPI processes.")
    dims = tuple(g0.dims[:N])
    size_A = tuple(A.shape) + (1,) * (N - N2)
    if tuple(A_global.shape) != tuple(d * s for d, s in zip(dims, size_A)):
        raise IGGError("The size of the global array `size(A_global)` must be equal to the product of `size(A)` "
                       "and `dims`.")
    for r, g in enumerate(grids):
        c = g.coords[:N]
        sl = tuple(slice(ci * si, (ci + 1) * si) for ci, si in zip(c, size_A))
        A_global[sl] = blocks[r].reshape(size_A, order="F")
    return None


# --------------------------------------------------------------------------------------------------
# Global size / index / coordinate helpers (src/tools.jl:47-380)
# --------------------------------------------------------------------------------------------------

def _n_g_dims(gg: GlobalGrid, d: int, dimd: Optional[int] = None) -> int:
    """_nx_g(dimx) (src/tools.jl: `_nx_g(dimx) = dimx*(nx-olx) + olx*!periodx`)."""
    dimd = gg.dims[d] if dimd is None else dimd
    return dimd * (gg.nxyz[d] - gg.overlaps[d]) + gg.overlaps[d] * (0 if gg.periods[d] else 1)


def n_g(gg: GlobalGrid, dim: int, nA: Optional[int] = None) -> int:
    """nx_g() / nx_g(A) (`_adjust_nx_g`, src/tools.jl)."""
    d = dim - 1
    base = gg.nxyz_g[d]
    if nA is None:
        return base
    diff = nA - gg.nxyz[d]
    p = 1 if gg.periods[d] else 0
    return base + diff * (1 - p) + min(diff + gg.overlaps[d], 0) * p


def i_g(gg: GlobalGrid, dim: int, i: int, nA: int, wrap_periodic: bool = True) -> int:
    """ix_g / iy_g / iz_g (`_ix_g`, src/tools.jl).  `÷` truncates toward zero."""
    d = dim - 1
    diff = nA - gg.nxyz[d]
    olA = gg.overlaps[d] + diff
    nga = n_g(gg, dim, nA)
    i0 = -int(olA / 2) if gg.periods[d] else 0
    ii = gg.coords[d] * (gg.nxyz[d] - gg.overlaps[d]) + i + i0
    if gg.periods[d] and wrap_periodic:
        if ii > nga:
            ii -= nga
        if ii < 1:
            ii += nga
    return ii


def x_g(gg: GlobalGrid, dim: int, i: int, dx: float, nA: int, wrap_periodic: bool = True) -> float:
    """x_g / y_g / z_g (`_x_g`, src/tools.jl).  The Julia range x1:dx:x2 is evaluated as x1 + (k-1)*dx, which
    is exact for the dyadic steps the reference tests use."""
    d = dim - 1
    ig = i_g(gg, dim, i, nA, wrap_periodic)
    diff = nA - gg.nxyz[d]
    ng = _n_g_dims(gg, d)
    l = dx * ng if gg.origin_on_vertex else dx * (ng - 1)
    lA = l + dx * diff
    centershift = -l * 0.5 if gg.centerxyz[d] else 0.0
    vertexshift = 0.5 if gg.origin_on_vertex else 0.0
    localshift = -diff * 0.5 if not gg.periods[d] else 0.0
    staggershift = -0.5 if (gg.periods[d] and diff % 2 != 0) else 0.0
    x1 = gg.origin[d] + centershift + dx * (vertexshift + localshift + staggershift)
    npts = int(np.floor(lA / dx + 1e-9)) + 1            # length(x1:dx:x1+lA)
    if ig > npts:
        return x1 + (npts - 1) * dx + dx * (ig - npts)
    if ig < 1:
        return x1 - dx * (1 - ig)
    return x1 + (ig - 1) * dx


def coord_field(gg: GlobalGrid, shape, dtype=np.float64, dxyz=(1.0, 1.0, 1.0), scale=(1.0, 10.0, 100.0)):
    """The reference tests' coordinate-coded field z_g*1e2 + y_g*1e1 + x_g (test/test_update_halo.jl:1007)."""
    s = size3(shape)
    nd = len(tuple(shape))
    out = np.zeros(s, dtype=np.float64, order="F")
    for d in range(nd):
        v = np.array([x_g(gg, d + 1, i, dxyz[d], s[d]) for i in range(1, s[d] + 1)])
        sh = [1, 1, 1]
        sh[d] = s[d]
        out += scale[d] * v.reshape(sh)
    return np.asfortranarray(out.reshape(tuple(shape), order="F").astype(dtype))


# --------------------------------------------------------------------------------------------------
# Synthetic inputs used by the parity tests and the bench
# --------------------------------------------------------------------------------------------------

def random_bits_field(shape, elsize: int, seed: int):
    """Full-range random bit patterns (includes NaN/denormal encodings) as an unsigned-int array, order='F'."""
    rng = np.random.default_rng(seed)
    dt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[elsize]
    a = rng.integers(0, np.iinfo(dt).max, size=tuple(shape), dtype=dt, endpoint=True)
    return np.asfortranarray(a)


def all_offsets():
    return [o for o in itertools.product((-1, 0, 1), repeat=3) if o != (0, 0, 0)]
