"""Process topology and argument normalisation of the host side (no GPU, no library calls).

In the Julia host these are MPI.jl calls; here they are restated so that the Python harness needs no MPI:
`MPI.Dims_create!`, `MPI.Cart_create` (row-major ranks), `MPI.Cart_coords`, `MPI.Cart_shift`
(ref: src/create_global_grid.jl:93-101) and `normalize_input` (ref: src/shared.jl:156-207).
"""
from __future__ import annotations

PROC_NULL = -1


class ArgumentError(RuntimeError):
    """Stand-in for the reference's ErrorException raised by `error(...)`."""


def _factorizations(n, k, maxf):
    if k == 1:
        if n <= maxf:
            yield (n,)
        return
    f = min(n, maxf)
    while f >= 1:
        if n % f == 0:
            for tail in _factorizations(n // f, k - 1, f):
                yield (f,) + tail
        f -= 1


def dims_create(nprocs, dims):
    """MPI_Dims_create: zero entries are filled, as balanced as possible, in non-increasing order."""
    dims = [int(d) for d in dims]
    if any(d < 0 for d in dims):
        raise ArgumentError("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
    fixed = 1
    for d in dims:
        if d > 0:
            fixed *= d
    if nprocs % fixed:
        raise ArgumentError(f"Dims_create: {nprocs} processes cannot be arranged with the fixed dims {dims}.")
    free = [i for i, d in enumerate(dims) if d == 0]
    rest = nprocs // fixed
    if not free:
        if rest != 1:
            raise ArgumentError(f"Dims_create: prod(dims)={fixed} does not match the number of processes {nprocs}.")
        return tuple(dims)
    best = min(_factorizations(rest, len(free), rest))
    for i, f in zip(free, best):
        dims[i] = f
    return tuple(dims)


def cart_rank(coords, dims):
    r = 0
    for c, d in zip(coords, dims):
        r = r * d + c
    return r


def cart_coords(rank, dims):
    out = []
    for d in reversed(dims):
        out.append(rank % d)
        rank //= d
    return tuple(reversed(out))


def cart_shift(coords, dims, periods, direction, disp):
    res = []
    for s in (-disp, disp):
        c = list(coords)
        c[direction] += s
        if periods[direction]:
            c[direction] %= dims[direction]
            res.append(cart_rank(c, dims))
        elif 0 <= c[direction] < dims[direction]:
            res.append(cart_rank(c, dims))
        else:
            res.append(PROC_NULL)
    return tuple(res)


def normalize_input(nx, ny, nz, dimx, dimy, dimz, periodx, periody, periodz, origin, origin_on_vertex, centerx,
                    centery, centerz, overlaps, halowidths, disp):
    """ref: src/shared.jl:156-207 (same checks, same messages)."""
    dims = [int(dimx), int(dimy), int(dimz)]
    periods = [int(periodx), int(periody), int(periodz)]
    nxyz = [int(nx), int(ny), int(nz)]
    if isinstance(origin, (int, float)):
        origin = (float(origin), 0.0, 0.0)
    origin = tuple(float(o) for o in origin) + (0.0,) * (3 - len(tuple(origin)))
    if any(d < 0 for d in dims):
        raise ArgumentError("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
    if any(p not in (0, 1) for p in periods):
        raise ArgumentError("Invalid arguments: periodx, periody, and periodz must be either 0 or 1.")
    if len(origin) != 3:
        raise ArgumentError("Invalid argument: the length of the origin tuple must be at most 3.")
    if any(h < 1 for h in halowidths):
        raise ArgumentError("Invalid arguments: halowidths cannot be less than 1.")
    if any(o > 0 and h > o // 2 for o, h in zip(overlaps, halowidths)):
        raise ArgumentError("Incoherent arguments: if overlap is greater than 0, then halowidth cannot be greater than "
                            "overlap÷2, in each dimension.")
    if any(n < 1 for n in nxyz):
        raise ArgumentError("Invalid arguments: nx, ny, and nz cannot be less than 1.")
    for c, n, ax in zip((centerx, centery, centerz), nxyz, "xyz"):
        if c and origin_on_vertex and n % 2 == 1:
            raise ArgumentError(f"Incoherent arguments: the grid cannot be centered on the origin with the constraint to "
                                f"have the origin on the cell vertex and n{ax} being odd; set either "
                                f"`origin_on_vertex=false` or make n{ax} even.")
        if c and not origin_on_vertex and n % 2 == 0:
            raise ArgumentError(f"Incoherent arguments: the grid cannot be centered on the origin with the constraint to "
                                f"have the origin the cell center and n{ax} being even; set either "
                                f"`origin_on_vertex=true` or make n{ax} odd.")
    if nxyz[0] == 1:
        raise ArgumentError("Invalid arguments: nx can never be 1.")
    if nxyz[1] == 1 and nxyz[2] > 1:
        raise ArgumentError("Invalid arguments: ny cannot be 1 if nz is greater than 1.")
    if any(n == 1 and d > 1 for n, d in zip(nxyz, dims)):
        raise ArgumentError("Incoherent arguments: if nx, ny, or nz is 1, then the corresponding dimx, dimy or dimz must "
                            "not be set (or set 0 or 1).")
    if any(n < 2 * o - 1 and p > 0 for n, o, p in zip(nxyz, overlaps, periods)):
        raise ArgumentError("Incoherent arguments: if nx, ny, or nz is smaller than 2*overlaps[1]-1, 2*overlaps[2]-1 or "
                            "2*overlaps[3]-1, respectively, then the corresponding periodx, periody or periodz must not "
                            "be set (or set 0).")
    for i in range(3):
        if nxyz[i] == 1 and dims[i] == 0:
            dims[i] = 1
    return tuple(nxyz), tuple(dims), tuple(periods), origin
