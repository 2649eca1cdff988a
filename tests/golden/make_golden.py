"""Generates tests/golden/halo_cases.{npz,json}: closed-form known answers for update_halo! on 1, 2, 4 and 8 ranks.

    python tests/golden/make_golden.py          # rewrites the two files (deterministic)

The Julia reference cannot run here (no Julia, no MPI), so these vectors are NOT outputs of a reference run.  They
are built from the property the reference's own tests use (test/test_update_halo.jl:975-1344): a field whose cells
hold a function of their GLOBAL cell index is consistent across ranks, so after `update_halo!` every halo cell that
has a neighbour must hold the value of its global cell again, and every halo cell without a neighbour keeps what it
held before.  The generator does not import the oracle and never copies a halo: it only evaluates

  * the global index of local cell ix of an array of size nA on the rank at Cartesian coordinate c
    (src/tools.jl:369-380):   ix_g = c*(n-ol) + ix + (periodic ? -(ol_A div 2) : 0),  ol_A = ol + (nA-n),
    wrapped once into 1..n_g_A when periodic, n_g_A from src/tools.jl:47,78;
  * the row-major rank <-> coordinate map of MPI_Cart_create (confirmed by src/gather.jl:40-46) and the
    MPI_Dims_create factorisations 1 -> 1x1x1, 2 -> 2x1x1, 4 -> 2x2x1, 8 -> 2x2x2 (2-D: 8 -> 4x2, 1-D: n);
  * has-halo: ol_A >= 2*hw (src/update_halo.jl:239 with src/shared.jl:220);  receive slabs: the hw outermost
    planes (src/update_halo.jl:292-302).

Input  = pattern, with every receive slab of an updated dimension overwritten by the sentinel 0xA5.
Output = pattern, except cells lying in a receive slab whose side has no neighbour (non-periodic edge of the process
grid, or a dimension left out by `dims=`): those keep the sentinel, also where slabs cross.
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SENT = 0xA5

DIMS_3D = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
DIMS_2D = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (4, 2, 1)}


def dims_for(nprocs, gsz):
    if gsz[1] == 1 and gsz[2] == 1:
        return (nprocs, 1, 1)
    return (DIMS_2D if gsz[2] == 1 else DIMS_3D)[nprocs]


def coords_of(rank, dims):
    return (rank // (dims[1] * dims[2]), (rank // dims[2]) % dims[1], rank % dims[2])


def tdiv(a, b):
    """Julia's `div` (truncation towards zero)."""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b > 0) else -q


def global_index(ix, nA, n, ol, c, D, periodic):
    """1-based global index of local cell ix (1-based); src/tools.jl:369-380 with n_g_A of :47,:78."""
    diff = nA - n
    ol_A = ol + diff
    n_g = D * (n - ol) + (0 if periodic else ol)
    n_g_A = n_g + (min(diff + ol, 0) if periodic else diff)
    i = c * (n - ol) + ix + (-tdiv(ol_A, 2) if periodic else 0)
    if periodic:
        if i > n_g_A:
            i -= n_g_A
        if i < 1:
            i += n_g_A
    return i


def pattern(elsize, fid, ig):
    """Bit pattern of a cell from its global indices ig = (ix_g, iy_g, iz_g) (broadcast arrays)."""
    ix, iy, iz = [np.asarray(a, dtype=np.int64) & 0xFFFF for a in ig]
    if elsize == 8:
        return ((np.int64(fid) << 48) | (iz << 32) | (iy << 16) | ix).astype(np.uint64)
    if elsize == 4:
        return ((np.int64(fid & 1) << 30) | ((iz & 0x3FF) << 20) | ((iy & 0x3FF) << 10) | (ix & 0x3FF)).astype(np.uint32)
    if elsize == 2:
        return (((iz & 0x1F) << 10) | ((iy & 0x1F) << 5) | (ix & 0x1F)).astype(np.uint16)
    return ((ix + 7 * iy + 31 * iz + 101 * fid) % 160).astype(np.uint8)


def build_case(case):
    gsz, nprocs = case["grid"], case["nprocs"]
    per = [case["kw"].get(k, 0) for k in ("periodx", "periody", "periodz")]
    ol = list(case["kw"].get("overlaps", (2, 2, 2)))
    ghw = list(case["kw"].get("halowidths", [max(1, o // 2) for o in ol]))
    dims = dims_for(nprocs, gsz)
    updated = [(d + 1) in case["dims"] for d in range(3)]
    dt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[case["elsize"]]
    sent = np.frombuffer(bytes([SENT]) * case["elsize"], dtype=dt)[0]
    out = {}
    for rank in range(nprocs):
        c = coords_of(rank, dims)
        for fi, shape in enumerate(case["shapes"]):
            sh3 = tuple(shape) + (1,) * (3 - len(shape))
            hw = (list(case["hws"][fi]) if case["hws"] else ghw)
            ig = []
            for d in range(3):
                v = np.array([global_index(i, sh3[d], gsz[d], ol[d], c[d], dims[d], per[d]) for i in range(1, sh3[d] + 1)])
                ig.append(v.reshape([-1 if k == d else 1 for k in range(3)]))
            full = np.asfortranarray(np.broadcast_to(pattern(case["elsize"], fi, ig), sh3)).copy(order="F")
            fin, fout = full.copy(order="F"), full.copy(order="F")
            for d in range(len(shape)):
                ol_A = ol[d] + (sh3[d] - gsz[d])
                is_halo_dim = ol_A >= 2 * hw[d]
                filled = is_halo_dim or not updated[d]          # slabs of left-out dimensions get the sentinel too
                if not filled or sh3[d] < 2 * hw[d]:
                    continue
                for side in (0, 1):
                    idx = [slice(None)] * 3
                    idx[d] = slice(0, hw[d]) if side == 0 else slice(sh3[d] - hw[d], sh3[d])
                    fin[tuple(idx)] = sent
                    has_nb = per[d] or (c[d] > 0 if side == 0 else c[d] < dims[d] - 1)
                    if dims[d] == 1 and not per[d]:
                        has_nb = False
                    if not (is_halo_dim and updated[d] and has_nb):
                        fout[tuple(idx)] = sent
            out[f"r{rank}_f{fi}_in"] = fin.reshape(shape, order="F")
            out[f"r{rank}_f{fi}_out"] = fout.reshape(shape, order="F")
    return out


def case_table():
    nx, ny, nz = 7, 5, 6             # test/test_update_halo.jl:111-113
    P3 = dict(periodx=1, periody=1, periodz=1)
    non_default = dict(P3, overlaps=(4, 2, 3), halowidths=(2, 1, 1))
    t = []

    def add(name, grid, kw, shapes, hws=None, elsize=8, dims=(3, 1, 2), nprocs=(1, 2, 4, 8), ref=""):
        for n in nprocs:
            t.append(dict(name=f"{name}_n{n}", grid=list(grid), kw=kw, shapes=[list(s) for s in shapes],
                          hws=[list(h) for h in hws] if hws else None, elsize=elsize, dims=list(dims), nprocs=n, ref=ref))

    add("periodic_1d", (nx, 1, 1), dict(periodx=1), [(nx,), (nx + 1,)], nprocs=(1, 2, 4), ref="test/test_update_halo.jl:977,1057")
    add("periodic_2d", (nx, ny, 1), dict(periodx=1, periody=1), [(nx, ny), (nx, ny + 1)], ref="test/test_update_halo.jl:990,1070")
    add("periodic_3d", (nx, ny, nz), P3, [(nx, ny, nz)], ref="test/test_update_halo.jl:1004")
    add("periodic_3d_ol423_hw211", (nx, ny, nz), non_default, [(nx, ny, nz), (nx + 1, ny, nz)], ref="test/test_update_halo.jl:1019,1099")
    add("not_periodic_3d", (nx, ny, nz), dict(), [(nx, ny, nz), (nx, ny, nz + 1)], ref="test/test_update_halo.jl:1034,1114")
    add("staggered_two_fields", (nx, ny, nz), P3, [(nx, ny, nz + 1), (nx + 1, ny, nz)], ref="test/test_update_halo.jl:1293")
    add("no_halo_in_x", (nx, ny, 1), dict(periodx=1, periody=1), [(nx - 1, ny + 2)], ref="test/test_update_halo.jl:1135")
    add("no_halo_in_y", (nx, ny, nz), P3, [(nx + 2, ny - 1, nz + 1)], ref="test/test_update_halo.jl:1150")
    add("mixed_periods_f32", (nx, ny, nz), dict(periodx=1, periodz=1, overlaps=(4, 2, 3), halowidths=(2, 1, 1)),
        [(nx, ny, nz), (nx + 1, ny, nz), (nx - 1, ny + 2, nz + 1)], elsize=4, ref="test/test_update_halo.jl:1318 (mixed periods added)")
    add("per_field_halowidths", (13, 9, 9), dict(P3, overlaps=(6, 4, 4), halowidths=(3, 1, 2)),
        [(13, 9, 9), (13, 9, 9), (14, 9, 9)], hws=[(3, 1, 2), (1, 1, 2), (2, 1, 1)], nprocs=(1, 2, 8),
        ref="test/test_update_halo.jl:404-437,1194-1291 (halowidths per field)")
    add("dims_x_z_only_i16", (nx, ny, nz), P3, [(nx, ny, nz)], elsize=2, dims=(1, 3), nprocs=(1, 2, 8), ref="src/update_halo.jl:30 (dims keyword)")
    add("bytes_dims_y_x", (nx, ny, nz), dict(periody=1), [(nx, ny, nz)], elsize=1, dims=(2, 1), nprocs=(1, 4), ref="src/update_halo.jl:30 (dims keyword)")
    return t


def main():
    cases = case_table()
    arrays = {}
    for ci, case in enumerate(cases):
        for k, v in build_case(case).items():
            arrays[f"c{ci}_{k}"] = v
    np.savez_compressed(os.path.join(HERE, "halo_cases.npz"), **arrays)
    with open(os.path.join(HERE, "halo_cases.json"), "w") as f:
        json.dump(dict(sentinel=SENT, cases=cases), f, indent=1)
    print(f"{len(cases)} cases, {len(arrays)} arrays")


if __name__ == "__main__":
    main()
