"""The C restatement of the reference CPU path (oracle/cpu_ref.c, the timed CPU baseline) must agree bit for bit with
the numpy oracle (which is pinned by the reference's fixtures)."""
import numpy as np
import pytest

from oracle import cpu_ref, igg_oracle as O


@pytest.mark.parametrize("nprocs,kw", [
    (1, dict(periodx=1, periody=1, periodz=1)),
    (2, dict(periodx=1, periody=1, periodz=1)),
    (4, dict(periody=1)),
    (8, dict(periodx=1, periody=1, periodz=1)),
    (8, dict()),
    (6, dict(periodz=1)),
])
@pytest.mark.parametrize("olhw", [((2, 2, 2), (1, 1, 1)), ((4, 2, 3), (2, 1, 1)), ((6, 4, 4), (3, 1, 2))])
@pytest.mark.parametrize("threads", [1, 3])
def test_c_restatement_matches_numpy_oracle(nprocs, kw, olhw, threads):
    ol, hw = olhw
    nx, ny, nz = 13, 9, 10
    grids = O.make_grids(nprocs, nx, ny, nz, overlaps=ol, halowidths=hw, **kw)
    shapes = [(nx, ny, nz), (nx + 1, ny, nz), (nx, ny, nz + 1), (nx - 1, ny + 2, nz + 1)]
    base = [[O.random_bits_field(s, 8, 31 * r + i) for i, s in enumerate(shapes)] for r in range(nprocs)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, want)
    got = [[a.copy(order="F") for a in fr] for fr in base]
    g = grids[0]
    cpu_ref.update_halo(g.dims, g.periods, g.nxyz, g.overlaps, got, [hw] * len(shapes), copy_threads=threads, iters=1)
    for r in range(nprocs):
        for i in range(len(shapes)):
            assert np.array_equal(got[r][i], want[r][i]), (r, i)


def test_large_faces_take_the_threaded_paths():
    """Faces above GG_THREADCOPY_THRESHOLD (32 KiB) go through the threaded x-/z-face copies."""
    grids = O.make_grids(2, 80, 72, 70, periodx=1, periody=1, periodz=1)
    base = [[O.random_bits_field((80, 72, 70), 8, r)] for r in range(2)]
    want = [[a.copy(order="F") for a in fr] for fr in base]
    O.update_halo(grids, want)
    g = grids[0]
    cpu_ref.update_halo(g.dims, g.periods, g.nxyz, g.overlaps, base, [g.halowidths], copy_threads=4, iters=2)
    for r in range(2):
        assert np.array_equal(base[r][0], want[r][0])
