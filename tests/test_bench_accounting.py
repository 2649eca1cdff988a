"""bench.py's byte accounting against SURVEY 8(d) / Appendix C and against the plan the library actually executes
(CPU only): `B_send` of cfg2 is 6 faces x 512 KiB at every GPU count, the NVLink share is the faces of dimensions
with more than one rank, and the single-phase plan moves the same cells minus the doubly counted edges."""
import importlib.util
import os

from igg_b200 import _lib as L
from oracle import igg_oracle as O
from test_plan_vs_oracle import desc_of

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def test_cfg2_halo_bytes():
    assert bench.halo_bytes_per_gpu() == 3_145_728
    assert [bench.nvlink_bytes_per_gpu(n) for n in (1, 2, 4, 8)] == [0, 1 << 20, 2 << 20, 3 << 20]
    assert [tuple(bench.dims_for(n)) for n in (1, 2, 4, 8)] == [(1, 1, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2)]


def test_plan_moves_what_the_bench_accounts_for():
    n = 256
    for nprocs in (1, 2, 8):
        grids = O.make_grids(nprocs, n, n, n, periodx=1, periody=1, periodz=1)
        g = grids[0]
        ops, staging = L.plan_describe(desc_of(g), [L.make_field(0x10000, (n, n, n), g.halowidths, 8)], None)
        moved = sum(op.nbytes for op in ops if op.kind in (L.IGG_OP_PUSH, L.IGG_OP_LOCAL))
        assert moved == (n ** 3 - (n - 2) ** 3) * 8                     # every halo cell exactly once: 3 121 216 B
        assert moved <= bench.halo_bytes_per_gpu() < moved * 1.01       # the reference's face count includes edges twice
        remote = sum(op.nbytes for op in ops if op.kind == L.IGG_OP_PUSH)
        credited = bench.nvlink_bytes_per_gpu(nprocs)                    # what the NVLink roofline is computed from
        assert 0.97 * credited <= remote <= credited
        if nprocs == 1:
            assert staging == 0 and remote == 0
        if nprocs == 2:
            assert remote == 2 * n * n * 8                               # both x halo planes, whole planes
