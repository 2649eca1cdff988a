"""CPU checks against the committed golden vectors (tests/golden/halo_cases.*, closed-form known answers on 1, 2, 4
and 8 ranks; see tests/golden/make_golden.py for how they are derived from the reference's test property).

Three things are pinned to them here, without a GPU:
  * the numpy oracle (oracle/igg_oracle.py) -- including its MULTI-rank exchange, which no reference fixture covers;
  * the C restatement of the reference CPU path (oracle/cpu_ref.c), the timed CPU baseline;
  * the host side of the product: the copy plan `igg_plan_describe` returns for every rank, executed with numpy.
The GPU twins are tests/test_gpu_update_halo.py::test_golden_single_rank and the golden loop in tests/mp_worker.py.
"""
import importlib.util
import os

import numpy as np
import pytest

import golden_cases as G
from oracle import cpu_ref, igg_oracle as O
from test_plan_vs_oracle import execute_plans

ALL = G.cases()
IDS = [c["name"] for _, c in ALL]


def setup_case(ci, case):
    grids = O.make_grids(case["nprocs"], *case["grid"], **G.kwargs(case))
    hws = G.halowidths(case, grids[0].halowidths)
    fields = [G.arrays(ci, r, "in") for r in range(case["nprocs"])]
    want = [G.arrays(ci, r, "out") for r in range(case["nprocs"])]
    return grids, hws, fields, want


def check(fields, want, what):
    for r, (fr, wr) in enumerate(zip(fields, want)):
        for i, (a, w) in enumerate(zip(fr, wr)):
            assert a.dtype == w.dtype and a.shape == w.shape
            assert np.array_equal(a, w), f"{what}: rank {r} field {i}: {np.argwhere(a != w)[:4].tolist()}"


def test_generator_reproduces_the_committed_vectors():
    """The committed files are what make_golden.py produces (no drift between script and fixtures)."""
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(G.GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    import json
    table = json.loads(json.dumps(mg.case_table()))      # tuples -> lists, as stored
    cs, npz = G.load()
    assert table == cs
    for ci in range(0, len(table), 3):
        for k, v in mg.build_case(table[ci]).items():
            assert np.array_equal(npz[f"c{ci}_{k}"], v), (ci, k)


def test_vectors_are_not_trivial():
    for ci, case in ALL:
        changed = 0
        for r in range(case["nprocs"]):
            for a, b in zip(G.arrays(ci, r, "in"), G.arrays(ci, r, "out")):
                changed += int(np.count_nonzero(a != b))
        assert changed > 0 or case["nprocs"] == 1 and not any(case["kw"].get(k) for k in ("periodx", "periody", "periodz")), case["name"]


@pytest.mark.parametrize("ci,case", ALL, ids=IDS)
def test_oracle_matches_golden(ci, case):
    grids, hws, fields, want = setup_case(ci, case)
    O.update_halo(grids, [[(a, hw) for a, hw in zip(fr, hws)] for fr in fields], dims=tuple(case["dims"]))
    check(fields, want, "oracle")


@pytest.mark.parametrize("ci,case", ALL, ids=IDS)
def test_plan_execution_matches_golden(ci, case):
    grids, hws, fields, want = setup_case(ci, case)
    execute_plans(grids, fields, hws, dims=tuple(case["dims"]))
    check(fields, want, "plan")


@pytest.mark.parametrize("ci,case", ALL, ids=IDS)
def test_c_restatement_matches_golden(ci, case):
    grids, hws, fields, want = setup_case(ci, case)
    g = grids[0]
    cpu_ref.update_halo(g.dims, g.periods, g.nxyz, g.overlaps, fields, hws, dims_order=tuple(case["dims"]), copy_threads=2)
    check(fields, want, "cpu_ref.c")
