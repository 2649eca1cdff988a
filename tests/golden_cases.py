"""Loader for tests/golden/halo_cases.{npz,json} (written by tests/golden/make_golden.py; closed-form known answers
for update_halo! on 1-8 ranks, built without the oracle)."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

_cache = {}


def load():
    if not _cache:
        with open(os.path.join(GOLDEN, "halo_cases.json")) as f:
            meta = json.load(f)
        _cache["cases"] = meta["cases"]
        _cache["npz"] = np.load(os.path.join(GOLDEN, "halo_cases.npz"))
    return _cache["cases"], _cache["npz"]


def cases(nprocs=None):
    """[(index, case dict)] of all cases, or of those for `nprocs` ranks."""
    cs, _ = load()
    return [(i, c) for i, c in enumerate(cs) if nprocs is None or c["nprocs"] == nprocs]


def arrays(ci, rank, which):
    """F-ordered copies of the fields of case ci on `rank`; which = "in" (before) or "out" (after update_halo!)."""
    cs, npz = load()
    return [np.asfortranarray(npz[f"c{ci}_r{rank}_f{fi}_{which}"]).copy(order="F") for fi in range(len(cs[ci]["shapes"]))]


def kwargs(case):
    kw = dict(case["kw"])
    for k in ("overlaps", "halowidths"):
        if k in kw:
            kw[k] = tuple(kw[k])
    return kw


def halowidths(case, default):
    return [tuple(h) for h in case["hws"]] if case["hws"] else [tuple(default)] * len(case["shapes"])
