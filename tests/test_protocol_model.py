"""A small executable model of the exchange protocol under rank skew (CPU only).

The CUDA path lets ranks drift apart: a rank starts its next `update_halo!` as soon as its own unpacks are done.
What keeps a fast neighbour from overwriting staged data that is still unread is (a) two staging halves chosen by
the parity of the per-neighbour-offset exchange count and (b) one staging REGION per neighbour offset whose position
does not depend on the call.  This test takes the real plans of the library for a sequence of
different calls (field sets, `dims=`), replays them on R asynchronous model ranks under many random schedules, and
checks every unpack reads exactly what the matching push of the same call wrote.

It also replays the same schedules with the layout the library had before regions existed (every plan packs its
offsets from 0): there the model must FIND the overwrite -- which shows the model is sharp enough to see the bug the
regions were introduced for."""
import random

import numpy as np
import pytest

from igg_b200 import _lib as L
from oracle import igg_oracle as O
from test_plan_vs_oracle import desc_of
from test_tiles_cpu import OFFSETS, PUSH, PUSH_LL, peer_of

UNIT = 128          # staging granule of the model (all slot sizes are multiples of 128 bytes)


def plans_for(grids, calls, mode):
    """Per call and rank: {offset index: (peer rank, need in units)} and the packed starts of that plan."""
    out = []
    for shapes, dims in calls:
        per_rank = []
        for g in grids:
            cf = [L.make_field(0x10000 * (i + 1), s, g.halowidths, 8) for i, s in enumerate(shapes)]
            tiles, _, reg = L.plan_tiles(desc_of(g), cf, dims, mode, regions=True)
            active = sorted({t.nbr for t in tiles if t.kind in (PUSH, PUSH_LL)})
            per_rank.append(dict(active={o: peer_of(g, o) for o in active},
                                 pack=[b // UNIT for b in reg["byte_pack"]], need=[b // UNIT for b in reg["byte_need"]]))
        out.append(per_rank)
    return out


def run_model(nranks, plans, layout, rng, steps_limit=100000):
    """Returns None if every unpack saw its own call's data, else a description of the first violation."""
    ncalls = len(plans)
    cap = [max(plans[c][0]["need"][o] for c in range(ncalls)) for o in range(27)]       # identical on all ranks
    region = np.concatenate([[0], np.cumsum(cap)])[:27]
    half_units = int(sum(cap)) + 1
    mem = [np.zeros((2, half_units), dtype=np.int64) for _ in range(nranks)]            # tag = (call + 1) * 100 + sender
    ex = [[0] * 27 for _ in range(nranks)]                                              # exchanges per offset
    call = [0] * nranks
    todo = [None] * nranks                                                              # offsets still to unpack

    def start(o, c, r):
        return int(region[o]) if layout == "regions" else plans[c][r]["pack"][o]

    for _ in range(steps_limit):
        ready = [r for r in range(nranks) if call[r] < ncalls]
        if not ready:
            return None
        r = rng.choice(ready)
        c = call[r]
        me = plans[c][r]
        if todo[r] is None:                               # PUSH phase: all pushes of the call, fire and forget
            for o, peer in me["active"].items():
                ro = 26 - o
                lo = start(ro, c, peer)
                n = plans[c][peer]["need"][ro]
                mem[peer][ex[r][o] & 1, lo:lo + n] = (c + 1) * 100 + r
            todo[r] = list(me["active"])
            rng.shuffle(todo[r])
            continue
        progressed = False
        for o in list(todo[r]):                           # UNPACK phase: any offset whose data has arrived
            peer = me["active"][o]
            lo, n = start(o, c, r), me["need"][o]
            got = mem[r][ex[r][o] & 1, lo:lo + n]
            want = (c + 1) * 100 + peer
            if np.all(got == want):
                todo[r].remove(o)
                progressed = True
            elif np.any(got // 100 > c + 1) or (np.any(got == want) and not np.all((got == want) | (got // 100 < c + 1))):
                return f"rank {r} call {c} offset {OFFSETS[o]}: expected data of rank {peer}, found tags {sorted(set(got.tolist()))}"
        if not todo[r]:
            for o in me["active"]:
                ex[r][o] += 1
            call[r] += 1
            todo[r] = None
        elif not progressed:
            continue                                      # still waiting: somebody else must move first
    return "model did not finish"


def scenario(nprocs, periodic):
    kw = dict(periodx=1, periody=1, periodz=1) if periodic else {}
    n = (12, 10, 9)
    grids = O.make_grids(nprocs, *n, **kw)
    A, B, C = n, (n[0] + 1, n[1], n[2]), (n[0], n[1], n[2] + 1)
    calls = [([A], (1,)), ([A, B], (3, 1, 2)), ([B], (2,)), ([A], (3, 1, 2)), ([A, B, C], (1, 3)), ([C], (3,)),
             ([B, C], (2, 1)), ([A], (1,)), ([A, B, C], (3, 1, 2)), ([B], (3, 1, 2))] * 2
    return grids, calls


@pytest.mark.parametrize("mode", [L.IGG_MODE_IPC_1PHASE, L.IGG_MODE_IPC_LL128], ids=["fence", "ll128"])
@pytest.mark.parametrize("nprocs,periodic", [(2, True), (4, True), (8, True), (8, False), (12, True)])
def test_regions_and_parity_keep_staged_data_intact_under_skew(nprocs, periodic, mode):
    grids, calls = scenario(nprocs, periodic)
    plans = plans_for(grids, calls, mode)
    rng = random.Random(1234 + nprocs)
    for trial in range(40):
        bad = run_model(nprocs, plans, "regions", rng)
        assert bad is None, f"trial {trial}: {bad}"


def test_model_finds_the_overwrite_with_per_plan_layouts():
    """The layout before regions: every plan packs its offsets from 0, so different neighbours' slots of different
    calls share addresses.  Some schedule must show a neighbour that is a call ahead destroying unread data."""
    grids, calls = scenario(8, True)
    plans = plans_for(grids, calls, L.IGG_MODE_IPC_LL128)
    rng = random.Random(7)
    found = None
    for trial in range(300):
        found = run_model(8, plans, "packed", rng)
        if found and found != "model did not finish":
            break
    assert found and found != "model did not finish", "the model never produced the known hazard"
