"""igg_update_halo_noalloc: the halo update for callers that may not allocate on the GPU.

Some users of this library run where cudaMalloc is forbidden -- their own memory pool owns the device, or they sit
inside a time loop that has promised "no allocation after start-up".  For them the library has one extra C entry point
whose only scratch (the kernel's tile table) is a fixed array built into the library.  This file checks, through the
C ABI, that it produces exactly what the regular igg_update_halo produces (bit for bit -- the data path only copies),
on two calls in a row (the second call reuses the shared array and must not be disturbed by what the first left
behind), and that it refuses work that does not fit the array.

The file name makes pytest collect it last, so that a run stopping at the first failure has already run everything
else; the test does not depend on that order and also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import igg_b200 as igg
from igg_b200 import _lib as L

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu_and_clean_grid():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    yield
    if igg.grid_is_initialized():
        igg.finalize_global_grid()


def _random_field(shape, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    t = igg.zeros(shape, dtype=torch.int64)
    t.copy_(torch.randint(-2**62, 2**62, tuple(shape), dtype=torch.int64, device="cuda", generator=g))
    return t


def _call_noalloc(desc, tensors, hws, dims=(3, 1, 2), flags=L.IGG_SYNC):
    lib = L.load()
    cf = [L.make_field(t.data_ptr(), tuple(t.shape), hw, t.element_size()) for t, hw in zip(tensors, hws)]
    arr = (L.Field * len(cf))(*cf)
    d = (C.c_int * len(dims))(*dims)
    return lib.igg_update_halo_noalloc(C.byref(desc), len(cf), arr, d, len(dims),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream), flags)


@pytest.mark.parametrize("gsz,kw,shapes", [
    ((40, 24, 18), dict(periodx=1, periody=1, periodz=1), [(40, 24, 18), (41, 24, 18), (40, 24, 19)]),
    ((33, 17, 20), dict(periodx=1, periodz=1, overlaps=(4, 2, 4), halowidths=(2, 1, 2)), [(33, 17, 20), (33, 18, 20)]),
    ((128, 96, 1), dict(periodx=1, periody=1), [(128, 96)]),
])
def test_matches_regular_entry_point_on_two_calls_in_a_row(gsz, kw, shapes):
    igg.init_global_grid(*gsz, quiet=True, **kw)
    g = igg.global_grid()
    hws = [g.halowidths] * len(shapes)
    a = [_random_field(s, 100 + i) for i, s in enumerate(shapes)]       # goes through igg_update_halo
    b = [t.clone(memory_format=torch.preserve_format) for t in a]       # goes through igg_update_halo_noalloc
    assert all(x.stride() == y.stride() for x, y in zip(a, b))
    for call in range(2):
        igg.update_halo(*a)
        L.check(_call_noalloc(g.desc, b, hws))
        torch.cuda.synchronize()
        for x, y in zip(a, b):
            assert torch.equal(x, y), f"call {call}: igg_update_halo_noalloc differs from igg_update_halo"
        # new interior data for the second call (identical in both copies); halos keep the old values until updated
        for i, (x, y) in enumerate(zip(a, b)):
            fresh = _random_field(tuple(x.shape), 200 + 10 * call + i)
            inner = tuple(slice(2, -2) for _ in range(x.dim()))
            x[inner] = fresh[inner]
            y[inner] = fresh[inner]
    # a third, stream-ordered (non-blocking) call followed by a different, smaller plan: the shared table is rewritten
    L.check(_call_noalloc(g.desc, b, hws, flags=L.IGG_ASYNC))
    L.check(_call_noalloc(g.desc, b[:1], hws[:1], dims=(1,), flags=L.IGG_ASYNC))
    igg.update_halo(*a)
    igg.update_halo(a[0], dims=(1,))
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_refuses_what_does_not_fit_and_multi_rank_grids():
    lib = L.load()
    assert lib.igg_noalloc_scratch_bytes() < (1 << 20)
    n = 512
    igg.init_global_grid(n, n, n, periodx=1, periody=1, periodz=1, overlaps=(4, 4, 4), halowidths=(2, 2, 2), quiet=True)
    g = igg.global_grid()
    fields = [igg.zeros((n, n, n), dtype=torch.float32) for _ in range(3)]   # ~36 MiB of halo data > 25 MiB limit
    before = [t.clone() for t in fields]
    rc = _call_noalloc(g.desc, fields, [g.halowidths] * 3)
    assert rc == L.IGG_ERR_ARG
    msg = lib.igg_last_error().decode()
    assert "scratch holds at most" in msg and "igg_update_halo" in msg
    for t, b in zip(fields, before):
        assert torch.equal(t, b)                                              # nothing was touched
    assert _call_noalloc(g.desc, fields[:1], [g.halowidths]) == 0              # one field fits
    import copy
    d2 = copy.copy(g.desc)
    d2 = type(g.desc).from_buffer_copy(g.desc)
    d2.nprocs = 2
    rc = _call_noalloc(d2, fields[:1], [g.halowidths])
    assert rc == L.IGG_ERR_UNSUPPORTED and "nprocs == 1" in lib.igg_last_error().decode()


def test_allocates_nothing_and_refuses_stream_capture():
    """The point of the entry point: device memory in use does not change across calls (first call of a plan, repeated
    call, a different plan), and a call on a capturing stream is refused instead of recording a copy from host memory
    that is gone when the graph is replayed."""
    igg.init_global_grid(40, 24, 18, periodx=1, periody=1, periodz=1, quiet=True)
    g = igg.global_grid()
    a = [_random_field((40, 24, 18), 7), _random_field((41, 24, 18), 8)]
    hws = [g.halowidths] * 2
    L.check(_call_noalloc(g.desc, a, hws))            # the first call on this device in this process may set up the runtime
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        L.check(_call_noalloc(g.desc, a, hws))
    L.check(_call_noalloc(g.desc, a[:1], hws[:1], dims=(2, 3)))
    L.check(_call_noalloc(g.desc, a, hws, flags=L.IGG_ASYNC))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    s = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    lib = L.load()
    with torch.cuda.stream(s):
        cf = [L.make_field(t.data_ptr(), tuple(t.shape), hw, 8) for t, hw in zip(a, hws)]
        arr = (L.Field * 2)(*cf)
        with torch.cuda.graph(graph, stream=s):
            rc = lib.igg_update_halo_noalloc(C.byref(g.desc), 2, arr, None, 0, C.c_void_p(s.cuda_stream), L.IGG_ASYNC)
    assert rc == L.IGG_ERR_UNSUPPORTED and "captured" in lib.igg_last_error().decode()
