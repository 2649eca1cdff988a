"""Host-side mirror of the reference API for the halo path, on top of the C ABI of libigg_b200.

Same names, argument meaning and error behaviour as the Julia functions (Julia's `!` becomes a trailing
underscore-free name: `update_halo`, `gather`):

    init_global_grid   ref: src/init_global_grid.jl:83-113, src/create_global_grid.jl:35-111
    update_halo        ref: src/update_halo.jl:30-42       (-> igg_update_halo, one C call)
    gather             ref: src/gather.jl:18-54            (-> igg_gather, one C call)
    finalize_global_grid  ref: src/finalize_global_grid.jl:15-28
    select_device      ref: src/select_device.jl:16-39
    nx_g ... x_g ... ix_g ...  ref: src/tools.jl:47-380

Arrays are `torch` tensors holding a column-major array: a Julia array of size (nx,ny,nz) is a tensor of shape
(nx,ny,nz) with strides (1,nx,nx*ny) -- `igg_b200.zeros((nx,ny,nz))` makes one.  torch is used for device memory,
streams and (when there is more than one rank) for the one-time all-gather of the bootstrap blobs; it plays the role
MPI.jl plays in the Julia host.  All data movement happens in libigg_b200's CUDA kernel.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

from . import _lib as L
from . import topology as T
from .topology import ArgumentError

NDIMS_MPI = 3
NNEIGHBORS_PER_DIM = 2


@dataclass
class GlobalGrid:
    """ref: src/shared.jl:58-81."""
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
    neighbors: Tuple[Tuple[int, int, int], Tuple[int, int, int]]   # neighbors[n][dim]
    periods: Tuple[int, int, int]
    disp: int
    reorder: int
    comm: object
    cuda_enabled: bool
    quiet: bool
    device_id: int = 0
    ctx: Optional[int] = None           # igg_ctx*
    desc: Optional[L.GridDesc] = None


_grid: Optional[GlobalGrid] = None
_t0 = 0.0


def _torch():
    import torch
    return torch


def grid_is_initialized() -> bool:
    return _grid is not None


def check_initialized():
    if _grid is None:
        raise ArgumentError("No function of the module can be called before init_global_grid().")


def global_grid() -> GlobalGrid:
    if _grid is None:
        raise ArgumentError("No global grid has been created and activated yet, or the package has not been initialized.")
    return _grid


def get_global_grid() -> GlobalGrid:
    check_initialized()
    import copy
    g = copy.copy(_grid)
    g.ctx = None
    return g


class _Comm:
    """What `comm` is in this harness: the torch.distributed world (or a single process)."""

    def __init__(self, rank, size):
        self.rank, self.size = rank, size

    def allgather_bytes(self, b: bytes):
        if self.size == 1:
            return [b]
        dist = _torch().distributed
        out = [None] * self.size
        dist.all_gather_object(out, b)
        return out

    def barrier(self):
        if self.size > 1:
            _torch().distributed.barrier()


def _world():
    torch = _torch()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        return torch.distributed.get_rank(), torch.distributed.get_world_size()
    return 0, 1


def select_device(gg: Optional[GlobalGrid] = None) -> int:
    """Device = node-local rank (ref: src/select_device.jl:28-35)."""
    torch = _torch()
    g = gg or global_grid()
    if not g.cuda_enabled:
        raise ArgumentError("Cannot select a device because neither CUDA nor AMDGPU is enabled (meaning that the "
                            "corresponding module was not imported before ImplicitGlobalGrid).")
    me_l = int(os.environ.get("LOCAL_RANK", g.me))
    nl = int(os.environ.get("LOCAL_WORLD_SIZE", g.nprocs))
    if nl > torch.cuda.device_count():
        raise ArgumentError("More processes have been launched per node than there are GPUs available.")
    torch.cuda.set_device(me_l)
    return me_l


def init_global_grid(nx, ny=1, nz=1, *, dimx=0, dimy=0, dimz=0, periodx=0, periody=0, periodz=0,
                     origin=(0.0, 0.0, 0.0), origin_on_vertex=False, centerx=False, centery=False, centerz=False,
                     overlaps=(2, 2, 2), halowidths=None, disp=1, reorder=1, comm=None, init_MPI=True,
                     device_type="auto", select_device=True, quiet=False):
    """Returns (me, dims, nprocs, coords, comm) like the reference."""
    global _grid
    if _grid is not None:
        raise ArgumentError("init_global_grid() can only be called once before finalize_global_grid().")
    if device_type not in ("none", "auto", "CUDA", "AMDGPU"):
        raise ArgumentError(f"Argument `device_type`: invalid value obtained ({device_type}). Valid values are: CUDA, "
                            "AMDGPU, none, auto")
    if device_type in ("none", "AMDGPU"):
        raise ArgumentError("libigg_b200 is a CUDA-only implementation of the halo path: device_type must be \"auto\" "
                            "or \"CUDA\" (there is no CPU or AMDGPU path).")
    overlaps = tuple(int(o) for o in overlaps)
    if halowidths is None:
        halowidths = tuple(max(1, o // 2) for o in overlaps)
    halowidths = tuple(int(h) for h in halowidths)
    nxyz, dims, periods, origin = T.normalize_input(nx, ny, nz, dimx, dimy, dimz, periodx, periody, periodz, origin,
                                                    origin_on_vertex, centerx, centery, centerz, overlaps, halowidths,
                                                    disp)
    torch = _torch()
    me, nprocs = _world() if comm is None else (comm.rank, comm.size)
    comm = comm or _Comm(me, nprocs)
    dims = T.dims_create(nprocs, dims)
    coords = T.cart_coords(me, dims)
    left, right = [], []
    for i in range(NDIMS_MPI):
        l, r = T.cart_shift(coords, dims, periods, i, disp)
        left.append(l)
        right.append(r)
    nxyz_g = tuple(d * (n - o) + o * (p == 0) for d, n, o, p in zip(dims, nxyz, overlaps, periods))
    cuda_enabled = torch.cuda.is_available()
    g = GlobalGrid(nxyz_g, nxyz, dims, overlaps, halowidths, origin, origin_on_vertex, (centerx, centery, centerz),
                   nprocs, me, coords, (tuple(left), tuple(right)), periods, disp, reorder, comm, cuda_enabled, quiet)
    if not quiet and me == 0:
        print(f"Global grid: {nxyz_g[0]}x{nxyz_g[1]}x{nxyz_g[2]} (nprocs: {nprocs}, dims: {dims[0]}x{dims[1]}x{dims[2]}; "
              f"device support: {'CUDA (libigg_b200, NVLink peer stores)' if cuda_enabled else 'none'})")
    if not cuda_enabled:
        raise L.IGGError(L.IGG_ERR_CUDA, "No usable CUDA device; libigg_b200 has no CPU fallback.")
    g.device_id = globals()["select_device"](g) if select_device else torch.cuda.current_device()
    # C context + one-time rank bootstrap (replaces the MPI transport of the halo path)
    lib = L.load()
    g.desc = L.make_desc(nxyz, dims, coords, g.neighbors, periods, overlaps, halowidths, me, nprocs, g.device_id)
    ctx = C.c_void_p()
    L.check(lib.igg_create(C.byref(g.desc), C.byref(ctx)))
    g.ctx = ctx.value
    if nprocs > 1:
        try:
            n = lib.igg_bootstrap_size()
            blob = C.create_string_buffer(n)
            L.check(lib.igg_bootstrap_export(g.ctx, blob))
            blobs = comm.allgather_bytes(blob.raw)
            allb = C.create_string_buffer(b"".join(blobs), n * nprocs)
            L.check(lib.igg_bootstrap_import(g.ctx, allb, nprocs))
        except BaseException:
            lib.igg_destroy(g.ctx)          # a failed bootstrap must not leave a context (and its device memory) behind
            raise
    _grid = g
    tic()
    toc()
    return g.me, g.dims, g.nprocs, g.coords, g.comm


def finalize_global_grid(*, finalize_MPI=True):
    global _grid
    check_initialized()
    _call_cache.clear()
    _sym_fields.clear()          # igg_destroy releases the symmetric allocations
    if _grid.ctx:
        L.check(L.load().igg_destroy(_grid.ctx))
    _grid = None


# ---- fields --------------------------------------------------------------------------------------------------

def zeros(shape, dtype=None, device=None, pin_memory=False):
    """A zero-initialised column-major array of Julia size `shape` as a torch tensor (shape == `shape`)."""
    torch = _torch()
    shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list)) else (shape,)))
    dtype = dtype or torch.float64
    if device is None:
        device = "cpu" if pin_memory else f"cuda:{torch.cuda.current_device()}"
    t = torch.zeros(tuple(reversed(shape)), dtype=dtype, device=device, pin_memory=pin_memory)
    return t.permute(*reversed(range(len(shape))))


class _DevBlob:
    """Raw device memory as something torch.as_tensor understands (CUDA array interface)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


_sym_fields = {}


def alloc_field(shape, dtype=None):
    """A zero-initialised column-major array of Julia size `shape` inside a SYMMETRIC allocation (igg_alloc_field):
    collective -- every rank calls it with the same shape and dtype in the same order.  update_halo! then stores the
    y/z faces of such a field straight into the neighbours' halos over NVLink (direct-to-halo) instead of staging
    and unpacking them, for calls that stage 32 MiB or more per rank (IGG_DIRECT_THRESHOLD; smaller calls are faster
    through LL128 staging).  Release it with free_field (collective) before finalize_global_grid, or let finalize do it."""
    torch = _torch()
    g = global_grid()
    shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list)) else (shape,)))
    dtype = dtype or torch.float64
    el = torch.empty((), dtype=dtype).element_size()
    n = 1
    for s_ in shape:
        n *= s_
    ptr = C.c_void_p()
    L.check(L.load().igg_alloc_field(g.ctx, n * el, C.byref(ptr)))
    raw = torch.as_tensor(_DevBlob(ptr.value, n * el), device=f"cuda:{torch.cuda.current_device()}")
    t = raw.view(dtype).reshape(tuple(reversed(shape))).permute(*reversed(range(len(shape))))
    _sym_fields[t.data_ptr()] = ptr.value
    return t


def free_field(t):
    """Releases an array made by alloc_field (collective).  The tensor must not be used afterwards."""
    g = global_grid()
    ptr = _sym_fields.pop(t.data_ptr())
    L.check(L.load().igg_free_field(g.ctx, C.c_void_p(ptr)))


def from_numpy(a, device=None):
    """Column-major torch tensor holding the values of numpy array `a` (any memory order), same shape."""
    torch = _torch()
    import numpy as np
    dt = a.dtype
    view_as = {np.dtype(np.uint16): np.int16, np.dtype(np.uint32): np.int32, np.dtype(np.uint64): np.int64}
    b = np.ascontiguousarray(a.transpose(tuple(reversed(range(a.ndim)))))
    if dt in view_as:
        b = b.view(view_as[dt])
    t = torch.from_numpy(b)
    if device is None:
        device = f"cuda:{torch.cuda.current_device()}"
    return t.to(device).permute(*reversed(range(a.ndim)))


def to_numpy(t, dtype=None):
    """numpy array (order='F') with the values of column-major tensor `t`; `dtype` reinterprets the bits."""
    import numpy as np
    b = t.permute(*reversed(range(t.dim()))).contiguous().cpu().numpy()
    if dtype is not None:
        b = b.view(dtype)
    return np.asfortranarray(b.transpose(tuple(reversed(range(b.ndim)))))


def _is_column_major(t) -> bool:
    exp = 1
    for s, st in zip(t.shape, t.stride()):
        if s != 1 and st != exp:
            return False
        exp *= s
    return True


def _wrap(A):
    """extract / wrap_field (ref: src/shared.jl:246-256): tensor | (tensor, halowidths) | {'A':..,'halowidths':..}."""
    if isinstance(A, dict):
        return A["A"], tuple(int(h) for h in A["halowidths"])
    if isinstance(A, (tuple, list)) and len(A) == 2 and not hasattr(A, "shape"):
        return A[0], tuple(int(h) for h in A[1])
    return A, tuple(global_grid().halowidths)


def _c_fields(fields):
    out = []
    for i, (A, hw) in enumerate(fields):
        if not hasattr(A, "data_ptr"):
            raise ArgumentError(f"The field at position {i + 1} does not have a supported array type.")
        if A.dim() < 1 or A.dim() > 3:
            raise ArgumentError(f"The field at position {i + 1} does not have a supported array type.")
        if not _is_column_major(A):
            raise ArgumentError(f"The field at position {i + 1} is non-contiguous (in indexing)..")
        out.append(L.make_field(A.data_ptr(), tuple(A.shape), hw, A.element_size()))
    return out


_call_cache = {}


def _cell_of(a):
    """The CellArray inside an update_halo argument (bare, (A, halowidths) or {'A':.., 'halowidths':..}), or None."""
    x = a["A"] if isinstance(a, dict) else (a[0] if isinstance(a, (tuple, list)) and len(a) == 2 else a)
    return x if hasattr(x, "bitsarrays") else None


def _extract_all(A):
    """extract (ref: src/shared.jl:246-250): CellArray arguments become their bits arrays (B = 0: one field per cell
    component; B = 1: one field of whole cells), with the halowidths the argument carried."""
    if not any(_cell_of(a) is not None for a in A):
        return A
    out = []
    for a in A:
        c = _cell_of(a)
        if c is None:
            out.append(a)
        elif isinstance(a, dict):
            out.extend({"A": b, "halowidths": a["halowidths"]} for b in c.bitsarrays())
        elif isinstance(a, (tuple, list)):
            out.extend((b, a[1]) for b in c.bitsarrays())
        else:
            out.extend(c.bitsarrays())
    return tuple(out)


def _field_key(a):
    if isinstance(a, dict):
        t, hw = a["A"], tuple(a["halowidths"])
    elif isinstance(a, (tuple, list)) and len(a) == 2 and not hasattr(a, "shape"):
        t, hw = a[0], tuple(a[1])
    else:
        t, hw = a, None
    return (t.data_ptr(), t.shape, t.stride(), t.dtype, hw)


def update_halo(*A, dims=(3, 1, 2), stream=None, blocking=True):
    """update_halo!(A...; dims=(3,1,2))  -- ref: src/update_halo.jl:30-42.

    Extra keywords: `stream` (a torch.cuda.Stream, default: the current stream) and `blocking=False` for a
    stream-ordered call (the overlap API of SURVEY section 8f); the default blocks like the reference.
    """
    g = _grid
    if g is None:
        check_initialized()
    lib = L.load()
    A = _extract_all(A)
    try:
        key = (tuple(_field_key(a) for a in A), tuple(dims))
        prepared = _call_cache.get(key)
    except AttributeError:
        key, prepared = None, None
    if prepared is None:
        fields = [_wrap(a) for a in A]
        # checks that need type information stay on the host side (ref: src/update_halo.jl:465-479)
        diff = [i + 1 for i in range(1, len(fields))
                if fields[i][0].dtype != fields[0][0].dtype or fields[i][0].dim() != fields[0][0].dim()
                or fields[i][0].device.type != fields[0][0].device.type]
        cf = _c_fields(fields)
        arr = (L.Field * len(cf))(*cf)
        L.check(lib.igg_check_fields(C.byref(g.desc), len(cf), arr))
        if len(diff) > 1:
            raise ArgumentError("The fields at positions " + ", ".join(map(str, diff[:-1])) + f" and {diff[-1]} are of "
                                "different type than the first field; make sure that in a same call all fields are of "
                                "the same type.")
        if diff:
            raise ArgumentError(f"The field at position {diff[0]} is of different type than the first field; make sure "
                                "that in a same call all fields are of the same type.")
        d = (C.c_int * len(dims))(*[int(x) for x in dims])
        prepared = (arr, len(cf), d, len(dims))
        if key is not None:
            if len(_call_cache) > 256:
                _call_cache.clear()
            _call_cache[key] = prepared
    arr, nf, d, nd = prepared
    s = stream.cuda_stream if stream is not None else _torch().cuda.current_stream().cuda_stream
    rc = lib.igg_update_halo(g.ctx, nf, arr, d, nd, s, L.IGG_SYNC if blocking else L.IGG_ASYNC)
    if rc:
        L.check(rc)
    return None


def set_exchange_mode(mode):
    """Selects how update_halo moves data between ranks.  "ipc": the mode new contexts start in (one fused kernel, peer
    stores over NVLink; "auto" unless the environment variable IGG_MODE names another).  "auto": per call, "ll128"
    below IGG_BULK_THRESHOLD staged bytes (default 64 MiB), "fence" above.  "ll128": every 128-byte line of staged data
    carries its own arrival flag, so there is no release fence and no separate flag (latency regime).  "fence":
    16-byte accesses, one release fence and one arrival flag per tile of 8-128 KiB (bandwidth regime).  "nccl":
    comparison path (per-dimension pack kernel, grouped ncclSend/ncclRecv, unpack kernel).  Every rank must select
    the same mode."""
    g = global_grid()
    lib = L.load()
    if mode == "ipc":
        env = os.environ.get("IGG_MODE")
        mode = "fence" if env in ("fence", "ipc") else env if env in ("ll128", "auto") else "auto"
    if mode in ("fence", "ll128", "auto"):
        L.check(lib.igg_set_mode(g.ctx, {"fence": L.IGG_MODE_IPC_1PHASE, "ll128": L.IGG_MODE_IPC_LL128,
                                         "auto": L.IGG_MODE_IPC_AUTO}[mode]))
        return
    if mode != "nccl":
        raise ArgumentError(f"unknown exchange mode {mode!r} (use 'ipc', 'auto', 'fence', 'll128' or 'nccl')")
    if g.nprocs > 1:
        n = lib.igg_nccl_unique_id_size()
        buf = C.create_string_buffer(n)
        if g.me == 0:
            L.check(lib.igg_nccl_get_unique_id(buf))
        ids = g.comm.allgather_bytes(buf.raw)
        idb = C.create_string_buffer(ids[0], n)
        L.check(lib.igg_nccl_init(g.ctx, idb))
    L.check(lib.igg_set_mode(g.ctx, L.IGG_MODE_NCCL_3PHASE))


class PreparedCall:
    """A validated `update_halo!(A...)` kept for reuse (igg_prepare): calling it performs the update through one
    foreign call that skips argument checks, pointer resolution and plan lookup."""

    def __init__(self, A, dims, stream, blocking):
        g = global_grid()
        lib = L.load()
        A = _extract_all(A)
        fields = [_wrap(a) for a in A]
        diff = [i + 1 for i in range(1, len(fields))
                if fields[i][0].dtype != fields[0][0].dtype or fields[i][0].dim() != fields[0][0].dim()]
        if diff:
            raise ArgumentError(f"The field at position {diff[0]} is of different type than the first field; make sure "
                                "that in a same call all fields are of the same type.")
        cf = _c_fields(fields)
        arr = (L.Field * len(cf))(*cf)
        d = (C.c_int * len(dims))(*[int(x) for x in dims])
        h = C.c_void_p()
        L.check(lib.igg_prepare(g.ctx, len(cf), arr, d, len(dims), C.byref(h)))
        self._h, self._lib, self._fn = h, lib, lib.igg_update_halo_prepared
        self._keep = (A, arr, d)                                  # the tensors must stay alive and in place
        self._s = C.c_void_p(stream.cuda_stream if stream is not None else _torch().cuda.current_stream().cuda_stream)
        self._flags = L.IGG_SYNC if blocking else L.IGG_ASYNC
        self._ctx = g.ctx
        import functools
        self._go = functools.partial(self._fn, self._h, self._s, self._flags)      # one foreign call, arguments prebuilt

    def __call__(self):
        if self._go():
            L.check(self._fn(self._h, self._s, self._flags) or 0)

    def release(self):
        if self._h is not None:
            self._lib.igg_release_prepared(self._h)
            self._h = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def prepare_update_halo(*A, dims=(3, 1, 2), stream=None, blocking=True, warm=True):
    """Validates the arguments once and returns a zero-argument callable that performs `update_halo!(A...)` with
    minimal host overhead (igg_prepare / igg_update_halo_prepared).  For time loops that update the same arrays every
    step.  The arrays must stay alive and in place; `stream` defaults to the stream current at preparation time.
    With `warm` (default) one host-synchronous update is performed right away, like a first `update_halo!` call: it
    is collective, and it makes the context grow its staging memory now rather than inside the time loop or a CUDA
    graph capture (updating twice gives the same halos as updating once).  `warm=False` only prepares."""
    call = PreparedCall(A, dims, stream, blocking)
    if warm:
        rc = call._fn(call._h, call._s, L.IGG_SYNC)
        if rc:
            L.check(rc)
    return call


def gather(A, A_global, *, root=0, stream=None):
    """gather!(A, A_global; root=0)  -- ref: src/gather.jl:18-54.  `A_global` may be None on non-root ranks."""
    g = global_grid()
    torch = _torch()
    lib = L.load()
    if not _is_column_major(A) or (A_global is not None and not _is_column_major(A_global)):
        raise ArgumentError("gather!: arrays must be contiguous column-major arrays.")
    if A_global is not None and A_global.dtype != A.dtype:
        raise ArgumentError("gather!: `A` and `A_global` must have the same element type.")
    sA = (C.c_int64 * 3)(*(tuple(A.shape) + (1,) * (3 - A.dim())))
    if A_global is not None:
        sG = (C.c_int64 * 3)(*(tuple(A_global.shape) + (1,) * (3 - A_global.dim())))
        gp, gnd = C.c_void_p(A_global.data_ptr()), A_global.dim()
    else:
        sG, gp, gnd = None, None, 0
    s = stream if stream is not None else torch.cuda.current_stream()
    L.check(lib.igg_gather(g.ctx, C.c_void_p(A.data_ptr()), sA, A.dim(), gp, sG, gnd, A.element_size(), int(root),
                           C.c_void_p(s.cuda_stream)))
    return None


# ---- tools (ref: src/tools.jl) ---------------------------------------------------------------------------------

def _size(A, d):
    shp = tuple(A.shape) if hasattr(A, "shape") else tuple(A)
    return shp[d] if d < len(shp) else 1


def _n_g(d, nA=None):
    g = global_grid()
    base = g.nxyz_g[d]
    if nA is None:
        return base
    diff = nA - g.nxyz[d]
    p = 1 if g.periods[d] else 0
    return base + diff * (1 - p) + min(diff + g.overlaps[d], 0) * p


def nx_g(A=None): return _n_g(0, None if A is None else (A if isinstance(A, int) else _size(A, 0)))
def ny_g(A=None): return _n_g(1, None if A is None else (A if isinstance(A, int) else _size(A, 1)))
def nz_g(A=None): return _n_g(2, None if A is None else (A if isinstance(A, int) else _size(A, 2)))


def _i_g(d, i, nA, wrap_periodic=True, coord=None):
    g = global_grid()
    coord = g.coords[d] if coord is None else coord
    diff = nA - g.nxyz[d]
    olA = g.overlaps[d] + diff
    nga = _n_g(d, nA)
    i0 = -int(olA / 2) if g.periods[d] else 0
    ii = coord * (g.nxyz[d] - g.overlaps[d]) + i + i0
    if g.periods[d] and wrap_periodic:
        if ii > nga:
            ii -= nga
        if ii < 1:
            ii += nga
    return ii


def ix_g(ix, A, wrap_periodic=True): return _i_g(0, ix, _size(A, 0), wrap_periodic)
def iy_g(iy, A, wrap_periodic=True): return _i_g(1, iy, _size(A, 1), wrap_periodic)
def iz_g(iz, A, wrap_periodic=True): return _i_g(2, iz, _size(A, 2), wrap_periodic)


def _x_g(d, i, dx, nA, wrap_periodic=True):
    g = global_grid()
    ig = _i_g(d, i, nA, wrap_periodic)
    diff = nA - g.nxyz[d]
    ng = g.dims[d] * (g.nxyz[d] - g.overlaps[d]) + g.overlaps[d] * (0 if g.periods[d] else 1)
    l = dx * ng if g.origin_on_vertex else dx * (ng - 1)
    lA = l + dx * diff
    x1 = g.origin[d] + (-l * 0.5 if g.centerxyz[d] else 0.0) + dx * (
        (0.5 if g.origin_on_vertex else 0.0) + (-diff * 0.5 if not g.periods[d] else 0.0)
        + (-0.5 if (g.periods[d] and diff % 2 != 0) else 0.0))
    npts = int(lA / dx + 1e-9) + 1
    if ig > npts:
        return x1 + (npts - 1) * dx + dx * (ig - npts)
    if ig < 1:
        return x1 - dx * (1 - ig)
    return x1 + (ig - 1) * dx


def x_g(ix, dx, A, wrap_periodic=True): return _x_g(0, ix, dx, _size(A, 0), wrap_periodic)
def y_g(iy, dy, A, wrap_periodic=True): return _x_g(1, iy, dy, _size(A, 1), wrap_periodic)
def z_g(iz, dz, A, wrap_periodic=True): return _x_g(2, iz, dz, _size(A, 2), wrap_periodic)


def tic():
    """ref: src/tools.jl:892-898 (MPI.Barrier + time())."""
    global _t0
    check_initialized()
    if _grid.cuda_enabled:
        _torch().cuda.synchronize()
    _grid.comm.barrier()
    _t0 = time.time()


def toc():
    check_initialized()
    if _grid.cuda_enabled:
        _torch().cuda.synchronize()
    _grid.comm.barrier()
    return time.time() - _t0
