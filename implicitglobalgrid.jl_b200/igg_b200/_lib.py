"""ctypes binding of libigg_b200.so (the C ABI declared in include/igg_b200.h).

This is the Python twin of the `ccall` stubs a Julia host would use.  The library is looked
up in-tree next to the package (`implicitglobalgrid.jl_b200/libigg_b200.so`); it is never pip-installed.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libigg_b200.so")

IGG_OK = 0
IGG_ERR_ARG, IGG_ERR_CUDA, IGG_ERR_STATE, IGG_ERR_FIELDS, IGG_ERR_IPC = -1, -2, -3, -4, -5
IGG_ERR_NCCL, IGG_ERR_TIMEOUT, IGG_ERR_UNSUPPORTED, IGG_ERR_GRID = -6, -7, -8, -9
IGG_SYNC, IGG_ASYNC = 0, 1
IGG_MODE_IPC_1PHASE, IGG_MODE_NCCL_3PHASE, IGG_MODE_IPC_LL128, IGG_MODE_IPC_AUTO = 0, 1, 2, 3
IGG_NOFFSETS = 27
IGG_OP_PUSH, IGG_OP_LOCAL, IGG_OP_UNPACK = 0, 1, 2
IGG_PROC_NULL = -1
IGG_MAX_FIELDS = 32


class GridDesc(C.Structure):
    _fields_ = [
        ("nxyz", C.c_int64 * 3), ("dims", C.c_int64 * 3), ("coords", C.c_int64 * 3),
        ("neighbors", (C.c_int64 * 2) * 3), ("periods", C.c_int64 * 3), ("overlaps", C.c_int64 * 3),
        ("halowidths", C.c_int64 * 3), ("me", C.c_int32), ("nprocs", C.c_int32), ("device", C.c_int32),
        ("reserved", C.c_int32),
    ]


class Field(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("size", C.c_int64 * 3), ("halowidths", C.c_int64 * 3),
                ("elsize", C.c_int32), ("ndims", C.c_int32)]


class PlanOp(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("field", C.c_int32), ("offset", C.c_int32 * 3), ("peer", C.c_int32),
        ("src_lo", C.c_int64 * 3), ("src_hi", C.c_int64 * 3), ("dst_lo", C.c_int64 * 3), ("dst_hi", C.c_int64 * 3),
        ("staging_off", C.c_int64), ("nbytes", C.c_int64),
        ("src_off", C.c_int64), ("dst_off", C.c_int64), ("src_sy", C.c_int64), ("src_sz", C.c_int64),
        ("dst_sy", C.c_int64), ("dst_sz", C.c_int64), ("row_bytes", C.c_int64), ("ny", C.c_int64), ("nz", C.c_int64),
        ("vec_bytes", C.c_int32), ("ntiles", C.c_int32), ("flag_base", C.c_int32), ("reserved", C.c_int32),
    ]


class IGGError(RuntimeError):
    """Raised where the reference raises ErrorException (`error("...")`); carries the C status code."""

    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


_lib = None


def load():
    """Loads the shared library (once).  Fails loudly if it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python implicitglobalgrid.jl_b200/build.py` "
                          "(libigg_b200 has no Python or CPU fallback).")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, i64p, i32p = C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int)
    gd, fp = C.POINTER(GridDesc), C.POINTER(Field)
    sigs = {
        "igg_version": (C.c_int, []),
        "igg_last_error": (C.c_char_p, []),
        "igg_create": (C.c_int, [gd, C.POINTER(vp)]),
        "igg_destroy": (C.c_int, [vp]),
        "igg_bootstrap_size": (C.c_size_t, []),
        "igg_bootstrap_export": (C.c_int, [vp, vp]),
        "igg_bootstrap_import": (C.c_int, [vp, vp, i32]),
        "igg_update_halo": (C.c_int, [vp, i32, fp, i32p, i32, vp, i32]),
        "igg_prepare": (C.c_int, [vp, i32, fp, i32p, i32, C.POINTER(vp)]),
        "igg_update_halo_prepared": (C.c_int, [vp, vp, i32]),
        "igg_release_prepared": (C.c_int, [vp]),
        "igg_alloc_field": (C.c_int, [vp, C.c_size_t, C.POINTER(vp)]),
        "igg_free_field": (C.c_int, [vp, vp]),
        "igg_update_halo_noalloc": (C.c_int, [gd, i32, fp, i32p, i32, vp, i32]),
        "igg_noalloc_scratch_bytes": (C.c_int64, []),
        "igg_set_mode": (C.c_int, [vp, i32]),
        "igg_nccl_unique_id_size": (C.c_size_t, []),
        "igg_nccl_get_unique_id": (C.c_int, [vp]),
        "igg_nccl_init": (C.c_int, [vp, vp]),
        "igg_gather": (C.c_int, [vp, vp, i64p, i32, vp, i64p, i32, i32, i32, vp]),
        "igg_diffusion3d_step": (C.c_int, [vp, vp, vp, i64p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                           i32, vp]),
        "igg_sendranges": (C.c_int, [gd, fp, i32, i32, i64p]),
        "igg_recvranges": (C.c_int, [gd, fp, i32, i32, i64p]),
        "igg_check_fields": (C.c_int, [gd, i32, fp]),
        "igg_plan_describe": (C.c_int, [gd, i32, fp, i32p, i32, C.POINTER(PlanOp), i32, i32p, i64p]),
        "igg_plan_tiles": (C.c_int, [gd, i32, fp, i32p, i32, i32, vp, i32, i32p, i64p, i64p]),
        "igg_ctrl_selftest_create": (C.c_int, [i32, i32, C.c_char_p, C.POINTER(vp)]),
        "igg_ctrl_selftest_run": (C.c_int, [vp, i32, i32, C.c_char_p, i32]),
        "igg_debug_state": (C.c_int, [vp, C.POINTER(C.c_uint64)]),
        "igg_debug_issue_loop": (C.c_int, [vp, vp, i32, C.POINTER(C.c_double)]),
        "igg_debug_peer_copy": (C.c_int, [vp, vp, i32, C.c_size_t, i32, C.POINTER(C.c_double)]),
        "igg_debug_timed_calls": (C.c_int, [C.POINTER(vp), i32, vp, i32, i32, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
        "igg_trace_enable": (C.c_int, [vp, i32]),
        "igg_trace_read": (C.c_int, [vp, C.POINTER(C.c_uint64), i32, i32p]),
        "igg_launch_count": (C.c_int64, [vp]),
        "igg_kernel_name": (C.c_char_p, []),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


DECLARED_SYMBOLS = [
    "igg_version", "igg_last_error", "igg_create", "igg_destroy", "igg_bootstrap_size", "igg_bootstrap_export",
    "igg_bootstrap_import", "igg_update_halo", "igg_prepare", "igg_update_halo_prepared", "igg_release_prepared", "igg_alloc_field", "igg_free_field", "igg_update_halo_noalloc", "igg_noalloc_scratch_bytes", "igg_set_mode", "igg_nccl_unique_id_size", "igg_nccl_get_unique_id",
    "igg_nccl_init", "igg_gather", "igg_diffusion3d_step", "igg_sendranges", "igg_recvranges", "igg_check_fields", "igg_plan_describe", "igg_plan_tiles",
    "igg_ctrl_selftest_create", "igg_ctrl_selftest_run", "igg_debug_state", "igg_debug_issue_loop", "igg_debug_peer_copy", "igg_debug_timed_calls", "igg_trace_enable", "igg_trace_read",
    "igg_launch_count", "igg_kernel_name",
]


def check(rc: int):
    if rc != IGG_OK:
        raise IGGError(rc, load().igg_last_error().decode("utf-8", "replace"))


def make_desc(nxyz, dims, coords, neighbors, periods, overlaps, halowidths, me, nprocs, device) -> GridDesc:
    """neighbors[n][d] as in the oracle / reference (n = side, d = dimension) -> C layout [d][n]."""
    d = GridDesc()
    for i in range(3):
        d.nxyz[i], d.dims[i], d.coords[i] = int(nxyz[i]), int(dims[i]), int(coords[i])
        d.periods[i], d.overlaps[i], d.halowidths[i] = int(periods[i]), int(overlaps[i]), int(halowidths[i])
        for n in range(2):
            v = int(neighbors[n][i])
            d.neighbors[i][n] = v if v >= 0 else IGG_PROC_NULL
    d.me, d.nprocs, d.device, d.reserved = int(me), int(nprocs), int(device), 0
    return d


def make_field(ptr: int, shape, halowidths, elsize: int) -> Field:
    f = Field()
    f.ptr = ptr
    nd = len(shape)
    for i in range(3):
        f.size[i] = int(shape[i]) if i < nd else 1
        f.halowidths[i] = int(halowidths[i])
    f.elsize, f.ndims = int(elsize), nd
    return f


def plan_describe(desc: GridDesc, fields, dims=None):
    """Returns (list[PlanOp], staging_bytes) for this rank (host only, no GPU needed)."""
    lib = load()
    arr = (Field * len(fields))(*fields)
    n = C.c_int(0)
    stg = C.c_int64(0)
    dptr, nd = (None, 0) if dims is None else ((C.c_int * len(dims))(*dims), len(dims))
    check(lib.igg_plan_describe(C.byref(desc), len(fields), arr, dptr, nd, None, 0, C.byref(n), C.byref(stg)))
    ops = (PlanOp * max(1, n.value))()
    check(lib.igg_plan_describe(C.byref(desc), len(fields), arr, dptr, nd, ops, n.value, C.byref(n), C.byref(stg)))
    return [ops[i] for i in range(n.value)], stg.value


class TileRec(C.Structure):
    """Mirror of struct TileRec (csrc/igg_internal.h), 80 bytes; what one thread block iteration works on."""
    _fields_ = [("src_off", C.c_int64), ("dst_off", C.c_int64), ("src_sy", C.c_int64), ("src_sz", C.c_int64),
                ("dst_sy", C.c_int64), ("dst_sz", C.c_int64), ("row0", C.c_uint32), ("nrows", C.c_uint32),
                ("ncols", C.c_uint32), ("ny", C.c_uint32), ("field", C.c_uint16), ("vec_log2", C.c_uint8),
                ("kind", C.c_uint8), ("nbr", C.c_uint8), ("bulk", C.c_uint8), ("pad", C.c_uint8 * 2), ("flag_idx", C.c_uint32),
                ("pad2", C.c_uint32)]


def plan_tiles(desc: GridDesc, fields, dims=None, mode=IGG_MODE_IPC_1PHASE, regions=False):
    """Returns (list[TileRec], staging_bytes): the tile table of the exchange kernel of `mode` (host only).
    With regions=True a third value: dict of four 27-element lists (byte_pack, byte_need, flag_pack, flag_need)."""
    assert C.sizeof(TileRec) == 80
    lib = load()
    arr = (Field * len(fields))(*fields)
    n = C.c_int(0)
    stg = C.c_int64(0)
    reg = (C.c_int64 * (4 * IGG_NOFFSETS))()
    dptr, nd = (None, 0) if dims is None else ((C.c_int * len(dims))(*dims), len(dims))
    check(lib.igg_plan_tiles(C.byref(desc), len(fields), arr, dptr, nd, mode, None, 0, C.byref(n), C.byref(stg), None))
    tiles = (TileRec * max(1, n.value))()
    check(lib.igg_plan_tiles(C.byref(desc), len(fields), arr, dptr, nd, mode, tiles, n.value, C.byref(n), C.byref(stg), reg))
    out = [tiles[i] for i in range(n.value)]
    if not regions:
        return out, stg.value
    r = list(reg)
    k = IGG_NOFFSETS
    return out, stg.value, dict(byte_pack=r[:k], byte_need=r[k:2 * k], flag_pack=r[2 * k:3 * k], flag_need=r[3 * k:])
