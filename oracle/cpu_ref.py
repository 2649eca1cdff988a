"""ctypes wrapper of oracle/libigg_cpu_ref.so (C restatement of the reference CPU path) -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use this module."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libigg_cpu_ref.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            import subprocess
            subprocess.run(["make", "-s", "-C", _HERE], check=True)
        _lib = C.CDLL(_PATH)
        _lib.igg_cpu_update_halo.restype = C.c_int
        _lib.igg_cpu_update_halo.argtypes = [C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                             C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                             C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int,
                                             C.POINTER(C.c_double)]
        _lib.igg_cpu_ref_max_threads.restype = C.c_int
    return _lib


def max_threads():
    return load().igg_cpu_ref_max_threads()


def update_halo(dims, periods, nxyz, overlaps, fields_per_rank, hws, dims_order=(3, 1, 2), copy_threads=1, iters=1):
    """fields_per_rank[r][i]: numpy arrays, order='F', updated in place.  hws[i]: halowidths of field i.
    Returns the elapsed seconds of the `iters` calls (all ranks, wall clock)."""
    lib = load()
    nranks = len(fields_per_rank)
    nf = len(fields_per_rank[0])
    a3 = lambda v: (C.c_int64 * 3)(*[int(x) for x in v])
    ptrs = (C.c_void_p * (nranks * nf))()
    for r, fr in enumerate(fields_per_rank):
        for i, A in enumerate(fr):
            assert A.flags.f_contiguous and A.dtype.itemsize == fields_per_rank[0][0].dtype.itemsize
            ptrs[r * nf + i] = A.ctypes.data
    sizes = (C.c_int64 * (3 * nf))()
    hw = (C.c_int64 * (3 * nf))()
    for i, A in enumerate(fields_per_rank[0]):
        shp = tuple(A.shape) + (1,) * (3 - A.ndim)
        for d in range(3):
            sizes[3 * i + d] = shp[d]
            hw[3 * i + d] = int(hws[i][d])
    order = (C.c_int * len(dims_order))(*dims_order)
    sec = C.c_double(0.0)
    rc = lib.igg_cpu_update_halo(nranks, a3(dims), a3(periods), a3(nxyz), a3(overlaps), nf, ptrs, sizes, hw,
                                 fields_per_rank[0][0].dtype.itemsize, order, len(dims_order), int(copy_threads),
                                 int(iters), C.byref(sec))
    if rc:
        raise RuntimeError("igg_cpu_update_halo failed")
    return sec.value
