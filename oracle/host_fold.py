"""Fixed rank-order host reduction (TEST INFRASTRUCTURE).

Restates `/root/reference/mnist/main.py:126-127`:
    distributed.all_reduce(p.grad.data, op=distributed.reduce_op.SUM)
    p.grad.data /= float(world_size)
with the add order pinned to ((g0+g1)+g2)+... and IEEE binary32 division, in numpy and in C
(`fold.c`).  The two must agree bit for bit with each other and with the CUDA kernels.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_fold.so")
_lib = None


def build_c_oracle(force=False):
    """Compile fold.c with the committed Makefile (gcc only, no GPU)."""
    src = os.path.join(_HERE, "fold.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle_fold.so"])
    return _SO


def _c():
    global _lib
    if _lib is None:
        build_c_oracle()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_fold_average_f32.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                                 ctypes.c_size_t, ctypes.c_void_p]
        _lib.oracle_fold_average_f32.restype = None
        _lib.oracle_fold_sum_i32.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                             ctypes.c_size_t, ctypes.c_void_p]
        _lib.oracle_fold_sum_i32.restype = None
        _lib.oracle_fold_wide_average_f32.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                                      ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int]
        _lib.oracle_fold_wide_average_f32.restype = None
    return _lib


def fold_average(per_rank):
    """numpy: per_rank = list (rank order) of equal-shape float32 arrays -> averaged array."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in per_rank]
    acc = arrs[0].copy()
    for a in arrs[1:]:
        acc = acc + a                      # elementwise IEEE binary32 add, rank order
    return acc / np.float32(float(len(arrs)))   # mnist/main.py:127 divides by float(world_size)


def _ptr_array(arrs):
    return (ctypes.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def fold_average_c(per_rank):
    """Same as fold_average through fold.c (used for large sizes and as the cross-check)."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in per_rank]
    out = np.empty_like(arrs[0])
    _c().oracle_fold_average_f32(_ptr_array(arrs), len(arrs), arrs[0].size, out.ctypes.data)
    return out


def fold_sum_i32(per_rank):
    """toy/main.py:19-20: int32 SUM, no division (numpy, wrap-around)."""
    arrs = [np.ascontiguousarray(a, dtype=np.int32) for a in per_rank]
    acc = arrs[0].astype(np.uint32)
    for a in arrs[1:]:
        acc = acc + a.astype(np.uint32)
    return acc.astype(np.int32)


def fold_sum_i32_c(per_rank):
    arrs = [np.ascontiguousarray(a, dtype=np.int32) for a in per_rank]
    out = np.empty_like(arrs[0])
    _c().oracle_fold_sum_i32(_ptr_array(arrs), len(arrs), arrs[0].size, out.ctypes.data)
    return out


def fold_wide_average(per_rank, divide=True):
    """numpy oracle of the opt-in NVSwitch variant: binary64 sum over the ranks rounded once to
    binary32, then the binary32 division of mnist/main.py:127 (see fold.c)."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in per_rank]
    acc = arrs[0].astype(np.float64)
    for a in arrs[1:]:
        acc = acc + a.astype(np.float64)
    s = acc.astype(np.float32)
    return s / np.float32(float(len(arrs))) if divide else s


def fold_wide_average_c(per_rank, divide=True):
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in per_rank]
    out = np.empty_like(arrs[0])
    _c().oracle_fold_wide_average_f32(_ptr_array(arrs), len(arrs), arrs[0].size, out.ctypes.data, 1 if divide else 0)
    return out
