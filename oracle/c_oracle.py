"""ctypes binding of oracle/liboracle.so (the C/OpenMP twin of nds_oracle.py).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by the product package.
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "nds_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE] + (["-B"] if force else []))
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        i64, i32, vp = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p
        L.oracle_nds_rank_f32.argtypes = [vp, i64, i32, vp, i32]
        L.oracle_nds_rank_f32.restype = i32
        L.oracle_dominate_relation_f32.argtypes = [vp, i64, vp, i64, i32, vp]
        L.oracle_dominate_relation_f32.restype = i32
        L.oracle_crowding_f32.argtypes = [vp, vp, i64, i32, vp]
        L.oracle_crowding_f32.restype = i32
        L.oracle_nsga2_select_f32.argtypes = [vp, i64, i32, i64, vp, vp, vp, i32]
        L.oracle_nsga2_select_f32.restype = i32
        L.oracle_num_threads.restype = i32
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def non_dominated_sort(x, nthreads: int = 0, return_fronts: bool = False):
    x = _f32(x)
    n, m = x.shape
    rank = np.zeros(n, dtype=np.int32)
    rc = lib().oracle_nds_rank_f32(x.ctypes.data, n, m, rank.ctypes.data, nthreads)
    if rc < 0:
        raise RuntimeError(f"oracle_nds_rank_f32 failed: {rc}")
    return (rank, rc) if return_fronts else rank


def dominate_relation(x, y):
    x, y = _f32(x), _f32(y)
    out = np.zeros((x.shape[0], y.shape[0]), dtype=np.uint8)
    lib().oracle_dominate_relation_f32(x.ctypes.data, x.shape[0], y.ctypes.data, y.shape[0], x.shape[1], out.ctypes.data)
    return out.astype(np.bool_)


def crowding_distance(costs, mask=None):
    costs = _f32(costs)
    n, m = costs.shape
    dist = np.zeros(n, dtype=np.float32)
    mp = None
    if mask is not None:
        mask = np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
        mp = mask.ctypes.data
    rc = lib().oracle_crowding_f32(costs.ctypes.data, mp, n, m, dist.ctypes.data)
    if rc < 0:
        raise RuntimeError(f"oracle_crowding_f32 failed: {rc}")
    return dist


def nsga2_select(fit, pop: int, nthreads: int = 0):
    fit = _f32(fit)
    n, m = fit.shape
    order = np.zeros(pop, dtype=np.int32)
    rank = np.zeros(n, dtype=np.int32)
    crowd = np.zeros(n, dtype=np.float32)
    rc = lib().oracle_nsga2_select_f32(fit.ctypes.data, n, m, pop, order.ctypes.data, rank.ctypes.data, crowd.ctypes.data, nthreads)
    if rc < 0:
        raise RuntimeError(f"oracle_nsga2_select_f32 failed: {rc}")
    return order, rank, crowd
