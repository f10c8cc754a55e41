"""ctypes binding of libevoxb.so -- the C ABI declared in include/evoxb.h.

There is no fallback: if the shared library is missing or a call fails, an exception is
raised.  The library is never built implicitly on import (``__graft_entry__.build()`` or
``python evox-myc_b200/_build.py`` does that), so a GPU box without the .so fails loudly.
"""

from __future__ import annotations

import ctypes
import os

from ._build import LIB_PATH

_lib = None

# every symbol include/evoxb.h declares (tests check that the .so exports all of them)
EXPORTS = [
    "evoxb_version",
    "evoxb_last_error",
    "evoxb_dominate_relation_f32",
    "evoxb_nds_workspace_bytes",
    "evoxb_nds_rank_f32",
    "evoxb_crowding_workspace_bytes",
    "evoxb_crowding_f32",
    "evoxb_crowding_noalloc_f32",
    "evoxb_nsga2_select_workspace_bytes",
    "evoxb_nsga2_select_f32",
    "evoxb_nsga3_select_workspace_bytes",
    "evoxb_nsga3_select_f32",
    "evoxb_kth_rank_workspace_bytes",
    "evoxb_kth_rank_i32",
    "evoxb_non_dominate_select_f32",
    "evoxb_argsort_workspace_bytes",
    "evoxb_argsort_f32",
    "evoxb_gather_rows_f32",
    "evoxb_sbx_pm_clip_f32",
    "evoxb_dtlz2_f32",
    "evoxb_lsmop1_f32",
    "evoxb_comm_unique_id",
    "evoxb_comm_create",
    "evoxb_comm_destroy",
    "evoxb_comm_status",
    "evoxb_comm_poll_error",
    "evoxb_comm_set_timeout_ms",
    "evoxb_comm_last_path",
    "evoxb_nds_build_path",
    "evoxb_shard_plan",
    "evoxb_flow_plan",
    "evoxb_nds_sharded_workspace_bytes",
    "evoxb_nds_rank_sharded_f32",
    "evoxb_nsga2_select_sharded_f32",
    "evoxb_profile_enable",
    "evoxb_profile_read",
    "evoxb_launch_count",
]


class EvoxbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"evoxb error {code}: {msg}")
        self.code = code


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU or PyTorch fallback for this path)"
            )
        L = ctypes.CDLL(LIB_PATH)
        i64, i32, vp, sz = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t
        L.evoxb_version.restype = i32
        L.evoxb_last_error.restype = ctypes.c_char_p
        L.evoxb_dominate_relation_f32.argtypes = [vp, i64, vp, i64, i32, vp, vp]
        L.evoxb_nds_workspace_bytes.argtypes = [i64, i32, ctypes.POINTER(sz)]
        L.evoxb_nds_rank_f32.argtypes = [vp, i64, i32, vp, vp, sz, vp]
        L.evoxb_crowding_workspace_bytes.argtypes = [i64, i32, ctypes.POINTER(sz)]
        L.evoxb_crowding_f32.argtypes = [vp, vp, i64, i32, vp, vp, sz, vp]
        L.evoxb_crowding_noalloc_f32.argtypes = [vp, vp, i64, i32, vp, vp]
        L.evoxb_nsga2_select_workspace_bytes.argtypes = [i64, i32, ctypes.POINTER(sz)]
        L.evoxb_nsga2_select_f32.argtypes = [vp, i64, i32, i64, vp, vp, vp, vp, sz, vp]
        L.evoxb_nsga3_select_workspace_bytes.argtypes = [i64, i32, i64, ctypes.POINTER(sz)]
        L.evoxb_nsga3_select_f32.argtypes = [vp, i64, i32, i64, vp, i64, ctypes.POINTER(ctypes.c_uint32), vp,
                                             ctypes.POINTER(ctypes.c_uint32), vp, vp, vp, vp, vp, sz, vp]
        L.evoxb_kth_rank_workspace_bytes.argtypes = [i64, ctypes.POINTER(sz)]
        L.evoxb_kth_rank_i32.argtypes = [vp, i64, i64, vp, vp, sz, vp]
        L.evoxb_non_dominate_select_f32.argtypes = [vp, i64, i32, i64, vp, vp, sz, vp]
        L.evoxb_argsort_workspace_bytes.argtypes = [i64, ctypes.POINTER(sz)]
        L.evoxb_argsort_f32.argtypes = [vp, i64, vp, vp, sz, vp]
        L.evoxb_gather_rows_f32.argtypes = [vp, i64, i64, vp, i64, vp, vp]
        f32, u32 = ctypes.c_float, ctypes.c_uint32
        L.evoxb_sbx_pm_clip_f32.argtypes = [vp, i64, i32, vp, vp, f32, f32, f32, f32, u32, u32, vp, vp]
        L.evoxb_dtlz2_f32.argtypes = [vp, i64, i32, i32, vp, vp]
        L.evoxb_lsmop1_f32.argtypes = [vp, i64, i32, i32, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32), vp, vp]
        ip = ctypes.POINTER(ctypes.c_int)
        L.evoxb_comm_unique_id.argtypes = [vp]
        L.evoxb_comm_create.argtypes = [ctypes.c_char_p, i32, i32, ctypes.POINTER(vp)]
        L.evoxb_comm_destroy.argtypes = [vp]
        L.evoxb_comm_status.argtypes = [vp, ip, ip, ip]
        L.evoxb_comm_poll_error.argtypes = [vp]
        L.evoxb_comm_set_timeout_ms.argtypes = [vp, i64]
        L.evoxb_comm_last_path.argtypes = [vp]
        L.evoxb_nds_build_path.argtypes = [vp, i64, i32, vp, ip]
        L.evoxb_shard_plan.argtypes = [i64, i32, i32, ip, ip, ip, ip]
        L.evoxb_flow_plan.argtypes = [i64, i32, i32, i32, i32, ip, ip, ip, ctypes.POINTER(ctypes.c_int32), i32]
        L.evoxb_nds_sharded_workspace_bytes.argtypes = [i64, i32, i32, i32, ctypes.POINTER(sz)]
        L.evoxb_nds_rank_sharded_f32.argtypes = [vp, i64, i32, vp, vp, sz, vp, vp]
        L.evoxb_nsga2_select_sharded_f32.argtypes = [vp, i64, i32, i64, vp, vp, vp, vp, sz, vp, vp]
        L.evoxb_launch_count.restype = ctypes.c_ulonglong
        L.evoxb_profile_enable.argtypes = [i32]
        L.evoxb_profile_read.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_char_p), i32]
        for name in EXPORTS:
            getattr(L, name)  # raises AttributeError if a declared symbol is not exported
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise EvoxbError(rc, lib().evoxb_last_error().decode("utf-8", "replace"))


def launch_count() -> int:
    return int(lib().evoxb_launch_count())


def profile_enable(on: bool) -> None:
    lib().evoxb_profile_enable(1 if on else 0)


def profile_read() -> dict[str, float]:
    """Stage times (ms) of the last compute call made on this thread; synchronises its events."""
    cap = 24
    ms = (ctypes.c_float * cap)()
    names = (ctypes.c_char_p * cap)()
    n = lib().evoxb_profile_read(ms, names, cap)
    if n < 0:
        check(n)
    out: dict[str, float] = {}
    for i in range(min(n, cap)):
        key = names[i].decode()
        out[key] = out.get(key, 0.0) + float(ms[i])
    return out
