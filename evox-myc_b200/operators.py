"""Python mirror of the reference's operator interface for the environmental-selection path.

Same names, argument meaning and error behaviour as
``/root/reference/src/evox/operators/selection/non_dominate.py`` (exported at
``src/evox/operators/__init__.py:1-5``); every function body is a call through the C ABI of
``libevoxb.so`` (include/evoxb.h).  Inputs may be CUDA tensors (stay on the device) or host
arrays (NumPy / lists / host tensors: copied to the device and back inside the call).

Fitness values that are not float32 (the reference's monitors feed int32, tests/test_monitors.py:34)
are cast to float32 first, which is exact below 2**24.
"""

from __future__ import annotations

import ctypes

import torch

from . import _lib, _runtime as rt

_METHODS = ("auto", "scan", "full map-reduce", "host")


def _check_2d(x, name):
    if x.dev.dim() != 2:
        raise ValueError(f"{name} must have shape (n, m), got {tuple(x.dev.shape)}")
    if x.dev.shape[1] < 1:
        raise ValueError(f"{name} needs at least one objective")


def non_dominated_sort(x, method: str = "auto"):
    """Pareto rank by front peeling -- non_dominate.py:30-113.

    Parameters and result as in the reference: ``x`` is (n, m); returns int32 (n,), ranks start at 0.
    ``method`` is validated against the reference's four strings (:47-52); all of them name XLA
    lowering strategies for the same integers, so here it only is checked.
    """
    assert method in _METHODS, "method must be either 'auto', 'scan', or 'full map-reduce', 'host'"
    L = _lib.lib()
    a = rt.Arg(x, torch.float32, "nds_x")
    _check_2d(a, "x")
    n, m = a.dev.shape
    rank = torch.empty((n,), dtype=torch.int32, device=a.dev.device)
    if n > 0:
        need = ctypes.c_size_t()
        _lib.check(L.evoxb_nds_workspace_bytes(n, m, ctypes.byref(need)))
        ws = rt.workspace(need.value, a.dev.device)
        with torch.cuda.device(a.dev.device):
            _lib.check(
                L.evoxb_nds_rank_f32(a.dev.data_ptr(), n, m, rank.data_ptr(), ws.data_ptr(), ws.numel(),
                                     rt.stream_ptr(a.dev.device))
            )
    return rt.give_back(rank, a, "nds_rank")
