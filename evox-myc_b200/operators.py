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


def nds_build_path(n: int, m: int, device=None) -> str:
    """How the dominance matrix of the last ``non_dominated_sort`` call of this shape was built (no counterpart in
    the reference; tests and benchmark labels): ``"float"``, ``"levels"`` or ``"levels->float"`` (levels judged too
    coarse on the device, float kernel ran).  Reads the call's workspace, so ask right after the call."""
    L = _lib.lib()
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    need = ctypes.c_size_t()
    _lib.check(L.evoxb_nds_workspace_bytes(n, m, ctypes.byref(need)))
    ws = rt.workspace(need.value, device)
    path = ctypes.c_int(0)
    with torch.cuda.device(device):
        _lib.check(L.evoxb_nds_build_path(ws.data_ptr(), n, m, rt.stream_ptr(device), ctypes.byref(path)))
    return ("float", "levels", "levels->float")[path.value]


def _ws_call(size_fn, size_args, device):
    need = ctypes.c_size_t()
    _lib.check(size_fn(*size_args, ctypes.byref(need)))
    return rt.workspace(need.value, device)


def crowding_distance(costs, mask=None):
    """Crowding distance -- non_dominate.py:116-156.

    ``costs`` is (n, m); ``mask`` is None or a boolean (n,) array, True = the row takes part.
    Returns float32 (n,): ignored rows are ``-inf``, the two ends of every objective ``+inf``.
    """
    L = _lib.lib()
    a = rt.Arg(costs, torch.float32, "cd_x")
    _check_2d(a, "costs")
    n, m = a.dev.shape
    mk = None
    if mask is not None:
        mk = rt.Arg(mask, torch.uint8, "cd_mask", device=a.dev.device)
        if tuple(mk.dev.shape) != (n,):
            raise ValueError(f"mask must have shape ({n},), got {tuple(mk.dev.shape)}")
    dist = torch.empty((n,), dtype=torch.float32, device=a.dev.device)
    if n > 0:
        with torch.cuda.device(a.dev.device):
            ws = _ws_call(L.evoxb_crowding_workspace_bytes, (n, m), a.dev.device)
            _lib.check(
                L.evoxb_crowding_f32(a.dev.data_ptr(), mk.dev.data_ptr() if mk is not None else None, n, m,
                                     dist.data_ptr(), ws.data_ptr(), ws.numel(), rt.stream_ptr(a.dev.device))
            )
    return rt.give_back(dist, a, "cd_out")


def _argsort_f32(x_dev: torch.Tensor) -> torch.Tensor:
    L = _lib.lib()
    n = x_dev.shape[0]
    idx = torch.empty((n,), dtype=torch.int32, device=x_dev.device)
    if n > 0:
        with torch.cuda.device(x_dev.device):
            ws = _ws_call(L.evoxb_argsort_workspace_bytes, (n,), x_dev.device)
            _lib.check(L.evoxb_argsort_f32(x_dev.data_ptr(), n, idx.data_ptr(), ws.data_ptr(), ws.numel(),
                                           rt.stream_ptr(x_dev.device)))
    return idx


def crowding_distance_sort(x, mask=None):
    """Indices that sort ``x`` by crowding distance -- non_dominate.py:159-178 (``jnp.argsort(distance)``)."""
    a = rt.Arg(x, torch.float32, "cd_x")
    mk = None if mask is None else rt.Arg(mask, torch.uint8, "cd_mask", device=a.dev.device).dev
    dist = crowding_distance(a.dev, mk)
    return rt.give_back(_argsort_f32(dist), a, "cds_out")


def _gather_rows(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    L = _lib.lib()
    flat = src.reshape(src.shape[0], -1)
    if flat.dtype != torch.float32:
        return src[idx.long()]  # non-float payloads (never on the hot path) use plain indexing
    flat = flat.contiguous()
    out = torch.empty((idx.shape[0], flat.shape[1]), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        _lib.check(L.evoxb_gather_rows_f32(flat.data_ptr(), flat.shape[0], flat.shape[1], idx.data_ptr(), idx.shape[0],
                                           out.data_ptr(), rt.stream_ptr(src.device)))
    return out.reshape((idx.shape[0],) + tuple(src.shape[1:]))


def non_dominate(population, fitness, topk: int):
    """Select the ``topk`` individuals by Pareto rank with crowding as tie-breaker -- non_dominate.py:181-195.

    Returns ``(population[order], fitness[order])``.  Follows the reference to the letter, including the
    cut at ``rank[order[topk-1]]`` and the use of the argsort *indices* as secondary key (:190-194).
    """
    L = _lib.lib()
    f = rt.Arg(fitness, torch.float32, "sel_f")
    p = rt.Arg(population, torch.float32, "sel_p", device=f.dev.device)
    _check_2d(f, "fitness")
    n, m = f.dev.shape
    order = torch.empty((topk,), dtype=torch.int32, device=f.dev.device)
    with torch.cuda.device(f.dev.device):
        ws = _ws_call(L.evoxb_nsga2_select_workspace_bytes, (n, m), f.dev.device)
        _lib.check(L.evoxb_non_dominate_select_f32(f.dev.data_ptr(), n, m, topk, order.data_ptr(), ws.data_ptr(),
                                                   ws.numel(), rt.stream_ptr(f.dev.device)))
    return (rt.give_back(_gather_rows(p.dev, order), p, "sel_po"), rt.give_back(_gather_rows(f.dev, order), f, "sel_fo"))


class NonDominate:
    """Operator form of :func:`non_dominate` -- non_dominate.py:198-204."""

    def __init__(self, topk):
        self.topk = topk

    def __call__(self, population, fitness):
        return non_dominate(population, fitness, self.topk)


def nsga2_select(merged_fitness, pop_size: int, return_details: bool = False):
    """Survivor indices of ``NSGA2.tell`` -- src/evox/algorithms/mo/nsga2.py:90-96.

    ``combined_order = lexsort((-crowding_dis, rank))[:pop_size]`` with
    ``worst_rank = rank[argsort(rank)[pop_size]]`` (index ``pop_size``, as the reference has it).
    Returns int32 (pop_size,); with ``return_details`` also ``rank`` and ``crowding_dis``.
    """
    L = _lib.lib()
    f = rt.Arg(merged_fitness, torch.float32, "sel_f")
    _check_2d(f, "merged_fitness")
    n, m = f.dev.shape
    dev = f.dev.device
    order = torch.empty((pop_size,), dtype=torch.int32, device=dev)
    rank = torch.empty((n,), dtype=torch.int32, device=dev) if return_details else None
    crowd = torch.empty((n,), dtype=torch.float32, device=dev) if return_details else None
    with torch.cuda.device(dev):
        ws = _ws_call(L.evoxb_nsga2_select_workspace_bytes, (n, m), dev)
        _lib.check(
            L.evoxb_nsga2_select_f32(f.dev.data_ptr(), n, m, pop_size, order.data_ptr(),
                                     rank.data_ptr() if rank is not None else None,
                                     crowd.data_ptr() if crowd is not None else None,
                                     ws.data_ptr(), ws.numel(), rt.stream_ptr(dev))
        )
    if return_details:
        return (rt.give_back(order, f, "sel_o"), rt.give_back(rank, f, "sel_r"), rt.give_back(crowd, f, "sel_c"))
    return rt.give_back(order, f, "sel_o")


def nsga3_last_rank(merged_fitness, pop_size: int):
    """The part of ``NSGA3.tell`` that is on this path -- src/evox/algorithms/mo/nsga3.py:96-98.

    ``rank = non_dominated_sort(merged_fitness); last_rank = rank[argsort(rank)[pop_size]]``.
    Returns ``(rank, last_rank)``; ``last_rank`` is a Python int for host inputs, a 0-d tensor otherwise.
    """
    L = _lib.lib()
    f = rt.Arg(merged_fitness, torch.float32, "n3_f")
    _check_2d(f, "merged_fitness")
    n = f.dev.shape[0]
    rank = non_dominated_sort(f.dev)
    out = torch.empty((1,), dtype=torch.int32, device=f.dev.device)
    with torch.cuda.device(f.dev.device):
        ws = _ws_call(L.evoxb_kth_rank_workspace_bytes, (n,), f.dev.device)
        _lib.check(L.evoxb_kth_rank_i32(rank.data_ptr(), n, pop_size, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                        rt.stream_ptr(f.dev.device)))
    if f.kind == "cuda":
        return rank, out[0]
    return rt.give_back(rank, f, "n3_r"), int(out.item())
