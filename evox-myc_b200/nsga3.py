"""NSGA-III with the whole of ``tell`` on the GPU.

Mirrors ``src/evox/algorithms/mo/nsga3.py`` of the reference (same constructor arguments and
``setup / init_ask / init_tell / ask / tell`` protocol) and ``UniformSampling`` of
``src/evox/operators/sampling/uniform.py``.  ``tell`` is one call through the C ABI
(``evoxb_nsga3_select_f32``: ranking, normalisation, association, niche-preserving selection).

Random streams: the reference's ``state.key`` is a JAX threefry key (two uint32 words).  The niche loop
consumes it exactly like the reference (``jax.random.split`` / ``randint`` restated on the device), so with
the same key and the same fitness the survivors are the reference's.  ``ask`` (crossover / mutation) uses
the Philox streams of nsga2.variation, as in the NSGA-II mirror: those cannot match JAX bit for bit.
"""

from __future__ import annotations

import ctypes
from itertools import combinations
from math import comb

import numpy as np
import torch

from . import _lib, _runtime as rt
from .nsga2 import Polynomial, SimulatedBinary, State, _as_dev, variation
from .operators import _check_2d, _gather_rows


class UniformSampling:
    """Das and Dennis reference points, two layers when h1 < m -- operators/sampling/uniform.py:10-60.
    Host-side set-up code (NumPy, float32 like the reference's jnp arrays); ``__call__`` returns ``(w, n)``."""

    def __init__(self, n=None, m=None):
        self.n, self.m = n, m

    def __call__(self, key=None):
        n, m = self.n, self.m
        f32 = np.float32
        h1 = 1
        while comb(h1 + m, m - 1) <= n:
            h1 += 1

        def layer(h):
            w = (np.array(list(combinations(range(1, h + m), m - 1))) - np.tile(np.arange(m - 1), (comb(h + m - 1, m - 1), 1)) - 1)
            return ((np.c_[w, np.zeros((w.shape[0], 1)) + h] - np.c_[np.zeros((w.shape[0], 1)), w]) / h).astype(f32)

        w = layer(h1)
        if h1 < m:
            h2 = 0
            while comb(h1 + m - 1, m - 1) + comb(h2 + m, m - 1) <= n:
                h2 += 1
            if h2 > 0:
                w = np.r_[w, (layer(h2) / f32(2.0) + f32(1.0 / (2.0 * m))).astype(f32)]
        w = np.maximum(w, f32(1e-6)).astype(f32)
        return w, w.shape[0]


def nsga3_select(merged_fitness, pop_size: int, ref, key, return_details: bool = False):
    """``NSGA3.tell`` from the ranking on -- nsga3.py:96-211.

    ``ref`` is the (r, m) array of reference points, ``key`` the two uint32 words of the JAX PRNG key.
    Returns ``(selected_idx, next_key)``; with ``return_details`` also a dict with ``rank``, ``group_id``
    (r = no direction), ``group_dist`` and ``norm`` (ideal point, nadir point, extreme row indices).
    """
    L = _lib.lib()
    f = rt.Arg(merged_fitness, torch.float32, "n3_f")
    _check_2d(f, "merged_fitness")
    dev = f.dev.device
    n, m = f.dev.shape
    refd = ref if isinstance(ref, torch.Tensor) and ref.is_cuda else torch.as_tensor(np.asarray(ref, dtype=np.float32), device=dev)
    refd = refd.to(torch.float32).contiguous()
    r = refd.shape[0]
    sel = torch.empty((pop_size,), dtype=torch.int32, device=dev)
    rank = torch.empty((n,), dtype=torch.int32, device=dev) if return_details else None
    gid = torch.empty((n,), dtype=torch.int32, device=dev) if return_details else None
    gdist = torch.empty((n,), dtype=torch.float32, device=dev) if return_details else None
    norm = torch.empty((3 * m,), dtype=torch.float32, device=dev) if return_details else None
    kin = (ctypes.c_uint32 * 2)(int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF)
    kout = (ctypes.c_uint32 * 2)()
    need = ctypes.c_size_t()
    _lib.check(L.evoxb_nsga3_select_workspace_bytes(n, m, r, ctypes.byref(need)))
    ws = rt.workspace(need.value, dev)
    ptr = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
    with torch.cuda.device(dev):
        _lib.check(L.evoxb_nsga3_select_f32(f.dev.data_ptr(), n, m, pop_size, refd.data_ptr(), r, kin, sel.data_ptr(), kout,
                                            ptr(rank), ptr(gid), ptr(gdist), ptr(norm), ws.data_ptr(), ws.numel(),
                                            rt.stream_ptr(dev)))
    next_key = (int(kout[0]), int(kout[1]))
    out = rt.give_back(sel, f, "n3_sel")
    if not return_details:
        return out, next_key
    details = {"rank": rt.give_back(rank, f, "n3_rank"), "group_id": rt.give_back(gid, f, "n3_gid"),
               "group_dist": rt.give_back(gdist, f, "n3_gd"), "norm": rt.give_back(norm, f, "n3_norm")}
    return out, next_key, details


class NSGA3:
    """NSGA-III -- src/evox/algorithms/mo/nsga3.py:27-211."""

    def __init__(self, lb, ub, n_objs, pop_size, selection_op=None, mutation_op=None, crossover_op=None, device=None):
        rt.require_cuda()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.lb, self.ub = _as_dev(lb, self.device), _as_dev(ub, self.device)
        self.n_objs, self.dim, self.pop_size = n_objs, self.lb.shape[0], pop_size
        self.mutation = mutation_op or Polynomial((self.lb, self.ub))
        self.crossover = crossover_op or SimulatedBinary()
        self.sampling = UniformSampling(self.pop_size, self.n_objs)  # nsga3.py:60

    def setup(self, key):
        """nsga3.py:62-76.  ``key``: an integer seed (PRNGKey(seed) = (0, seed)) or a pair of uint32 words."""
        if isinstance(key, int):
            key = ((key >> 32) & 0xFFFFFFFF, key & 0xFFFFFFFF)
        g = torch.Generator(device=self.device).manual_seed(int(key[1]))
        population = torch.rand((self.pop_size, self.dim), generator=g, device=self.device) * (self.ub - self.lb) + self.lb
        self.ref = torch.as_tensor(self.sampling()[0], device=self.device)
        return State(population=population, fitness=torch.zeros((self.pop_size, self.n_objs), device=self.device),
                     next_generation=population, key=tuple(key), generation=0)

    def init_ask(self, state):
        return state.population, state

    def init_tell(self, state, fitness):
        return state.update(fitness=fitness)

    def ask(self, state):  # nsga3.py:85-91
        nxt = variation(state.population, self.lb, self.ub, state.key[1], state.generation, self.crossover, self.mutation)
        return nxt, state.update(next_generation=nxt, generation=state.generation + 1)

    def tell(self, state, fitness):  # nsga3.py:93-211
        merged_pop = torch.cat([state.population, state.next_generation], dim=0)
        merged_fitness = torch.cat([state.fitness, fitness], dim=0)
        idx, next_key = nsga3_select(merged_fitness, self.pop_size, self.ref, state.key)
        return state.update(population=_gather_rows(merged_pop, idx), fitness=_gather_rows(merged_fitness, idx), key=next_key)
