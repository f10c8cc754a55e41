"""NSGA-II with the environmental selection, variation and DTLZ2 evaluation on the GPU.

Mirrors ``src/evox/algorithms/mo/nsga2.py`` of the reference (same constructor arguments, same
``setup / init_ask / init_tell / ask / tell`` protocol, ``State`` with ``population``, ``fitness``,
``next_generation``), plus the two operators its defaults use and the DTLZ2 problem -- enough to run
BASELINE config 3 (pop = 16384, M = 3, D = 12) end to end.  The reference's core runtime (State
pytrees, jit_class, StdWorkflow, monitors) is out of scope: ``State`` here
is a plain attribute bag and ``run`` is a bare generation loop.

Random streams: the reference threads a JAX PRNG key through the state; here the key is a
``(seed, generation)`` pair feeding Philox inside the kernels, so trajectories cannot be compared
bit for bit with JAX (statistical parity; tests/test_nsga2_gpu.py).
"""

from __future__ import annotations

import torch

from . import _lib, _runtime as rt
from .operators import _gather_rows, nsga2_select


class State:
    """Attribute bag with the reference's ``update`` idiom (core/state.py:178-182 forbids setattr)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def update(self, **kw) -> "State":
        d = dict(self.__dict__)
        d.update(kw)
        return State(**d)


def _as_dev(x, device):
    return x.to(device=device, dtype=torch.float32).contiguous() if isinstance(x, torch.Tensor) else torch.as_tensor(
        x, dtype=torch.float32, device=device).contiguous()


class SimulatedBinary:
    """SBX parameters -- operators/crossover/sbx.py:44-62 (type 1 only: offspring as many as parents)."""

    def __init__(self, pro_c=1, dis_c=20, type=1):
        if type != 1:
            raise NotImplementedError("only type=1 (the NSGA-II default) is implemented")
        self.pro_c, self.dis_c, self.type = float(pro_c), float(dis_c), type


class Polynomial:
    """Polynomial mutation parameters -- operators/mutation/pm_mutation.py:57-73."""

    def __init__(self, boundary, pro_m=1, dis_m=20):
        self.boundary, self.pro_m, self.dis_m = boundary, float(pro_m), float(dis_m)


def variation(population: torch.Tensor, lb: torch.Tensor, ub: torch.Tensor, seed: int, generation: int,
              crossover: SimulatedBinary | None = None, mutation: Polynomial | None = None) -> torch.Tensor:
    """``clip(mutation(crossover(population)), lb, ub)`` -- nsga2.py:76-84, one fused kernel."""
    crossover = crossover or SimulatedBinary()
    mutation = mutation or Polynomial((lb, ub))
    L = _lib.lib()
    n, d = population.shape
    out = torch.empty_like(population)
    with torch.cuda.device(population.device):
        _lib.check(L.evoxb_sbx_pm_clip_f32(population.data_ptr(), n, d, lb.data_ptr(), ub.data_ptr(), crossover.pro_c,
                                           crossover.dis_c, mutation.pro_m, mutation.dis_m, seed & 0xFFFFFFFF,
                                           generation & 0xFFFFFFFF, out.data_ptr(), rt.stream_ptr(population.device)))
    return out


class DTLZ2:
    """DTLZ2 -- problems/numerical/dtlz.py:70-108 (``evaluate`` only; m = 3, d = m + 9 by default)."""

    def __init__(self, d=None, m=None):
        self.m = 3 if m is None else m
        self.d = self.m + 9 if d is None else d

    def evaluate(self, state, X: torch.Tensor):
        L = _lib.lib()
        n, d = X.shape
        f = torch.empty((n, self.m), dtype=torch.float32, device=X.device)
        with torch.cuda.device(X.device):
            _lib.check(L.evoxb_dtlz2_f32(X.data_ptr(), n, d, self.m, f.data_ptr(), rt.stream_ptr(X.device)))
        return f, state


class LSMOP1:
    """LSMOP1 -- problems/numerical/lsmop.py:17-58 (group sizes) and :114-141 (``evaluate``).

    ``d`` defaults as in the reference: the base class sizes the groups for ``d = 100 m`` (lsmop.py:40-43), then
    ``LSMOP1.__init__`` resets ``self.d = m + 4`` when no ``d`` is given (:122-125) -- the groups keep the sizes
    computed from 100 m.  Pass ``d`` explicitly (as the reference's own test does, tests/test_lsmop.py:19-20).
    """

    NK = 5

    def __init__(self, d=None, m=None, ref_num=1000):
        import numpy as np

        self.m = 3 if m is None else m
        group_d = self.m * 100 if d is None else d
        self.d = self.m + 4 if d is None else d
        self.ref_num = ref_num
        c = [3.8 * 0.1 * (1 - 0.1)]
        for _ in range(1, self.m):
            c.append(3.8 * c[-1] * (1 - c[-1]))
        c = np.asarray(c, dtype=np.float32)  # the reference evaluates these lines in float32 (jnp arrays)
        sublen = np.floor(c / np.sum(c, dtype=np.float32) * np.float32(group_d) / np.float32(self.NK)).astype(np.float32)
        length = np.concatenate([[0], np.cumsum(sublen * self.NK)])
        self.sublen = tuple(int(v) for v in sublen)
        self.len = tuple(int(v) for v in length)

    def evaluate(self, state, X: torch.Tensor):
        import ctypes

        L = _lib.lib()
        X = X.to(torch.float32).contiguous()
        n, d = X.shape
        f = torch.empty((n, self.m), dtype=torch.float32, device=X.device)
        sub = (ctypes.c_int32 * self.m)(*self.sublen)
        start = (ctypes.c_int32 * self.m)(*self.len[: self.m])
        with torch.cuda.device(X.device):
            _lib.check(L.evoxb_lsmop1_f32(X.data_ptr(), n, d, self.m, sub, start, f.data_ptr(), rt.stream_ptr(X.device)))
        return f, state


class NSGA2:
    """NSGA-II -- src/evox/algorithms/mo/nsga2.py:22-100."""

    def __init__(self, lb, ub, n_objs, pop_size, selection_op=None, mutation_op=None, crossover_op=None, device=None):
        rt.require_cuda()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.lb, self.ub = _as_dev(lb, self.device), _as_dev(ub, self.device)
        self.n_objs, self.dim, self.pop_size = n_objs, self.lb.shape[0], pop_size
        self.mutation = mutation_op or Polynomial((self.lb, self.ub))  # nsga2.py:50-51
        self.crossover = crossover_op or SimulatedBinary()              # nsga2.py:52-53

    def setup(self, key: int) -> State:
        """nsga2.py:55-67 -- uniform population in [lb, ub], zero fitness; ``key`` is an integer seed."""
        g = torch.Generator(device=self.device).manual_seed(int(key))
        population = torch.rand((self.pop_size, self.dim), generator=g, device=self.device) * (self.ub - self.lb) + self.lb
        return State(population=population,
                     fitness=torch.zeros((self.pop_size, self.n_objs), device=self.device),
                     next_generation=population, key=int(key), generation=0)

    def init_ask(self, state):  # nsga2.py:69-70
        return state.population, state

    def init_tell(self, state, fitness):  # nsga2.py:72-74
        return state.update(fitness=fitness)

    def ask(self, state):  # nsga2.py:76-84
        nxt = variation(state.population, self.lb, self.ub, state.key, state.generation, self.crossover, self.mutation)
        return nxt, state.update(next_generation=nxt, generation=state.generation + 1)

    def tell(self, state, fitness):  # nsga2.py:86-100
        merged_pop = torch.cat([state.population, state.next_generation], dim=0)
        merged_fitness = torch.cat([state.fitness, fitness], dim=0)
        order = nsga2_select(merged_fitness, self.pop_size)
        return state.update(population=_gather_rows(merged_pop, order), fitness=_gather_rows(merged_fitness, order))


def run(algorithm: NSGA2, problem, generations: int, key: int = 0) -> State:
    """Bare generation loop: what ``StdWorkflow.step`` does for this algorithm (std_workflow.py:129-199)
    without monitors, transforms or sharding."""
    state = algorithm.setup(key)
    pop, state = algorithm.init_ask(state)
    fit, _ = problem.evaluate(None, pop)
    state = algorithm.init_tell(state, fit)
    for _ in range(generations):
        pop, state = algorithm.ask(state)
        fit, _ = problem.evaluate(None, pop)
        state = algorithm.tell(state, fit)
    return state
