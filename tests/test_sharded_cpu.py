"""Host-side logic of the multi-GPU path on CPU: the shard plan, and the per-front exchange protocol
run by two gloo ranks on NumPy arrays (the oracle supplies each rank's matrix columns)."""

import os
import socket
import sys

import numpy as np
import pytest

import helpers as h

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_plan_partitions_every_chunk_once():
    from evox_myc_b200 import sharded

    for n in (1, 4096, 4097, 65536, 262144, 100000):
        for nranks in (1, 2, 3, 4, 8):
            seen = []
            for r in range(nranks):
                first, stride, nlocal, total = sharded.shard_plan(n, r, nranks)
                assert total == -(-n // 4096) and stride == nranks and first == r
                seen += list(range(first, first + nlocal * stride, stride))
            assert sorted(seen) == list(range(total))
            cols = np.concatenate([sharded.owned_columns(n, r, nranks) for r in range(nranks)])
            assert sorted(cols.tolist()) == list(range(n))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, m, seed, out, flow=False):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    from evox_myc_b200 import sharded
    from oracle import nds_oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import helpers

    x = helpers.nasty(n, m, seed)

    def dominates_cols(cols):
        return nds_oracle.dominate_relation(x, x[cols])

    def all_gather(bits):
        t = torch.from_numpy(bits)
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t)
        return [o.numpy() for o in outs]

    want = nds_oracle.non_dominated_sort(x)
    if flow:
        # two calls in a row: the second continues from the first one's epoch (buffers are never cleared)
        ranks, fronts, e1 = sharded.simulate_flow_exchange(dominates_cols, n, rank, world, all_gather)
        ranks2, fronts2, e2 = sharded.simulate_flow_exchange(dominates_cols, n, rank, world, all_gather, epoch0=e1)
        ok = bool(np.array_equal(ranks, want) and np.array_equal(ranks2, want) and fronts == fronts2 and e2 == 2 * e1)
        out[rank] = (ok, fronts + 1, int(want.max()) + 1)
    else:
        ranks, rounds = sharded.simulate_exchange(dominates_cols, n, rank, world, all_gather)
        out[rank] = (bool(np.array_equal(ranks, want)), rounds, int(want.max()) + 1)
    dist.destroy_process_group()


@pytest.mark.parametrize("n,m", [(9000, 3), (5000, 2)])
def test_exchange_protocol_two_gloo_ranks(n, m):
    import torch.multiprocessing as mp

    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, n, m, 7, out), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        ok, rounds, fronts = out[r]
        assert ok, f"rank {r}: ranks differ from the oracle"
        assert fronts < rounds <= fronts + 4  # stops within one batch after the first empty front


@pytest.mark.parametrize("n", [1, 4097, 32768, 65536, 150001, 262144, 524288])
@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_flow_plan_tiles_every_column_once(n, nranks):
    """Column ownership of the data-flow peel kernel (host arithmetic of csrc/peel_flow.cuh): over all ranks the
    pieces must cover every 256-column unit of every chunk exactly once, each rank only its own chunks, every CTA
    at least one piece, and the largest CTA must fit the kernel's shared-memory budget."""
    import ctypes

    from evox_myc_b200 import _lib

    L = _lib.lib()
    nchunks = -(-n // 4096)
    seen = np.zeros((nchunks, 16), dtype=np.int32)
    for rank in range(nranks):
        grid, npieces, max_words = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        cap = 16 * nchunks + 16
        buf = (ctypes.c_int32 * (4 * cap))()
        _lib.check(L.evoxb_flow_plan(n, rank, nranks, 1, 148, ctypes.byref(grid), ctypes.byref(npieces), ctypes.byref(max_words),
                                     buf, cap))
        pieces = np.frombuffer(buf, dtype=np.int32)[: 4 * npieces.value].reshape(-1, 4)
        own = [c for c in range(nchunks) if c % nranks == rank]
        if not own:
            assert grid.value == 0 and npieces.value == 0
            continue
        assert 1 <= grid.value <= 148 and grid.value == min(148, npieces.value)
        assert sorted(set(pieces[:, 0])) == list(range(grid.value))            # no idle CTA among those launched
        assert set(pieces[:, 1]) == set(own)
        for cta, chunk, u0, nu in pieces:
            assert nu in (1, 2, 4, 8, 16) and u0 % nu == 0 and u0 + nu <= 16
            seen[chunk, u0 : u0 + nu] += 1
        words = np.bincount(pieces[:, 0], weights=pieces[:, 3] * 8, minlength=grid.value)
        assert int(words.max()) == max_words.value
        assert max_words.value * 128 <= 96 * 1024                              # counts of a CTA's columns in shared memory
    assert np.all(seen == 1)


@pytest.mark.parametrize("n,m", [(9000, 3), (12300, 2)])
def test_flow_exchange_protocol_two_gloo_ranks(n, m):
    """The data-flow peel's exchange (tagged words, epoch parity, every rank stops on the same empty front)."""
    import torch.multiprocessing as mp

    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, m, 11, out, True), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        ok, rounds, fronts = out[r]
        assert ok, f"rank {r}: ranks differ from the oracle"
        assert rounds == fronts + 1  # exactly one empty front is exchanged
