"""GPU tests of the multi-GPU path.  World size 1 runs on the single-GPU box (same kernels, same host
loop, a one-rank NCCL communicator); world size 2+ needs as many GPUs and is skipped otherwise."""

import os
import socket
import sys

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, cases, out, no_p2p=False):
    if no_p2p:
        os.environ["EVOXB_NO_P2P"] = "1"
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    import helpers
    from evox_myc_b200 import sharded
    from oracle import c_oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl" if world > 1 else "gloo", rank=rank, world_size=world,
                            **({"device_id": torch.device("cuda", rank)} if world > 1 else {}))
    res = []
    for kind, n, m, seed in cases:
        x = getattr(helpers, kind)(n, m, seed)
        sel = sharded.ShardedSelector(n, m, torch.device("cuda", rank))
        r1 = sel.non_dominated_sort(x)
        r2 = sel.non_dominated_sort(torch.from_numpy(x).cuda(rank)).cpu().numpy()  # second call, same workspace
        pop = n // 2
        order, rk, crowd = sel.nsga2_select(x, pop, return_details=True)
        w_order, w_rank, w_crowd = c_oracle.nsga2_select(x, pop)
        fin = np.isfinite(w_crowd)
        p2p, err, fronts = sel.status()
        why = []
        if not np.array_equal(r1, w_rank): why.append(f"ranks differ in {int(np.sum(r1 != w_rank))} rows (host input)")
        if not np.array_equal(r2, w_rank): why.append(f"ranks differ in {int(np.sum(r2 != w_rank))} rows (device input)")
        if not np.array_equal(rk, w_rank): why.append("select: ranks differ")
        if not np.array_equal(order, w_order): why.append("select: survivor order differs")
        if not (np.array_equal(np.isfinite(crowd), fin) and np.allclose(crowd[fin], w_crowd[fin], rtol=1e-6, atol=0)):
            why.append("select: crowding differs")
        if err: why.append("a peer timed out")
        if p2p == no_p2p: why.append(f"peer-memory path enabled={p2p}, expected {not no_p2p}")
        if p2p and fronts != int(w_rank.max()) + 1: why.append(f"fronts {fronts} != {int(w_rank.max()) + 1}")
        res.append((kind, n, m, "; ".join(why)))
        sel.close()
    out[rank] = res
    dist.destroy_process_group()


CASES = [("uniform", 5000, 3, 1), ("nasty", 9001, 4, 2), ("uniform", 20000, 2, 3), ("integer_grid", 12000, 3, 4),
         ("uniform", 40000, 5, 5)]


def _run(world, no_p2p=False):
    import torch.multiprocessing as mp

    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), CASES, out, no_p2p), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        for kind, n, m, why in out[r]:
            assert not why, f"rank {r} of {world}: {kind} {n}x{m}: {why}"


@pytest.mark.parametrize("no_p2p", [False, True], ids=["peer_memory", "nccl_allgather"])
def test_sharded_world_size_1(no_p2p):
    _run(1, no_p2p)


@pytest.mark.parametrize("no_p2p", [False, True], ids=["peer_memory", "nccl_allgather"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_multi_gpu(world, no_p2p):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world, no_p2p)
