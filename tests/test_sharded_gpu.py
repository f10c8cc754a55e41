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


PATH_ENV = {
    "data_flow": {"EVOXB_PEEL_MODE": "flow"},      # peel_flow_kernel across the GPUs (where every rank owns a chunk)
    "grid_peer": {"EVOXB_PEEL_MODE": "grid"},      # peel_p2p_kernel
    "nccl_allgather": {"EVOXB_NO_P2P": "1"},       # one ncclAllGather per front
    "auto": {},                                    # whatever the library picks
    "levels": {"EVOXB_BUILD_MODE": "levels"},      # matrix built from 15-bit levels at every size (csrc/levels.cuh)
}


def _worker(rank, world, port, cases, out, path="auto"):
    os.environ.update(PATH_ENV[path])
    no_p2p = path == "nccl_allgather"
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
        used = sel.last_path()
        nchunks = -(-n // 4096)
        if path == "data_flow" and world <= nchunks <= 128 and not used.startswith("data-flow"): why.append(f"ran on '{used}'")
        if path == "grid_peer" and not used.startswith("grid-wide"): why.append(f"ran on '{used}'")
        if path == "nccl_allgather" and not used.startswith("ncclAllGather"): why.append(f"ran on '{used}'")
        res.append((kind, n, m, "; ".join(why)))
        sel.close()
    out[rank] = res
    dist.destroy_process_group()


# small populations (fewer chunks than ranks at the larger world sizes), then enough chunks for every rank
CASES = [("uniform", 5000, 3, 1), ("nasty", 9001, 4, 2), ("uniform", 20000, 2, 3), ("integer_grid", 12000, 3, 4),
         ("uniform", 40000, 5, 5), ("uniform", 70000, 3, 6), ("nasty", 150001, 4, 7)]
# BASELINE config 4 at full size: 64 chunks, 8 per rank at world 8
FULL = [("uniform", 262144, 5, 8)]


def _run(world, path="auto", cases=CASES):
    import torch.multiprocessing as mp

    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), cases, out, path), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        for kind, n, m, why in out[r]:
            assert not why, f"rank {r} of {world} ({path}): {kind} {n}x{m}: {why}"


@pytest.mark.parametrize("path", ["data_flow", "grid_peer", "nccl_allgather", "auto", "levels"])
def test_sharded_world_size_1(path):
    _run(1, path)


@pytest.mark.parametrize("path", ["data_flow", "grid_peer", "nccl_allgather", "auto", "levels"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_multi_gpu(world, path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world, path)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_config4_full_size(world):
    """N = 262144, M = 5 sharded (ranks, crowding and survivor order against the C oracle on every rank)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world, "auto", FULL)
