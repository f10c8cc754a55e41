"""Multi-GPU environmental selection: one process per GPU, dominance matrix sharded by column chunk.

The reference never shards this path (``tell`` is replicated on every device,
src/evox/workflows/std_workflow.py:194, :341); this is new capability behind the same results.
``torch.distributed`` is used for plumbing only -- to hand the 128-byte NCCL id from rank 0 to the
others.  The exchange itself (one bit per individual per front) is an ``ncclAllGather`` issued by
libevoxb.so on its own communicator, between its own kernels.

``shard_plan`` / ``simulate_exchange`` are the host-side logic and run on CPU (the latter is the
executable specification of the protocol, used by the world_size-2 gloo test with NumPy arrays).
"""

from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib, _runtime as rt

CHUNK = 4096  # columns per chunk (csrc/peel.cuh: kChunkCols)
FRONT_BATCH = 4  # fronts between two termination checks (csrc/evoxb.cu: kFrontBatch)


def shard_plan(n: int, rank: int, nranks: int):
    """(first_chunk, chunk_stride, local_chunks, total_chunks) of ``rank`` -- evoxb_shard_plan."""
    L = _lib.lib()
    a, b, c, d = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _lib.check(L.evoxb_shard_plan(n, rank, nranks, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c), ctypes.byref(d)))
    return a.value, b.value, c.value, d.value


def owned_columns(n: int, rank: int, nranks: int) -> np.ndarray:
    first, stride, nlocal, _ = shard_plan(n, rank, nranks)
    cols = [np.arange(c * CHUNK, min(n, (c + 1) * CHUNK)) for c in range(first, first + nlocal * stride, stride)]
    return np.concatenate(cols) if cols else np.zeros(0, dtype=np.int64)


def simulate_exchange(dominates_cols, n: int, rank: int, nranks: int, all_gather):
    """The per-front protocol of ``run_nds_sharded`` on host arrays.

    ``dominates_cols(cols) -> bool (n, len(cols))`` gives this rank's columns of the dominance matrix;
    ``all_gather(bits) -> list of every rank's bits``.  Returns the full rank vector (identical on
    every rank) and the number of exchange rounds.
    """
    first, stride, nlocal, nchunks = shard_plan(n, rank, nranks)
    slice_chunks = -(-nchunks // nranks)
    cols = owned_columns(n, rank, nranks)
    d = dominates_cols(cols)
    count = d.sum(axis=0).astype(np.int64)  # counts of OWN columns only: no reduction across ranks
    ranks = np.zeros(n, dtype=np.int32)
    r = 0
    rounds = 0
    while True:
        sizes = []
        for _ in range(FRONT_BATCH):
            bits = np.zeros(slice_chunks * CHUNK, dtype=np.uint8)  # one entry per column slot of my slice
            mine = count == 0
            count[mine] = -1
            lc = (cols // CHUNK - first) // stride
            bits[lc[mine] * CHUNK + cols[mine] % CHUNK] = 1
            gathered = all_gather(bits)
            rounds += 1
            members = []
            for c in range(nchunks):  # front_expand_kernel: chunk c lives in slice c % P at slot c // P
                sl = gathered[c % nranks][(c // nranks) * CHUNK : (c // nranks + 1) * CHUNK]
                members.append(c * CHUNK + np.nonzero(sl)[0])
            members = np.concatenate(members)
            ranks[members] = r
            sizes.append(len(members))
            if len(members):
                count -= np.where(count >= 0, d[members].sum(axis=0), 0)
            r += 1
        if min(sizes) == 0:
            return ranks, rounds


def simulate_flow_exchange(dominates_cols, n: int, rank: int, nranks: int, all_gather, epoch0: int = 0):
    """The protocol of ``peel_flow_kernel`` across ranks, on host arrays (executable specification, CPU tests).

    Every rank owns the counts of its own column chunks.  Per front it publishes, for each of its 32-column
    words, the pair ``(word, epoch)`` into one of two buffers chosen by the epoch's parity (here: one
    ``all_gather`` of the tagged words), reads the COMPLETE bitmask of that front -- every word must carry the
    current epoch, which is what allows a buffer to be reused two fronts later -- writes the ranks, subtracts the
    members' rows from its counts, and stops on the first front that is empty everywhere.  Returns the rank
    vector, the number of fronts and the epoch reached (the next call continues from it).
    """
    first, stride, nlocal, nchunks = shard_plan(n, rank, nranks)
    cols = owned_columns(n, rank, nranks)
    d = dominates_cols(cols)
    count = d.sum(axis=0).astype(np.int64)
    nw = nchunks * (CHUNK // 32)
    word_of = cols // 32
    ranks = np.full(n, -1, dtype=np.int32)
    buffers = np.zeros((2, nw, 2), dtype=np.int64)  # [parity][word] = (bits, epoch); epoch 0 = never written
    e, r = epoch0, 0
    while True:
        e += 1
        mine = count == 0
        count[mine] = -1  # leaves for good (non_dominate.py:103)
        words = np.zeros((nw, 2), dtype=np.int64)
        own_words = np.unique(np.concatenate([np.arange(c * (CHUNK // 32), (c + 1) * (CHUNK // 32))
                                              for c in range(first, first + nlocal * stride, stride)])) if nlocal else np.zeros(0, np.int64)
        words[own_words, 1] = e  # every owned word is published every front, zero or not
        np.add.at(words[:, 0], word_of[mine], np.left_shift(1, cols[mine] % 32).astype(np.int64))
        gathered = all_gather(words)
        slot = buffers[e & 1]
        for q, w in enumerate(gathered):
            sel = w[:, 1] == e
            assert not np.any(sel & (slot[:, 1] == e)), "two ranks published the same word"
            slot[sel] = w[sel]
        assert np.all(slot[:, 1] == e), "a word of this front is missing: the reader would spin"
        bits = slot[:, 0]
        members = np.nonzero((bits[:, None] >> np.arange(32)) & 1)
        members = members[0] * 32 + members[1]
        members = members[members < n]
        if len(members) == 0:
            return ranks, r, e
        ranks[members] = r
        count -= np.where(count >= 0, d[members].sum(axis=0), 0)
        r += 1


class ShardedSelector:
    """Holds the NCCL communicator and workspace of one rank; all ranks must make the same calls."""

    def __init__(self, n: int, m: int, device: torch.device | None = None):
        import torch.distributed as dist

        rt.require_cuda()
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised (one process per GPU)")
        self.rank, self.nranks = dist.get_rank(), dist.get_world_size()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.n, self.m = n, m
        L = _lib.lib()
        backend = dist.get_backend()
        id_dev = self.device if backend == "nccl" else torch.device("cpu")
        idbuf = (ctypes.c_ubyte * 128)()
        if self.rank == 0:
            _lib.check(L.evoxb_comm_unique_id(idbuf))
        t = torch.tensor(list(idbuf), dtype=torch.uint8, device=id_dev)
        dist.broadcast(t, src=0)
        raw = bytes(t.cpu().tolist())
        self._comm = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(L.evoxb_comm_create(raw, self.rank, self.nranks, ctypes.byref(self._comm)))
            need = ctypes.c_size_t()
            _lib.check(L.evoxb_nds_sharded_workspace_bytes(n, m, self.rank, self.nranks, ctypes.byref(need)))
            self._ws = torch.empty(need.value, dtype=torch.uint8, device=self.device)

    def status(self):
        """(peer_memory_path, error, fronts_of_last_call) -- synchronises the device."""
        a, b, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        torch.cuda.synchronize(self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().evoxb_comm_status(self._comm, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return bool(a.value), bool(b.value), c.value

    PATHS = {0: "ncclAllGather per front", 1: "grid-wide peer-memory kernel", 2: "data-flow peer-memory kernel"}

    def last_path(self) -> str:
        """Which exchange the last call used (see include/evoxb.h, section e)."""
        return self.PATHS.get(int(_lib.lib().evoxb_comm_last_path(self._comm)), "unknown")

    def set_timeout_ms(self, ms: int) -> None:
        _lib.check(_lib.lib().evoxb_comm_set_timeout_ms(self._comm, int(ms)))

    def _raise_if_peer_failed(self, synced: bool) -> None:
        """A peer time-out leaves ranks of -1; that must be an exception, not a value somebody has to inspect.
        The kernels raise a host word, so after the call's stream has been synchronised this is exact; for
        results left on the device the check covers every EARLIER call (the library itself refuses further
        calls on a failed communicator)."""
        if _lib.lib().evoxb_comm_poll_error(self._comm):
            raise _lib.EvoxbError(-5, "a peer GPU did not answer within the time-out: the ranks of that call are -1 "
                                      "and this ShardedSelector is unusable (close it on every rank and make a new one)"
                                      + ("" if synced else " [detected on a later call]"))

    def close(self):
        if getattr(self, "_comm", None):
            _lib.lib().evoxb_comm_destroy(self._comm)
            self._comm = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def non_dominated_sort(self, x):
        L = _lib.lib()
        a = rt.Arg(x, torch.float32, "snds_x", device=self.device)
        n, m = a.dev.shape
        assert (n, m) == (self.n, self.m), "ShardedSelector was sized for another population"
        rank = torch.empty((n,), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(L.evoxb_nds_rank_sharded_f32(a.dev.data_ptr(), n, m, rank.data_ptr(), self._ws.data_ptr(),
                                                    self._ws.numel(), self._comm, rt.stream_ptr(self.device)))
        out = rt.give_back(rank, a, "snds_r")  # synchronises the stream unless the input was a CUDA tensor
        self._raise_if_peer_failed(synced=a.kind != "cuda")
        return out

    def nsga2_select(self, merged_fitness, pop_size: int, return_details: bool = False):
        L = _lib.lib()
        f = rt.Arg(merged_fitness, torch.float32, "ssel_f", device=self.device)
        n, m = f.dev.shape
        assert (n, m) == (self.n, self.m), "ShardedSelector was sized for another population"
        order = torch.empty((pop_size,), dtype=torch.int32, device=self.device)
        rank = torch.empty((n,), dtype=torch.int32, device=self.device) if return_details else None
        crowd = torch.empty((n,), dtype=torch.float32, device=self.device) if return_details else None
        with torch.cuda.device(self.device):
            _lib.check(
                L.evoxb_nsga2_select_sharded_f32(f.dev.data_ptr(), n, m, pop_size, order.data_ptr(),
                                                 rank.data_ptr() if rank is not None else None,
                                                 crowd.data_ptr() if crowd is not None else None,
                                                 self._ws.data_ptr(), self._ws.numel(), self._comm,
                                                 rt.stream_ptr(self.device))
            )
        if return_details:
            out = (rt.give_back(order, f, "ssel_o"), rt.give_back(rank, f, "ssel_r"), rt.give_back(crowd, f, "ssel_c"))
        else:
            out = rt.give_back(order, f, "ssel_o")
        self._raise_if_peer_failed(synced=f.kind != "cuda")
        return out
