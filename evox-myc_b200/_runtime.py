"""Device plumbing for the Python operators: workspaces, streams, host<->device staging.

PyTorch is used only for device memory, streams and (elsewhere) torch.distributed; every
computation is a call through the C ABI of libevoxb.so.
"""

from __future__ import annotations

import threading

import numpy as np
import torch

_tls = threading.local()


def require_cuda() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError(
            "evox_myc_b200 needs a CUDA device: this path has no CPU fallback "
            "(the CPU restatement under oracle/ is test infrastructure only)"
        )


def workspace(nbytes: int, device: torch.device, slot: str = "main") -> torch.Tensor:
    """Grow-only per-thread, per-device scratch buffer (256-byte aligned by the caching allocator)."""
    cache = getattr(_tls, "ws", None)
    if cache is None:
        cache = _tls.ws = {}
    key = (device.index if device.index is not None else torch.cuda.current_device(), slot)
    buf = cache.get(key)
    if buf is None or buf.numel() < nbytes:
        cache.pop(key, None)
        buf = None
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        cache[key] = buf
    return buf


def release_workspaces() -> None:
    if getattr(_tls, "ws", None):
        _tls.ws.clear()


def stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def pinned(shape, dtype: torch.dtype, slot: str) -> torch.Tensor:
    """Grow-only pinned host staging buffer, one per (slot, dtype)."""
    cache = getattr(_tls, "pin", None)
    if cache is None:
        cache = _tls.pin = {}
    n = int(np.prod(shape)) if len(shape) else 1
    key = (slot, dtype)
    buf = cache.get(key)
    if buf is None or buf.numel() < n:
        buf = torch.empty(max(n, 1), dtype=dtype, pin_memory=True)
        cache[key] = buf
    return buf[:n].view(*shape) if n else buf[:0].view(*shape)


class Arg:
    """One array argument, normalised to a contiguous CUDA tensor of a given dtype.

    Remembers where the caller's data lived so results can be handed back the same way:
    CUDA tensor in -> CUDA tensor out; NumPy / host tensor / list in -> NumPy (or host tensor) out,
    with the host->device and device->host copies made here.
    """

    def __init__(self, value, dtype: torch.dtype, slot: str, device: torch.device | None = None):
        self.kind = "cuda"
        if isinstance(value, torch.Tensor):
            if value.is_cuda:
                self.dev = value.to(dtype).contiguous()
                return
            self.kind = "host_torch"
            host = value.detach()
        else:
            self.kind = "numpy"
            host = torch.from_numpy(np.ascontiguousarray(np.asarray(value)))
        require_cuda()
        device = device or torch.device("cuda", torch.cuda.current_device())
        stage = pinned(tuple(host.shape), dtype, slot)
        stage.copy_(host)  # casts to the compute dtype on the host
        self.dev = stage.to(device, non_blocking=True)
        self.h2d_bytes = stage.numel() * stage.element_size()


def give_back(result: torch.Tensor, like: Arg, slot: str):
    """Returns `result` the way the caller supplied `like`."""
    if like.kind == "cuda":
        return result
    stage = pinned(tuple(result.shape), result.dtype, slot)
    stage.copy_(result, non_blocking=True)
    torch.cuda.current_stream(result.device).synchronize()
    if like.kind == "host_torch":
        return stage.clone()
    return stage.numpy().copy()
