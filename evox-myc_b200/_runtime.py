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
    """Grow-only scratch buffer per (thread, device, CUDA stream) -- 256-byte aligned by the caching allocator.
    Keyed by the current stream because calls issued under different ``torch.cuda.stream`` contexts overlap on
    the GPU and must not share the dominance matrix; a buffer that is outgrown goes back to the allocator,
    which keeps it tied to the stream it was allocated on."""
    cache = getattr(_tls, "ws", None)
    if cache is None:
        cache = _tls.ws = {}
    dev_index = device.index if device.index is not None else torch.cuda.current_device()
    key = (dev_index, torch.cuda.current_stream(device).cuda_stream, slot)
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


# ---- caller-owned host arrays: page-lock them in place instead of copying them to a staging buffer ----------
# A float32 C-contiguous NumPy array is registered with the CUDA driver once (cudaHostRegister) and then copied
# to the device straight from where it lives; the registration is remembered for the few most recent arrays
# (the typical caller hands over the same buffer every generation).  Anything else goes through the pinned
# staging buffer, which also does the cast.
_REG_MAX = 4
_reg_lock = threading.Lock()
_registered: "dict[tuple[int, int], object]" = {}  # (address, nbytes) -> the array (kept alive while registered)


def _unregister(key) -> None:
    try:
        torch.cuda.cudart().cudaHostUnregister(key[0])
    except Exception:
        pass


def _register_host(arr: np.ndarray) -> bool:
    """Page-locks ``arr``'s memory (idempotent).  False when the driver refuses (e.g. read-only pages)."""
    key = (arr.ctypes.data, arr.nbytes)
    with _reg_lock:
        if key in _registered:
            return True
        if arr.nbytes < (1 << 16):  # not worth a driver call
            return False
        try:
            rc = torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, arr.nbytes, 0)
            if int(rc) != 0:
                return False
        except Exception:
            return False
        while len(_registered) >= _REG_MAX:
            old = next(iter(_registered))
            _registered.pop(old)
            _unregister(old)
        _registered[key] = arr
        return True


def release_host_registrations() -> None:
    with _reg_lock:
        for key in list(_registered):
            _registered.pop(key)
            _unregister(key)


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
            arr = np.asarray(value)
            require_cuda()
            device = device or torch.device("cuda", torch.cuda.current_device())
            if (isinstance(value, np.ndarray) and dtype == torch.float32 and arr.dtype == np.float32
                    and arr.flags.c_contiguous and arr.flags.writeable and _register_host(arr)):
                self.dev = torch.from_numpy(arr).to(device, non_blocking=True)  # DMA from the caller's own pages
                self.h2d_bytes = arr.nbytes
                return
            host = torch.from_numpy(np.ascontiguousarray(arr))
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
