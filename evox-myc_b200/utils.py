"""Mirror of the one helper of ``src/evox/utils/common.py`` that is on the path (:95-104)."""

from __future__ import annotations

import torch

from . import _lib, _runtime as rt


def dominate_relation(x, y):
    """``A[i, j]`` is True iff ``x_i`` dominates ``y_j`` -- utils/common.py:101-104.

    ``x`` is (nx, m), ``y`` is (ny, m); returns a boolean (nx, ny) matrix.  This is the dense
    form kept for API parity (SPEA2 calls it directly, algorithms/mo/spea2.py:24); the ranking
    path never materialises bytes, it builds one bit per pair (csrc/dominance.cuh).
    """
    L = _lib.lib()
    a = rt.Arg(x, torch.float32, "dom_x")
    b = rt.Arg(y, torch.float32, "dom_y", device=a.dev.device)
    if a.dev.dim() != 2 or b.dev.dim() != 2 or a.dev.shape[1] != b.dev.shape[1]:
        raise ValueError("x and y must be (nx, m) and (ny, m)")
    nx, m = a.dev.shape
    ny = b.dev.shape[0]
    out = torch.empty((nx, ny), dtype=torch.uint8, device=a.dev.device)
    if nx and ny:
        with torch.cuda.device(a.dev.device):
            _lib.check(
                L.evoxb_dominate_relation_f32(a.dev.data_ptr(), nx, b.dev.data_ptr(), ny, m, out.data_ptr(),
                                              rt.stream_ptr(a.dev.device))
            )
    res = rt.give_back(out, a, "dom_out")
    return res.bool() if isinstance(res, torch.Tensor) else res.astype(bool)
