"""evoxb_crowding_noalloc_f32: the crowding distance for callers that cannot allocate.

A handler that runs inside someone else's compiled step (or a loop that has spent its memory
budget) has no workspace to give.  This entry point borrows an 8 MiB array that lives inside
libevoxb.so instead.  The test drives it through the C ABI exactly as such a caller would --
inputs and outputs only -- and compares it with evoxb_crowding_f32 (caller workspace) on two
calls in a row: the second call is smaller and differently masked than the first, so anything
the first call left behind in the shared array would show up as a wrong distance.

The file name sorts last on purpose: a run that stops at the first failure has then already
executed every other test.  It also passes on its own in a fresh process.
"""

import ctypes

import numpy as np
import pytest

import helpers as h

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

RTOL = 1e-6  # the tolerance of tests/test_crowding_gpu.py


def _both(L, x, mask):
    """(no-alloc result, workspace result) for one input, both through the C ABI."""
    n, m = x.shape
    xd = torch.from_numpy(x).cuda()
    md = None if mask is None else torch.from_numpy(mask.astype(np.uint8)).cuda()
    mp = None if md is None else md.data_ptr()
    stream = torch.cuda.current_stream().cuda_stream

    got = torch.full((n,), 123.0, dtype=torch.float32, device="cuda")
    rc = L.evoxb_crowding_noalloc_f32(xd.data_ptr(), mp, n, m, got.data_ptr(), stream)
    assert rc == 0, L.evoxb_last_error()

    need = ctypes.c_size_t()
    assert L.evoxb_crowding_workspace_bytes(n, m, ctypes.byref(need)) == 0
    ws = torch.empty(need.value, dtype=torch.uint8, device="cuda")
    want = torch.full((n,), -123.0, dtype=torch.float32, device="cuda")
    rc = L.evoxb_crowding_f32(xd.data_ptr(), mp, n, m, want.data_ptr(), ws.data_ptr(), ws.numel(), stream)
    assert rc == 0, L.evoxb_last_error()
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def _assert_same(got, want):
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(np.isposinf(got), np.isposinf(want))
    assert np.array_equal(np.isneginf(got), np.isneginf(want))
    fin = np.isfinite(want)
    np.testing.assert_allclose(got[fin], want[fin], rtol=RTOL, atol=0)


def test_two_calls_in_a_row_match_the_workspace_entry_point():
    from evox_myc_b200 import _lib

    L = _lib.lib()
    # first call: large, no mask
    x1 = h.uniform(50000, 3, 101)
    got, want = _both(L, x1, None)
    _assert_same(got, want)
    # second call straight after: smaller, salted with NaN/inf/ties, sparse mask
    x2 = h.nasty(7001, 5, 102)
    m2 = np.random.default_rng(103).random(7001) < 0.3
    got, want = _both(L, x2, m2)
    _assert_same(got, want)
    # and the answer is the oracle's, not merely self-consistent
    from oracle import c_oracle

    _assert_same(got, c_oracle.crowding_distance(x2, m2))


def test_sizes_beyond_the_builtin_array_are_rejected():
    from evox_myc_b200 import _lib

    L = _lib.lib()
    n, m = 262144, 5  # needs ~26 MiB of sort buffers, the array holds 8 MiB
    x = torch.zeros((n, m), dtype=torch.float32, device="cuda")
    out = torch.zeros((n,), dtype=torch.float32, device="cuda")
    rc = L.evoxb_crowding_noalloc_f32(x.data_ptr(), None, n, m, out.data_ptr(), None)
    assert rc == -2
    msg = L.evoxb_last_error().decode()
    assert "built-in array" in msg and "evoxb_crowding_f32" in msg
