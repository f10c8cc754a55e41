"""The C-ABI library loads on a CPU-only box and exports every symbol include/evoxb.h declares."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "evoxb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(evoxb_\w+)\s*\(", src)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for must in ("evoxb_nds_rank_f32", "evoxb_nds_workspace_bytes", "evoxb_dominate_relation_f32"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from evox_myc_b200 import _build, _lib

    if not os.path.exists(_build.LIB_PATH):
        _build.build()
    L = ctypes.CDLL(_build.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(L, s), f"{s} is declared in include/evoxb.h but not exported"
    assert sorted(_lib.EXPORTS) == declared_symbols()
    assert _lib.lib().evoxb_version() >= 100


def test_argument_validation_without_gpu():
    """Pure host-side checks: no kernel is launched for these calls."""
    from evox_myc_b200 import _lib

    L = _lib.lib()
    need = ctypes.c_size_t()
    assert L.evoxb_nds_workspace_bytes(262144, 5, ctypes.byref(need)) == 0
    # the packed matrix dominates: n * n / 8 bytes up to 262144 rows ...
    assert 262144 * 262144 // 8 <= need.value < 262144 * 262144 // 8 + (64 << 20)
    # ... and only its triangle beyond (row i keeps the chunks from its own on): n^2 / 16 + half a chunk column per row
    assert L.evoxb_nds_workspace_bytes(524288, 5, ctypes.byref(need)) == 0
    tri = 524288 * 524288 // 16
    assert tri <= need.value < tri + 524288 * 4096 // 16 + (128 << 20)
    assert L.evoxb_nds_workspace_bytes(5000, 20, ctypes.byref(need)) == 0  # m > 16: full matrix in input order
    assert 5000 * 8192 // 8 <= need.value < 5000 * 8192 // 8 + (8 << 20)
    assert L.evoxb_nds_workspace_bytes(-1, 3, ctypes.byref(need)) == -1
    assert L.evoxb_nds_workspace_bytes(10, 0, ctypes.byref(need)) == -1
    assert L.evoxb_nds_workspace_bytes(10, 33, ctypes.byref(need)) == -1
    assert b"m must be" in L.evoxb_last_error()
    # n == 0 is a no-op that must not touch the device
    assert L.evoxb_nds_rank_f32(None, 0, 3, None, None, 0, None) == 0
    # missing workspace is reported, not dereferenced
    assert L.evoxb_nds_rank_f32(ctypes.c_void_p(256), 8, 3, ctypes.c_void_p(256), None, 0, None) == -2


def test_product_has_no_cpu_fallback():
    import numpy as np
    import torch

    import evox_myc_b200 as e

    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        e.non_dominated_sort(np.zeros((4, 2), np.float32))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "evox-myc_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f
                assert "liboracle" not in text, f


def test_jax_binding_says_why_it_cannot_load():
    """The JAX side is shipped but cannot run here (no jax): importing it must say so, not fail obscurely."""
    try:
        import jax  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError, match="needs jax/jaxlib"):
            import evox_myc_b200.jax_ffi  # noqa: F401
    else:  # pragma: no cover
        pytest.skip("jax is installed: the binding can be exercised for real")
