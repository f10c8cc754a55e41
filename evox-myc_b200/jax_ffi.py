"""JAX binding of the path: ``jax.ffi`` custom calls into libevoxb.so (the drop-in behind StdWorkflow).

UNTESTED ON THIS IMAGE -- jax / jaxlib are not installed here and cannot be fetched, so neither this
module nor ``csrc/xla_ffi_handlers.cc`` has ever run.  It is the literal form of what INTEGRATION.md
section 2 describes; importing it without jax raises ImportError with that explanation.  Where jax
exists: build the handlers with ``python -c "from evox_myc_b200 import _build; _build.build_ffi()"``
and replace the two functions of ``evox.operators.selection.non_dominate`` with the ones below.
"""

from __future__ import annotations

try:
    import jax
    import jax.numpy as jnp
except ImportError as exc:  # pragma: no cover - jax is absent on this image
    raise ImportError(
        "evox_myc_b200.jax_ffi needs jax/jaxlib, which this image does not have; "
        "use evox_myc_b200.non_dominated_sort / crowding_distance (torch or NumPy buffers) instead"
    ) from exc

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_FFI_SO = os.path.join(_HERE, "libevoxb_ffi.so")
_registered = False


def _register():
    global _registered
    if _registered:
        return
    lib = ctypes.CDLL(_FFI_SO)  # links libevoxb.so; built by _build.build_ffi()
    jax.ffi.register_ffi_target("evoxb_nds", jax.ffi.pycapsule(lib.EvoxbNds), platform="CUDA")
    jax.ffi.register_ffi_target("evoxb_crowding", jax.ffi.pycapsule(lib.EvoxbCrowding), platform="CUDA")
    _registered = True


def non_dominated_sort(x, method="auto"):
    """Same signature as the reference (non_dominate.py:30-31); traceable under jit / cond / pmap."""
    assert method in ["auto", "scan", "full map-reduce", "host"], "method must be either 'auto', 'scan', or 'full map-reduce', 'host'"
    _register()
    return jax.ffi.ffi_call("evoxb_nds", jax.ShapeDtypeStruct((x.shape[0],), jnp.int32))(x.astype(jnp.float32))


def crowding_distance(costs, mask=None):
    """Same signature as the reference (non_dominate.py:116-117)."""
    _register()
    if mask is None:
        mask = jnp.ones((costs.shape[0],), dtype=jnp.bool_)
    return jax.ffi.ffi_call("evoxb_crowding", jax.ShapeDtypeStruct((costs.shape[0],), jnp.float32))(
        costs.astype(jnp.float32), mask.astype(jnp.bool_))
