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


def _nds_call(x):
    return jax.ffi.ffi_call("evoxb_nds", jax.ShapeDtypeStruct((x.shape[0],), jnp.int32))(x)


def _crowding_call(costs, mask):
    return jax.ffi.ffi_call("evoxb_crowding", jax.ShapeDtypeStruct((costs.shape[0],), jnp.float32))(costs, mask)


# ---- partitioning rule (jit over a device mesh, e.g. StdWorkflow under `jax.sharding`) -----------------------
# A custom call is opaque to the SPMD partitioner; without a rule it refuses sharded operands.  The ranking of one
# population is a whole-population computation, so the rule is the simplest correct one: every operand and the
# result are REPLICATED (each device ranks the full population, exactly what the reference does today:
# `tell` runs replicated per device, workflows/std_workflow.py:194,341).  Sharding the matrix by column chunk over
# the devices is what `evoxb_nds_rank_sharded_f32` does; binding that needs a communicator per process and is the
# next step after these handlers have been compiled once.  Like the rest of this file: never run here.
def _replicated(mesh, ndim):
    from jax.sharding import NamedSharding, PartitionSpec

    return NamedSharding(mesh, PartitionSpec(*([None] * ndim)))


def _with_replication_rule(fn, result_ndim):
    from jax.experimental.custom_partitioning import custom_partitioning

    wrapped = custom_partitioning(fn)

    def infer_sharding_from_operands(mesh, arg_shapes, result_shape):
        return _replicated(mesh, result_ndim)

    def partition(mesh, arg_shapes, result_shape):
        arg_shardings = tuple(_replicated(mesh, len(a.shape)) for a in arg_shapes)
        return mesh, fn, _replicated(mesh, result_ndim), arg_shardings

    wrapped.def_partition(infer_sharding_from_operands=infer_sharding_from_operands, partition=partition)
    return wrapped


_nds_partitioned = None
_crowding_partitioned = None


def non_dominated_sort(x, method="auto"):
    """Same signature as the reference (non_dominate.py:30-31); traceable under jit / cond / pmap, and under a
    sharded jit through the replication rule above."""
    global _nds_partitioned
    assert method in ["auto", "scan", "full map-reduce", "host"], "method must be either 'auto', 'scan', or 'full map-reduce', 'host'"
    _register()
    if _nds_partitioned is None:
        _nds_partitioned = _with_replication_rule(_nds_call, 1)
    return _nds_partitioned(x.astype(jnp.float32))


def crowding_distance(costs, mask=None):
    """Same signature as the reference (non_dominate.py:116-117)."""
    global _crowding_partitioned
    _register()
    if mask is None:
        mask = jnp.ones((costs.shape[0],), dtype=jnp.bool_)
    if _crowding_partitioned is None:
        _crowding_partitioned = _with_replication_rule(_crowding_call, 1)
    return _crowding_partitioned(costs.astype(jnp.float32), mask.astype(jnp.bool_))
