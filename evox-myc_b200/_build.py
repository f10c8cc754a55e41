"""Builds libevoxb.so in-tree with nvcc for sm_100a (no torch headers, plain C ABI)."""

from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
# EVOXB_LIB_PATH lets a developer load an experimental build; the product default is the in-tree .so
LIB_PATH = os.environ.get("EVOXB_LIB_PATH") or os.path.join(_HERE, "libevoxb.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-shared", "-cudart", "static", "-ldl",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libevoxb.so cannot be built")


def sources() -> list[str]:
    out = []
    for root, _, files in os.walk(CSRC):
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".h"))]
    out.append(os.path.join(os.path.dirname(_HERE), "include", "evoxb.h"))
    return out


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in sources() if os.path.exists(s))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/evoxb.cu -> libevoxb.so.  Cross-compiles without a GPU."""
    if not force and not is_stale():
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB_PATH + ".tmp", os.path.join(CSRC, "evoxb.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(LIB_PATH + ".tmp", LIB_PATH)
    if verbose:
        print(res.stderr)
    return LIB_PATH


def build_ffi() -> str:
    """Compile csrc/xla_ffi_handlers.cc -> libevoxb_ffi.so.  Needs jaxlib's XLA FFI headers, which are
    absent on this image (the function then raises); never called by build()."""
    try:
        import jax.ffi  # noqa: F401
    except Exception as exc:  # pragma: no cover
        raise RuntimeError("jax is not installed: the XLA FFI handlers cannot be built on this image") from exc
    inc = jax.ffi.include_dir()
    out = os.path.join(_HERE, "libevoxb_ffi.so")
    cmd = [_nvcc(), "-O2", "-std=c++17", "-Xcompiler", "-fPIC", "-shared", "-I", inc, "-o", out,
           os.path.join(CSRC, "xla_ffi_handlers.cc"), LIB_PATH, "-Xlinker", "-rpath", "-Xlinker", _HERE]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("building the XLA FFI handlers failed:\n" + res.stdout + res.stderr)
    return out


if __name__ == "__main__":
    import sys

    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
