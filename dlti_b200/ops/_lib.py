"""ctypes loader for the native libraries (C ABI in include/dlti_b200.h).

The GPU library is REQUIRED whenever CUDA is available: if it is missing we try to build it
in-tree (nvcc cross-compiles without a GPU) and otherwise fail loudly — there is no silent
PyTorch fallback on a GPU box. On a CPU-only box the ops use the pure-PyTorch oracles.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_PKG = Path(__file__).resolve().parent.parent
_GPU_LIB_PATH = _PKG / "lib" / "libdlti_b200.so"
_CPU_LIB_PATH = _PKG / "lib" / "libdlti_cpu.so"

_gpu_lib = None
_cpu_lib = None

DT_BF16, DT_FP16, DT_FP32 = 0, 1, 2
_DT = {torch.bfloat16: DT_BF16, torch.float16: DT_FP16, torch.float32: DT_FP32}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError(f"unsupported dtype for dlti kernels: {dt}") from None


class DltiError(RuntimeError):
    pass


def _try_build(which: str) -> None:
    from .. import build as _b

    if which == "gpu":
        _b.build_gpu()
    else:
        _b.build_cpu()


def gpu_lib() -> C.CDLL:
    """Load (building if needed) libdlti_b200.so. Raises if unavailable."""
    global _gpu_lib
    if _gpu_lib is None:
        if not _GPU_LIB_PATH.exists() or os.environ.get("DLTI_REBUILD"):
            _try_build("gpu")
        if not _GPU_LIB_PATH.exists():
            raise DltiError(f"native GPU library missing: {_GPU_LIB_PATH} (run python -m dlti_b200.build)")
        # DLTI_GPU_LIB: load an alternative build of the library (A/B experiments with compile-time knobs)
        lib = C.CDLL(os.environ.get("DLTI_GPU_LIB") or str(_GPU_LIB_PATH))
        lib.dlti_last_error.restype = C.c_char_p
        lib.dlti_error_string.restype = C.c_char_p
        lib.dlti_error_string.argtypes = [C.c_int]
        lib.dlti_launch_count.restype = C.c_longlong
        _gpu_lib = lib
    return _gpu_lib


def cpu_lib() -> C.CDLL:
    global _cpu_lib
    if _cpu_lib is None:
        if not _CPU_LIB_PATH.exists() or os.environ.get("DLTI_REBUILD"):
            _try_build("cpu")
        if not _CPU_LIB_PATH.exists():
            raise DltiError(f"native CPU library missing: {_CPU_LIB_PATH}")
        _cpu_lib = C.CDLL(str(_CPU_LIB_PATH))
    return _cpu_lib


def have_gpu_kernels() -> bool:
    """True when tensors on CUDA must go through the native kernels."""
    return torch.cuda.is_available()


def use_native(*tensors: torch.Tensor) -> bool:
    """Dispatch rule: CUDA tensors -> native kernels (mandatory); CPU tensors -> PyTorch oracle."""
    return any(t is not None and t.is_cuda for t in tensors)


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        lib = gpu_lib()
        msg = lib.dlti_error_string(rc).decode()
        detail = lib.dlti_last_error().decode()
        raise DltiError(f"{what} failed: {msg} (code {rc}) {detail}")


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: torch.Tensor | None) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def launch_count() -> int:
    return int(gpu_lib().dlti_launch_count()) if _GPU_LIB_PATH.exists() else 0


def reset_launch_count() -> None:
    if _GPU_LIB_PATH.exists():
        gpu_lib().dlti_reset_launch_count()


_bound: dict[str, object] = {}


def fn(name: str, argtypes: list, restype=C.c_int, cpu: bool = False):
    """Bind and cache a C function with explicit argtypes."""
    key = ("cpu:" if cpu else "gpu:") + name
    f = _bound.get(key)
    if f is None:
        lib = cpu_lib() if cpu else gpu_lib()
        f = getattr(lib, name)
        f.argtypes = argtypes
        f.restype = restype
        _bound[key] = f
    return f
