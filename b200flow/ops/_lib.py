"""ctypes binding to the in-tree native libraries.

The CUDA library is mandatory on a GPU box: if it is missing, or a kernel call
fails, the op raises — there is no silent PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIB_DIR = Path(__file__).resolve().parent.parent / "lib"
_cuda = None
_host = None

# The 16-bit type of activations / weights on the GPU, process-wide: bf16 (default) or IEEE fp16 (`B2F_HALF=fp16`, set by
# the front ends when a job asks for fp16: `"fp16": {"enabled": true}`, --fp16, mixed_precision="fp16").  The two are
# separate builds of the same kernels (b200flow/build.py), chosen when the library is first loaded.
HALF_NAME = os.environ.get("B2F_HALF", "bf16").lower()
assert HALF_NAME in ("bf16", "fp16"), f"B2F_HALF must be bf16 or fp16, not {HALF_NAME!r}"
HALF = torch.float16 if HALF_NAME == "fp16" else torch.bfloat16
_CUDA_LIB_NAME = os.environ.get("B2F_CUDA_LIB") or (
    "libb200flow_cuda_fp16.so" if HALF_NAME == "fp16" else "libb200flow_cuda.so")   # B2F_CUDA_LIB: an instrumented variant


def set_half(name: str) -> bool:
    """Select the 16-bit type before the first kernel call.  Returns False (and changes nothing) when the other build
    is already loaded in this process - the choice is per process; callers then keep computing in the loaded type."""
    global HALF_NAME, HALF, _CUDA_LIB_NAME
    name = name.lower()
    if name == HALF_NAME:
        return True
    if _cuda is not None:
        import warnings

        warnings.warn(f"b200flow: {name} kernels were asked for but the {HALF_NAME} build is already loaded in this "
                      f"process; continuing in {HALF_NAME} (start the job with B2F_HALF={name})")
        return False
    assert name in ("bf16", "fp16")
    HALF_NAME, HALF = name, (torch.float16 if name == "fp16" else torch.bfloat16)
    _CUDA_LIB_NAME = "libb200flow_cuda_fp16.so" if name == "fp16" else "libb200flow_cuda.so"
    os.environ["B2F_HALF"] = name
    return True


def half_dtype() -> torch.dtype:
    return HALF


class NativeError(RuntimeError):
    pass


def _load(name: str) -> ctypes.CDLL:
    path = _LIB_DIR / name
    if not path.exists():
        # Build on demand (nvcc is available on both the dev box and the GPU box).
        from .. import build as _build

        if "cuda" in name:
            _build.build_cuda(half="fp16" if "fp16" in name else "bf16")
        else:
            _build.build_host()
    if not path.exists():
        raise NativeError(f"native library {path} is missing; run `python -m b200flow.build`")
    return ctypes.CDLL(str(path))


def cuda_lib() -> ctypes.CDLL:
    global _cuda
    if _cuda is None:
        _cuda = _load(_CUDA_LIB_NAME)
        _cuda.b2f_last_error.restype = ctypes.c_char_p
    return _cuda


def host_lib() -> ctypes.CDLL:
    global _host
    if _host is None:
        _host = _load("libb200flow_host.so")
    return _host


def loaded_libraries() -> list[str]:
    return [str(_LIB_DIR / n) for n, l in ((_CUDA_LIB_NAME, _cuda), ("libb200flow_host.so", _host)) if l]


_launches = 0
call_hook = None      # optional callable(fn_name): observers of native entry points (obs/comms.py)
# kernels launched per entry point (default 1)
_NVTX = os.environ.get("B2F_NVTX", "0") == "1"
_KERNELS_PER_CALL = {"b2f_norm_bwd": 2, "b2f_norm_bwd_noalloc": 2, "b2f_sumsq": 2, "b2f_attn_bwd": 3,
                     "b2f_moe_router_top1": 2}


def launch_count() -> int:
    """Number of b200flow CUDA kernels launched through this binding."""
    return _launches


def reset_launch_count() -> None:
    global _launches
    _launches = 0


c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_ull = ctypes.c_ulonglong
c_float = ctypes.c_float


def ptr(t: torch.Tensor | None) -> c_void_p:
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream_ptr() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def call(fn_name: str, *args) -> None:
    """Call `int fn(...)` in the CUDA library and raise with its message on failure."""
    global _launches
    lib = cuda_lib()
    fn = getattr(lib, fn_name)
    if _NVTX:       # B2F_NVTX=1: one NVTX range per native entry point (shows up in nsys / torch.profiler)
        torch.cuda.nvtx.range_push(fn_name)
        try:
            rc = fn(*args)
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        rc = fn(*args)
    _launches += _KERNELS_PER_CALL.get(fn_name, 1)
    if call_hook is not None:
        call_hook(fn_name)
    if rc != 0:
        msg = lib.b2f_last_error()
        raise NativeError(f"{fn_name} failed (rc={rc}): {msg.decode() if msg else '?'}")


def call_int(fn_name: str, *args) -> int:
    fn = getattr(cuda_lib(), fn_name)
    fn.restype = ctypes.c_int
    return int(fn(*args))


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NativeError("b200flow ops run on sm_100a only; got a CPU tensor "
                              "(CPU oracles live in b200flow.ops.ref)")


def debug_sync() -> None:
    if os.environ.get("B2F_DEBUG_SYNC"):
        torch.cuda.synchronize()
