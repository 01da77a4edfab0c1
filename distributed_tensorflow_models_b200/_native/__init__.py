"""ctypes binding of ``libdtfb200.so`` (the hand-written sm_100a kernels; C ABI in ``include/dtf_b200.h``).

The library is loaded lazily.  On a machine with a GPU a missing or stale library is rebuilt
with ``nvcc`` when possible and is otherwise a hard error — GPU code paths never fall back
silently to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libdtfb200.so"

_lib = None
_lock = threading.Lock()

c_float_p = C.POINTER(C.c_float)


class ConvGemmParams(C.Structure):
    """Mirror of ``dtf_conv_gemm_params``."""
    _fields_ = [
        ("src", C.c_void_p), ("srcH", C.c_int), ("srcW", C.c_int), ("srcC", C.c_int), ("src_ld", C.c_int),
        ("N", C.c_int), ("OH", C.c_int), ("OW", C.c_int),
        ("R", C.c_int), ("S", C.c_int),
        ("a_h", C.c_int), ("b_h", C.c_int), ("c_h", C.c_int), ("div_h", C.c_int),
        ("a_w", C.c_int), ("b_w", C.c_int), ("c_w", C.c_int), ("div_w", C.c_int),
        ("wt", C.c_void_p),
        ("Nout", C.c_int),
        ("out", C.c_void_p), ("out_ld", C.c_int),
        ("in_scale", C.c_void_p), ("in_shift", C.c_void_p), ("in_relu", C.c_int),
        ("bias", C.c_void_p), ("residual", C.c_void_p), ("res_ld", C.c_int), ("out_relu", C.c_int),
        ("stats", C.c_void_p),
        ("bnb_x", C.c_void_p), ("bnb_ld", C.c_int), ("bnb_scale", C.c_void_p), ("bnb_shift", C.c_void_p),
        ("act_out", C.c_void_p),
        ("fin_stats", C.c_void_p), ("fin_gamma", C.c_void_p), ("fin_beta", C.c_void_p),
        ("fin_moving_mean", C.c_void_p), ("fin_moving_var", C.c_void_p), ("fin_mean", C.c_void_p),
        ("fin_invstd", C.c_void_p),
        ("fin_inv_count", C.c_float), ("fin_unbias", C.c_float), ("fin_eps", C.c_float), ("fin_decay", C.c_float),
        ("fin_update_moving", C.c_int),
        ("gate_flags", C.c_void_p), ("gate_n", C.c_int), ("gate_epoch", C.c_void_p), ("gate_status", C.c_void_p),
        ("gate_timeout_cycles", C.c_longlong),
    ]


class ConvWgradParams(C.Structure):
    """Mirror of ``dtf_conv_wgrad_params``."""
    _fields_ = [
        ("src", C.c_void_p), ("srcH", C.c_int), ("srcW", C.c_int), ("srcC", C.c_int), ("src_ld", C.c_int),
        ("N", C.c_int), ("OH", C.c_int), ("OW", C.c_int),
        ("R", C.c_int), ("S", C.c_int),
        ("stride_h", C.c_int), ("stride_w", C.c_int), ("pad_h", C.c_int), ("pad_w", C.c_int),
        ("dy", C.c_void_p), ("dy_ld", C.c_int), ("Nout", C.c_int),
        ("dw", C.c_void_p),
        ("in_scale", C.c_void_p), ("in_shift", C.c_void_p), ("in_relu", C.c_int),
    ]


def available() -> bool:
    """True when the shared object exists (it may still be unusable without a GPU)."""
    return LIB_PATH.exists()


def build_library(force: bool = False) -> Path:
    from . import build as _b
    return _b.build(force=force)


def _declare(lib: C.CDLL) -> None:
    lib.dtf_version.restype = C.c_int
    lib.dtf_last_error.restype = C.c_char_p
    # every other entry point returns int; arguments are passed as explicit ctypes values by the wrappers


def lib() -> C.CDLL:
    """Load (building first if necessary) and return the native library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not LIB_PATH.exists() or os.environ.get("DTF_REBUILD") == "1":
            build_library(force=os.environ.get("DTF_REBUILD") == "1")
        l = C.CDLL(str(LIB_PATH))
        _declare(l)
        if os.environ.get("DTF_PDL") == "0":          # plain stream order instead of programmatic dependent launch
            l.dtf_debug_set_pdl(0)
        _lib = l
    return _lib


class NativeError(RuntimeError):
    pass


_launches = [0]


def launch_count() -> int:
    """Number of native entry-point calls (each launches one kernel) since ``reset_launch_count``."""
    return _launches[0]


def reset_launch_count() -> None:
    _launches[0] = 0


def check(rc: int, what: str) -> None:
    _launches[0] += 1
    if rc != 0:
        msg = lib().dtf_last_error()
        raise NativeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


_pre_launch = [None]


def set_pre_launch(fn) -> None:
    """Install (or clear, with ``None``) a callable that runs in front of every native launch.  The symmetric-memory
    transport uses it to put parameter-server pull gates that no convolution has taken over ahead of whatever kernel
    comes next (parallel/symm.py)."""
    _pre_launch[0] = fn


def run_pre_launch() -> None:
    """What a native launch would trigger, for code that is about to read parameters with a PyTorch kernel."""
    if _pre_launch[0] is not None:
        fn, _pre_launch[0] = _pre_launch[0], None
        try:
            fn()
        finally:
            _pre_launch[0] = fn


def stream_ptr(hook: bool = True) -> C.c_void_p:
    """Current torch CUDA stream as a ``void*`` for the C ABI (every launch wrapper fetches it right before its
    call, which makes this the one place that sees all native launches)."""
    import torch
    if hook and _pre_launch[0] is not None:
        fn, _pre_launch[0] = _pre_launch[0], None        # the hook launches kernels itself: no re-entry
        try:
            fn()
        finally:
            _pre_launch[0] = fn
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
