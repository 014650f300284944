"""ctypes binding of ``libpipe_b200.so`` (C ABI declared in ``include/pipe_b200.h``)."""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_native", "libpipe_b200.so")

_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()


class NativeUnavailable(RuntimeError):
    pass


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("a", ctypes.c_void_p), ("b", ctypes.c_void_p),
        ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int),
        ("lda", ctypes.c_longlong), ("ldb", ctypes.c_longlong),
        ("out", ctypes.c_void_p), ("ld_out", ctypes.c_longlong),
        ("out_f32", ctypes.c_int), ("transpose_out", ctypes.c_int),
        ("bias", ctypes.c_void_p), ("bias_along_m", ctypes.c_int), ("relu", ctypes.c_int),
        ("drop_p", ctypes.c_float),
        ("seed", ctypes.c_ulonglong), ("rng_stream", ctypes.c_ulonglong),
        ("mask_src", ctypes.c_void_p), ("mask_scale", ctypes.c_float),
        ("addend", ctypes.c_void_p),
        ("accumulate", ctypes.c_int), ("alpha", ctypes.c_float),
        ("sumsq", ctypes.c_void_p),
        ("block_n", ctypes.c_int), ("splits", ctypes.c_int), ("stages", ctypes.c_int),
    ]


def load() -> ctypes.CDLL:
    """Load the native library (once).  Raises :class:`NativeUnavailable` with a
    build hint when it is missing — never falls back silently."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NativeUnavailable(
                f"{LIB_PATH} not found: build it with `python -m pipe_b200.ops._build` "
                "(or __graft_entry__.build())")
        lib = ctypes.CDLL(LIB_PATH)
        lib.pb2_last_error.restype = ctypes.c_char_p
        lib.pb2_launch_count.restype = ctypes.c_longlong
        lib.pb2_gemm_workspace_bytes.restype = ctypes.c_size_t
        _lib = lib
    return _lib


def available() -> bool:
    return os.path.exists(LIB_PATH)


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().pb2_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed: {msg}")


def stream_ptr(device: Optional[torch.device] = None) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()
