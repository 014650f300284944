"""ctypes binding of ``libpipe_b200.so`` (C ABI declared in ``include/pipe_b200.h``).

The reference has no native code of its own; this is the boundary between the PyTorch
framework layer and the hand-written sm_100a kernels."""
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
        ("reduce_mode", ctypes.c_int),
        ("a_static", ctypes.c_int),
        ("persistent", ctypes.c_int),
        ("a_mn_major", ctypes.c_int),
    ]


_vp, _ll, _ci, _cf = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_float


class EncoderLayerDesc(ctypes.Structure):
    """Mirror of ``pb2_encoder_layer`` (include/pipe_b200.h)."""
    _fields_ = (
        [(n, _ci) for n in ("T", "NB", "S", "E", "heads", "FF", "causal")]
        + [("drop_p", _cf), ("eps", _cf), ("seed", ctypes.c_ulonglong)]
        + [(n, _vp) for n in ("w_qkv", "w_o", "w_1", "w_2", "wt_qkv", "wt_o", "wt_1", "wt_2")]
        + [(n, _ll) for n in ("ldt_qkv", "ldt_o", "ldt_1", "ldt_2")]
        + [(n, _vp) for n in ("b_qkv", "b_o", "b_1", "b_2", "g1", "be1", "g2", "be2")]
        + [(n, _vp) for n in ("x", "qkv", "a", "o", "x1", "hmid", "f", "y", "mean1", "rstd1", "mean2", "rstd2")]
        + [(n, _vp) for n in ("dy", "dx", "dr2", "df", "dz1", "dx1", "dr1", "do_", "da", "dqkv")]
        + [(n, _vp) for n in ("dg1", "dbe1", "dg2", "dbe2", "db_qkv", "db_o", "db_1", "db_2")]
        + [(n, _vp) for n in ("xt_qkv", "dyt_qkv", "xt_o", "dyt_o", "xt_1", "dyt_1", "xt_2", "dyt_2")]
        + [(n, _ll) for n in ("ld_stash_qkv", "ld_stash_o", "ld_stash_1", "ld_stash_2")]
        + [(n, _ci) for n in ("stash_off_qkv", "stash_off_o", "stash_off_1", "stash_off_2")]
        + [("workspace", _vp), ("workspace_bytes", ctypes.c_size_t)]
        + [("next_weight", _vp), ("next_weight_bytes", ctypes.c_size_t), ("prefetch", _ci), ("side_stream", _vp)]
        + [("n_peers", _ci), ("peer_delta", _ll * 7)]
        + [("attn_lse", _vp), ("attn_scratch", _vp), ("attn_scratch_bytes", ctypes.c_size_t)]
    )


class EncoderStackDesc(ctypes.Structure):
    """Mirror of ``pb2_encoder_stack`` (include/pipe_b200.h)."""
    _fields_ = [
        ("n_layers", _ci), ("layers", _vp), ("T", _ci), ("NB", _ci), ("S", _ci), ("drop_p", _cf),
        ("seeds", _vp), ("x", _vp), ("y", _vp), ("act", _vp), ("stats", _vp), ("dy", _vp), ("dx", _vp),
        ("tmp", _vp), ("stash_off", _ci), ("workspace", _vp), ("workspace_bytes", ctypes.c_size_t),
        ("side_stream", _vp), ("attn_scratch", _vp), ("attn_scratch_bytes", ctypes.c_size_t),
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
        lib.pb2_encoder_stack_act_elems.restype = ctypes.c_size_t
        lib.pb2_flash_attention_bwd_scratch_bytes.restype = ctypes.c_size_t
        lib.pb2_encoder_stack_tmp_elems.restype = ctypes.c_size_t
        lib.pb2_encoder_stack_stats_floats.restype = ctypes.c_size_t
        _lib = lib
    return _lib


def available() -> bool:
    return os.path.exists(LIB_PATH)


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().pb2_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed: {msg}")


def raw_stream(device: torch.device) -> int:
    """The current CUDA stream of ``device`` as an integer handle (one C call; the
    ``torch.cuda.current_stream()`` object path costs several microseconds)."""
    idx = device.index
    if idx is None:
        idx = torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(idx)


def stream_ptr(device: Optional[torch.device] = None) -> ctypes.c_void_p:
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return ctypes.c_void_p(raw_stream(device))


class _Noop:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NOOP = _Noop()


def on_device(device: torch.device):
    """Context that makes ``device`` current for a raw launch; free when it already is (the usual
    case: pipeline workers and autograd threads run with their stage's device current)."""
    idx = device.index
    if idx is None or torch._C._cuda_getDevice() == idx:
        return _NOOP
    return torch.cuda.device(idx)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()
