"""Python face of the tcgen05 GEMM (``pb2_gemm_bf16``; ``csrc/gemm_sm100.cu``).

``gemm_nt(a, b)`` computes ``a @ b.T`` for bf16 ``a [M, K]`` and ``b [N, K]`` with
fp32 accumulation in TMEM.  ``a`` rides the 128-wide UMMA M axis and ``b`` the
flexible N axis, so callers put the *larger* dimension (usually the weight) in
``a`` and ask for a transposed store (``transpose_out=True``) to get the usual
``[tokens, features]`` result — see :func:`linear_forward`.

Reference counterpart: the cuBLAS GEMMs behind ``nn.Linear`` / ``nn.MultiheadAttention`` (``main.py:45,148``);
SURVEY.md K2 / K4 / K6 / K7, W5.
"""
from __future__ import annotations

import ctypes
import threading
from typing import Dict, Optional

import torch
from torch import Tensor

from . import _lib

__all__ = ["gemm_nt", "linear_forward", "workspace"]

_ws: Dict[tuple, Tensor] = {}
_ws_lock = threading.Lock()


def workspace(device: torch.device, stream: Optional[int] = None) -> Tensor:
    """Per-(device, stream) split-K scratch: zero-initialised once, then
    self-cleaning (the kernel re-arms its counters).  Keyed by stream too, because
    two streams running split-K GEMMs at once must not share partial tiles."""
    key = (device.index, _lib.raw_stream(device) if stream is None else stream)
    ws = _ws.get(key)
    if ws is None:
        with _ws_lock:
            ws = _ws.get(key)
            if ws is None:
                nbytes = int(_lib.load().pb2_gemm_workspace_bytes())
                ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
                _ws[key] = ws
    return ws


def gemm_nt(a: Tensor, b: Tensor, *, out: Optional[Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
            transpose_out: bool = False, bias: Optional[Tensor] = None, bias_along_m: bool = False,
            relu: bool = False, drop_p: float = 0.0, seed: int = 0, rng_stream: int = 0,
            mask_src: Optional[Tensor] = None, mask_scale: float = 1.0, addend: Optional[Tensor] = None,
            accumulate: bool = False, alpha: float = 1.0, sumsq: Optional[Tensor] = None,
            block_n: int = 0, splits: int = 0, stages: int = 0, reduce_mode: int = 0, a_static: bool = False, persistent: int = 0,
            a_mn: bool = False, stream: Optional[int] = None) -> Tensor:
    """``a_mn=True``: ``a`` is given transposed, ``[K, M]`` with M contiguous (computes ``a.T @ b.T``): the tensor core
    reads it as an MN-major operand — a dgrad GEMM passes the forward weight ``[out, in]`` as it is."""
    lib = _lib.load()
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "gemm_nt takes bf16 operands"
    assert a.dim() == 2 and b.dim() == 2 and (a.size(0) if a_mn else a.size(1)) == b.size(1)
    assert a.stride(1) == 1 and b.stride(1) == 1, "operands must be contiguous along their last dimension"
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    N = b.size(0)
    if out is None:
        shape = (N, M) if transpose_out else (M, N)
        out = torch.empty(shape, dtype=out_dtype, device=a.device)
    assert out.stride(-1) == 1
    assert out.dtype in (torch.bfloat16, torch.float32)
    for t in (mask_src, addend):
        if t is not None:
            assert t.dtype == torch.bfloat16 and t.stride() == out.stride() and t.shape == out.shape
    if bias is not None:
        assert bias.dtype == torch.float32

    d = _lib.GemmDesc()
    d.a, d.b = a.data_ptr(), b.data_ptr()
    d.M, d.N, d.K = M, N, K
    d.lda, d.ldb = a.stride(0), b.stride(0)
    d.out, d.ld_out = out.data_ptr(), out.stride(0)
    d.out_f32 = int(out.dtype == torch.float32)
    d.transpose_out = int(transpose_out)
    d.bias = _lib.ptr(bias)
    d.bias_along_m = int(bias_along_m)
    d.relu = int(relu)
    d.drop_p = float(drop_p)
    d.seed, d.rng_stream = seed & (2**64 - 1), rng_stream & (2**64 - 1)
    d.mask_src = _lib.ptr(mask_src)
    d.mask_scale = float(mask_scale)
    d.addend = _lib.ptr(addend)
    d.accumulate = int(accumulate)
    d.alpha = float(alpha)
    d.sumsq = _lib.ptr(sumsq)
    d.block_n, d.splits, d.stages, d.reduce_mode = block_n, splits, stages, reduce_mode
    d.a_static = int(a_static)
    d.persistent = persistent
    d.a_mn_major = int(a_mn)

    if stream is None:
        stream = _lib.raw_stream(a.device)
    ws = workspace(a.device, stream)
    with _lib.on_device(a.device):
        rc = lib.pb2_gemm_bf16(ctypes.byref(d), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(ws.numel()),
                               ctypes.c_void_p(stream))
    _lib.check(rc, "pb2_gemm_bf16")
    return out


def linear_forward(x: Tensor, weight: Tensor, bias: Optional[Tensor] = None, **epilogue) -> Tensor:
    """``x @ weight.T + bias`` for ``x [tokens, in]`` and ``weight [out, in]``
    (bf16), with the weight on the UMMA M axis (swap-AB)."""
    return gemm_nt(weight, x, transpose_out=True, bias=bias, bias_along_m=True, a_static=True, **epilogue)
