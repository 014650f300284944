"""Dependent skinny GEMMs and the LayerNorms between them as one persistent kernel (``csrc/chain_sm100.cu``).

A :class:`Chain` collects ops in program order and launches them with one ``pb2_chain_run``; every GEMM takes the
arguments of :func:`pipe_b200.ops.gemm.gemm_nt` (same fused epilogue).  The Transformer-layer executor in
``csrc/layer.cu`` builds its chains natively; this wrapper exists for tests, benchmarks and custom stage bodies.

Reference counterpart: none (separate cuBLAS / ATen calls under ``nn.TransformerEncoderLayer``, ``main.py:148``).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
from torch import Tensor

from . import _lib
from .gemm import workspace

__all__ = ["Chain", "ChainLn", "ChainOp"]

MAX_OPS = 8


class ChainLn(ctypes.Structure):
    _fields_ = [("x", ctypes.c_void_p), ("h", ctypes.c_void_p), ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
                ("y", ctypes.c_void_p), ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
                ("rows", ctypes.c_int), ("E", ctypes.c_int), ("eps", ctypes.c_float)]


class ChainOp(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int), ("gemm", _lib.GemmDesc), ("ln", ChainLn)]


class Chain:
    def __init__(self, device: torch.device) -> None:
        self.device = torch.device(device)
        self.ops = (ChainOp * MAX_OPS)()
        self.n = 0
        self._keep: List[Tensor] = []

    def gemm(self, a: Tensor, b: Tensor, out: Tensor, *, bias: Optional[Tensor] = None, relu: bool = False,
             drop_p: float = 0.0, seed: int = 0, rng_stream: int = 0, mask_src: Optional[Tensor] = None,
             mask_scale: float = 1.0, addend: Optional[Tensor] = None, a_mn: bool = False) -> "Chain":
        """``out[tokens, M] = b[tokens, K] @ a[M, K]^T`` (+ epilogue); ``a_mn``: ``a`` is ``[K, M]``."""
        assert self.n < MAX_OPS
        op = self.ops[self.n]
        op.kind = 0
        d = op.gemm
        if a_mn:
            K, M = a.shape
        else:
            M, K = a.shape
        assert b.size(1) == K and out.shape == (b.size(0), M) and out.dtype == torch.bfloat16
        d.a, d.b = a.data_ptr(), b.data_ptr()
        d.M, d.N, d.K = M, b.size(0), K
        d.lda, d.ldb = a.stride(0), b.stride(0)
        d.out, d.ld_out = out.data_ptr(), out.stride(0)
        d.out_f32, d.transpose_out = 0, 1
        d.bias, d.bias_along_m, d.relu = _lib.ptr(bias), 1, int(relu)
        d.drop_p = float(drop_p)
        d.seed, d.rng_stream = seed & (2**64 - 1), rng_stream & (2**64 - 1)
        d.mask_src, d.mask_scale, d.addend = _lib.ptr(mask_src), float(mask_scale), _lib.ptr(addend)
        d.alpha = 1.0
        d.a_static, d.a_mn_major = 1, int(a_mn)
        self._keep += [a, b, out]
        self.n += 1
        return self

    def layernorm(self, x: Tensor, h: Optional[Tensor], gamma: Tensor, beta: Tensor, y: Tensor, mean: Tensor,
                  rstd: Tensor, eps: float = 1e-5) -> "Chain":
        assert self.n < MAX_OPS
        op = self.ops[self.n]
        op.kind = 1
        l = op.ln
        l.x, l.h, l.gamma, l.beta = x.data_ptr(), _lib.ptr(h), gamma.data_ptr(), beta.data_ptr()
        l.y, l.mean, l.rstd = y.data_ptr(), mean.data_ptr(), rstd.data_ptr()
        l.rows, l.E, l.eps = x.size(0), x.size(1), float(eps)
        self._keep += [x, y]
        self.n += 1
        return self

    def supported(self) -> bool:
        return bool(_lib.load().pb2_chain_supported(self.ops, ctypes.c_int(self.n)))

    def run(self, stream: Optional[int] = None) -> None:
        if stream is None:
            stream = _lib.raw_stream(self.device)
        ws = workspace(self.device, stream)
        with _lib.on_device(self.device):
            rc = _lib.load().pb2_chain_run(self.ops, ctypes.c_int(self.n), ctypes.c_void_p(ws.data_ptr()),
                                           ctypes.c_size_t(ws.numel()), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_chain_run")
