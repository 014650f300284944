"""Thin, shape-checked Python wrappers over the C entry points (no autograd here).

No counterpart in the reference, whose hot operations are ``torch.nn`` call sites (``main.py:24-73,148,182-184``;
SURVEY.md §2.5 K1-K11): each wrapper here is the sm_100a replacement of one of those library calls."""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import _lib
from ..autograd import seeds as _seeds
from ._lib import ptr

__all__ = ["add_layernorm_fwd", "add_layernorm_bwd", "embed_fwd", "embed_bwd", "cross_entropy",
           "transpose_colsum", "transpose_multi", "sumsq", "attention_fwd", "attention_bwd", "draw_seed"]

c_f = ctypes.c_float
c_i = ctypes.c_int
c_ll = ctypes.c_longlong
c_ull = ctypes.c_ulonglong
c_p = ctypes.c_void_p


def draw_seed() -> int:
    """A 63-bit seed for one fused-dropout call site.  Inside a checkpointed cell the seeds come from
    the cell's own base (:mod:`pipe_b200.autograd.seeds`), so a recompute — on whichever thread —
    regenerates bit-identical masks; elsewhere it is a locked draw from torch's CPU generator."""
    return _seeds.draw_seed()


def _s(t: Tensor) -> c_p:
    return _lib.stream_ptr(t.device)


def _bf16_rows(t: Tensor, E: int) -> None:
    assert t.dtype == torch.bfloat16 and t.is_contiguous() and t.size(-1) == E, (t.dtype, t.shape, t.stride())


def add_layernorm_fwd(x: Tensor, h: Optional[Tensor], gamma: Tensor, beta: Tensor, eps: float
                      ) -> Tuple[Tensor, Tensor, Tensor]:
    E = x.size(-1)
    rows = x.numel() // E
    _bf16_rows(x, E)
    if h is not None:
        _bf16_rows(h, E)
    y = torch.empty_like(x)
    mean = torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    with _lib.on_device(x.device):
        rc = _lib.load().pb2_add_layernorm_fwd(c_p(x.data_ptr()), c_p(ptr(h)), c_p(gamma.data_ptr()),
                                               c_p(beta.data_ptr()), c_p(y.data_ptr()), c_p(mean.data_ptr()),
                                               c_p(rstd.data_ptr()), c_i(rows), c_i(E), c_f(eps), _s(x))
    _lib.check(rc, "pb2_add_layernorm_fwd")
    return y, mean, rstd


def add_layernorm_bwd(dy: Tensor, x: Tensor, h: Optional[Tensor], gamma: Tensor, mean: Tensor, rstd: Tensor,
                      dgamma: Tensor, dbeta: Tensor, *, want_dh: bool, drop_p: float = 0.0, seed: int = 0,
                      rng_stream: int = 0) -> Tuple[Tensor, Optional[Tensor]]:
    E = x.size(-1)
    rows = x.numel() // E
    _bf16_rows(dy, E)
    dres = torch.empty_like(x)
    dh = torch.empty_like(x) if want_dh else None
    assert dgamma.dtype == torch.float32 and dbeta.dtype == torch.float32
    with _lib.on_device(x.device):
        rc = _lib.load().pb2_add_layernorm_bwd(c_p(dy.data_ptr()), c_p(x.data_ptr()), c_p(ptr(h)),
                                               c_p(gamma.data_ptr()), c_p(mean.data_ptr()), c_p(rstd.data_ptr()),
                                               c_p(dres.data_ptr()), c_p(ptr(dh)), c_p(dgamma.data_ptr()),
                                               c_p(dbeta.data_ptr()), c_i(rows), c_i(E), c_f(drop_p), c_ull(seed),
                                               c_ull(rng_stream), _s(x))
    _lib.check(rc, "pb2_add_layernorm_bwd")
    return dres, dh


def embed_fwd(tokens: Tensor, weight: Tensor, pe: Tensor, scale: float, drop_p: float, seed: int,
              rng_stream: int) -> Tensor:
    assert tokens.dtype == torch.int64 and tokens.dim() == 2 and tokens.is_contiguous()
    N, S = tokens.shape
    E = weight.size(1)
    assert weight.dtype == torch.float32 and weight.is_contiguous()
    assert pe.dtype == torch.float32 and pe.is_contiguous() and pe.size(-1) == E and pe.numel() // E >= S
    out = torch.empty((S, N, E), dtype=torch.bfloat16, device=tokens.device)
    with _lib.on_device(tokens.device):
        rc = _lib.load().pb2_embed_fwd(c_p(tokens.data_ptr()), c_p(weight.data_ptr()), c_p(pe.data_ptr()),
                                       c_p(out.data_ptr()), c_i(N), c_i(S), c_i(E), c_f(scale), c_f(drop_p),
                                       c_ull(seed), c_ull(rng_stream), _s(tokens))
    _lib.check(rc, "pb2_embed_fwd")
    return out


def embed_bwd(tokens: Tensor, dout: Tensor, dweight: Tensor, scale: float, drop_p: float, seed: int,
              rng_stream: int) -> None:
    N, S = tokens.shape
    E = dweight.size(1)
    assert dout.dtype == torch.bfloat16 and dout.is_contiguous() and dweight.dtype == torch.float32
    with _lib.on_device(tokens.device):
        rc = _lib.load().pb2_embed_bwd(c_p(tokens.data_ptr()), c_p(dout.data_ptr()), c_p(dweight.data_ptr()),
                                       c_i(N), c_i(S), c_i(E), c_f(scale), c_f(drop_p), c_ull(seed),
                                       c_ull(rng_stream), _s(tokens))
    _lib.check(rc, "pb2_embed_bwd")


def cross_entropy(logits: Tensor, targets: Tensor, V: int, *, write_grad: bool,
                  row_scale: Optional[Tensor] = None, want_loss: bool = True) -> Optional[Tensor]:
    """Per-row losses for bf16 ``logits [rows, ld]``; with ``write_grad`` the
    logits are overwritten by ``(softmax - onehot) * row_scale``."""
    assert logits.dtype == torch.bfloat16 and logits.dim() == 2 and logits.stride(1) == 1
    assert targets.dtype == torch.int64 and targets.is_contiguous() and targets.numel() == logits.size(0)
    rows = logits.size(0)
    loss = torch.empty(rows, dtype=torch.float32, device=logits.device) if want_loss else None
    if row_scale is not None:
        assert row_scale.dtype == torch.float32 and row_scale.is_contiguous() and row_scale.numel() == rows
    with _lib.on_device(logits.device):
        rc = _lib.load().pb2_cross_entropy(c_p(logits.data_ptr()), c_p(targets.data_ptr()), c_p(ptr(loss)),
                                           c_p(ptr(row_scale)), c_i(rows), c_i(V), c_ll(logits.stride(0)),
                                           c_i(int(write_grad)), _s(logits))
    _lib.check(rc, "pb2_cross_entropy")
    return loss


def transpose_colsum(src: Tensor, dst: Tensor, colsum: Optional[Tensor] = None, scale: float = 1.0) -> None:
    """``dst[c, r] = src[r, c]``; ``dst`` is ``[C, ld]`` with ``ld >= R`` (zero-filled beyond R)."""
    assert src.dtype == torch.bfloat16 and dst.dtype == torch.bfloat16 and src.dim() == 2 and dst.dim() == 2
    R, C = src.shape
    assert dst.size(0) == C and dst.stride(1) == 1 and src.stride(1) == 1
    with _lib.on_device(src.device):
        rc = _lib.load().pb2_transpose_colsum(c_p(src.data_ptr()), c_ll(src.stride(0)), c_p(dst.data_ptr()),
                                              c_ll(dst.stride(0)), c_i(R), c_i(C), c_p(ptr(colsum)), c_f(scale),
                                              _s(src))
    _lib.check(rc, "pb2_transpose_colsum")


class _TransposeJob(ctypes.Structure):
    _fields_ = [("src", c_p), ("dst", c_p), ("ld_src", c_ll), ("ld_dst", c_ll), ("R", c_i), ("C", c_i),
                ("colsum", c_p), ("scale", c_f)]


class _TransposeJobs(ctypes.Structure):
    _fields_ = [("job", _TransposeJob * 8), ("count", c_i), ("n_peers", c_i), ("peer_delta", c_ll * 7)]


def transpose_multi(jobs: Sequence[Tuple[Tensor, Tensor, int, Optional[Tensor], float]],
                    peer_deltas: Sequence[int] = ()) -> None:
    """Each job is ``(src [R, C], dst_base [C, ld], dst_col_offset, colsum or None, scale)``:
    writes ``dst_base[c, off + r] = src[r, c]`` for all jobs in one launch.  ``peer_deltas`` (bytes): every
    stored element is also stored at the same address plus each delta — the same slot of a symmetric
    stash in the data-parallel replicas' memory."""
    assert 0 < len(jobs) <= 8 and len(peer_deltas) <= 7
    packed = _TransposeJobs()
    packed.count = len(jobs)
    packed.n_peers = len(peer_deltas)
    for q, delta in enumerate(peer_deltas):
        packed.peer_delta[q] = delta
    dev = jobs[0][0].device
    for k, (src, dst, off, colsum, scale) in enumerate(jobs):
        assert src.dtype == torch.bfloat16 and dst.dtype == torch.bfloat16 and src.stride(1) == 1 and dst.stride(1) == 1
        R, C = src.shape
        assert dst.size(0) == C and off + R <= dst.size(1), (src.shape, dst.shape, off)
        j = packed.job[k]
        j.src, j.dst = src.data_ptr(), dst.data_ptr() + off * 2
        j.ld_src, j.ld_dst = src.stride(0), dst.stride(0)
        j.R, j.C = R, C
        j.colsum = ptr(colsum)
        j.scale = scale
    with _lib.on_device(dev):
        rc = _lib.load().pb2_transpose_multi(ctypes.byref(packed), _lib.stream_ptr(dev))
    _lib.check(rc, "pb2_transpose_multi")


def sumsq(x: Tensor, out: Tensor) -> None:
    assert x.dtype == torch.float32 and x.is_contiguous() and out.dtype == torch.float32
    with _lib.on_device(x.device):
        rc = _lib.load().pb2_sumsq(c_p(x.data_ptr()), ctypes.c_size_t(x.numel()), c_p(out.data_ptr()), _s(x))
    _lib.check(rc, "pb2_sumsq")


def attention_fwd(qkv: Tensor, NB: int, S: int, H: int, *, causal: bool, drop_p: float, seed: int,
                  rng_stream: int) -> Tensor:
    T, E3 = qkv.shape
    E = E3 // 3
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and T == NB * S and E % H == 0
    out = torch.empty((T, E), dtype=torch.bfloat16, device=qkv.device)
    with _lib.on_device(qkv.device):
        rc = _lib.load().pb2_attention_fwd(c_p(qkv.data_ptr()), c_p(out.data_ptr()), c_i(NB), c_i(S), c_i(H),
                                           c_i(E // H), c_i(int(causal)), c_f(drop_p), c_ull(seed),
                                           c_ull(rng_stream), _s(qkv))
    _lib.check(rc, "pb2_attention_fwd")
    return out


def attention_bwd(qkv: Tensor, dout: Tensor, NB: int, S: int, H: int, *, causal: bool, drop_p: float, seed: int,
                  rng_stream: int) -> Tensor:
    T, E3 = qkv.shape
    E = E3 // 3
    assert dout.dtype == torch.bfloat16 and dout.is_contiguous() and dout.shape == (T, E)
    dqkv = torch.empty_like(qkv)
    with _lib.on_device(qkv.device):
        rc = _lib.load().pb2_attention_bwd(c_p(qkv.data_ptr()), c_p(dout.data_ptr()), c_p(dqkv.data_ptr()), c_i(NB),
                                           c_i(S), c_i(H), c_i(E // H), c_i(int(causal)), c_f(drop_p), c_ull(seed),
                                           c_ull(rng_stream), _s(qkv))
    _lib.check(rc, "pb2_attention_bwd")
    return dqkv


def flash_attention_fwd(qkv: Tensor, NB: int, S: int, H: int, *, causal: bool, drop_p: float, seed: int,
                        rng_stream: int, want_lse: bool = True) -> Tuple[Tensor, Optional[Tensor]]:
    """Blocked (flash) attention on the tcgen05 kernels: any ``S``, head_dim in {64, 128, 256}.  Returns
    ``(out [T, E] bf16, lse [NB * H, S] fp32 or None)``."""
    T, E3 = qkv.shape
    E = E3 // 3
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and T == NB * S and E % H == 0
    out = torch.empty((T, E), dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty((NB * H, S), dtype=torch.float32, device=qkv.device) if want_lse else None
    with _lib.on_device(qkv.device):
        rc = _lib.load().pb2_flash_attention_fwd(c_p(qkv.data_ptr()), c_p(out.data_ptr()), c_p(ptr(lse)), c_i(NB), c_i(S),
                                                 c_i(H), c_i(E // H), c_i(int(causal)), c_f(drop_p), c_ull(seed),
                                                 c_ull(rng_stream), _s(qkv))
    _lib.check(rc, "pb2_flash_attention_fwd")
    return out, lse


def flash_attention_bwd(qkv: Tensor, out: Tensor, dout: Tensor, lse: Tensor, NB: int, S: int, H: int, *, causal: bool,
                        drop_p: float, seed: int, rng_stream: int) -> Tensor:
    """Gradient of :func:`flash_attention_fwd` with respect to the packed QKV (head_dim 64 or 128)."""
    T, E3 = qkv.shape
    E = E3 // 3
    assert dout.dtype == torch.bfloat16 and dout.is_contiguous() and dout.shape == (T, E) and out.shape == (T, E)
    assert lse.dtype == torch.float32 and lse.is_contiguous() and lse.numel() == NB * H * S
    lib = _lib.load()
    nbytes = int(lib.pb2_flash_attention_bwd_scratch_bytes(c_i(NB), c_i(S), c_i(H), c_i(E // H)))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=qkv.device)
    dqkv = torch.empty_like(qkv)
    with _lib.on_device(qkv.device):
        rc = lib.pb2_flash_attention_bwd(c_p(qkv.data_ptr()), c_p(out.data_ptr()), c_p(dout.data_ptr()), c_p(lse.data_ptr()),
                                         c_p(dqkv.data_ptr()), c_p(scratch.data_ptr()), ctypes.c_size_t(nbytes), c_i(NB),
                                         c_i(S), c_i(H), c_i(E // H), c_i(int(causal)), c_f(drop_p), c_ull(seed),
                                         c_ull(rng_stream), _s(qkv))
    _lib.check(rc, "pb2_flash_attention_bwd")
    return dqkv
