"""The tutorial Transformer's blocks on the sm_100a kernels of :mod:`pipe_b200.ops`.

Drop-in bodies for :class:`~pipe_b200.models.transformer.Encoder`,
``nn.TransformerEncoderLayer`` and :class:`~pipe_b200.models.transformer.Decoder`
with the *same parameter names* (``self_attn.in_proj_weight``, ``linear1.weight``,
``norm1.weight``, ``encoder.weight``, ``decoder.weight`` ...), so a state dict
moves freely between the eager and the fused model.

Precision policy: parameters are fp32 masters (what the optimizer and the state
dict see); every GEMM reads a bf16 *shadow* of its weight (and the dgrad GEMM a
transposed bf16 shadow), accumulates in fp32 in TMEM, and activations travel
between layers in bf16.  :class:`~pipe_b200.ops.optim.FusedClipAdam` rewrites
master and both shadows in one pass; any other in-place update of a master is
noticed through the tensor version counter and the shadows are re-derived.

Per encoder layer the forward is 7 launches (QKV GEMM, attention, out-proj GEMM
with bias+dropout epilogue, add+LayerNorm, FFN1 GEMM with bias+ReLU+dropout, FFN2
GEMM with bias+dropout, add+LayerNorm) and the backward 8 (two LayerNorm
backwards that also undo the dropout, four dgrad GEMMs with fused
ReLU/dropout-mask and residual-add epilogues, attention backward, one
multi-transpose that stacks all wgrad operands); weight gradients are one GEMM per
Linear per *mini-batch* (:mod:`pipe_b200.ops.wgrad`).

Parity: reference ``main.py:24-73,115-120,139-157``; SURVEY.md K1-K9.
"""
from __future__ import annotations

import math
import os
import threading
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor, nn

import ctypes

from ..ops import _lib
from ..ops import functional as F2
from ..ops.gemm import gemm_nt, workspace
from ..ops.wgrad import LinearGrad, ensure_grad

__all__ = ["FusedEncoder", "FusedEncoderLayer", "FusedDecoder", "Shadow"]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


# The dgrad GEMMs read the forward shadow [out, in] as an MN-major tcgen05 operand, so no transposed copy is kept
# (2 of the optimizer's 32 bytes per parameter and step, and 2 bytes per parameter of memory).  PB2_DGRAD_SHADOW=1
# brings the transposed bf16 shadow back (the round-1 layout, both operands K-major).
_TRANSPOSED_SHADOW = os.environ.get("PB2_DGRAD_SHADOW", "0") == "1"


class Shadow:
    """bf16 copy ``w16 [out, in]`` of an fp32 master weight (and, with ``PB2_DGRAD_SHADOW=1``, a transposed copy
    ``w16t [in, ld_t]``), kept in sync lazily."""

    def __init__(self, param: nn.Parameter) -> None:
        self.param = param
        self.w16: Optional[Tensor] = None
        self.w16t: Optional[Tensor] = None
        self.version = -1

    def allocate(self) -> None:
        p = self.param
        if self.w16 is None or self.w16.device != p.device:
            rows, cols = p.shape
            self.w16 = torch.empty((rows, cols), dtype=torch.bfloat16, device=p.device)
            self.w16t = (torch.zeros((cols, _round_up(rows, 8)), dtype=torch.bfloat16, device=p.device)
                         if _TRANSPOSED_SHADOW else None)
            self.version = -1

    def ensure(self) -> Tuple[Tensor, Optional[Tensor]]:
        p = self.param
        self.allocate()
        if p._version != self.version:
            with torch.no_grad():
                self.w16.copy_(p)
                if self.w16t is not None:
                    self.w16t[:, : p.size(0)].copy_(p.t())
            self.version = p._version
        return self.w16, self.w16t

    def mark_synced(self) -> None:
        self.version = self.param._version


class _Linear(nn.Module):
    """Parameter holder with nn.Linear's names and init (no forward of its own)."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        ref = nn.Linear(in_features, out_features)
        self.weight = nn.Parameter(ref.weight.detach().clone())
        self.bias = nn.Parameter(ref.bias.detach().clone())


class _Attention(nn.Module):
    """Parameter holder with nn.MultiheadAttention's names and init."""

    def __init__(self, embed_dim: int):
        super().__init__()
        self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * embed_dim))
        self.out_proj = _Linear(embed_dim, embed_dim)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.zeros_(self.out_proj.bias)


class _Norm(nn.Module):
    def __init__(self, dim: int, eps: float = 1e-5):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.bias = nn.Parameter(torch.zeros(dim))
        self.eps = eps


class _FusedBase(nn.Module):
    """Common bookkeeping: shadows and deferred-wgrad objects per weight."""

    def _register_linear(self, key: str, weight: nn.Parameter, bias: Optional[nn.Parameter]) -> None:
        if "_pb2_lin" not in self.__dict__:
            self.__dict__["_pb2_lin"] = {}
        self.__dict__["_pb2_lin"][key] = (Shadow(weight), LinearGrad(weight, bias))

    def pb2_shadows(self) -> List[Shadow]:
        return [s for s, _ in self.__dict__.get("_pb2_lin", {}).values()]

    def pb2_linear_grads(self) -> List[LinearGrad]:
        return [g for _, g in self.__dict__.get("_pb2_lin", {}).values()]

    def _sh(self, key: str) -> Tuple[Tensor, Tensor]:
        return self.__dict__["_pb2_lin"][key][0].ensure()

    def _lg(self, key: str) -> LinearGrad:
        return self.__dict__["_pb2_lin"][key][1]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        # masters may have been replaced / moved: re-bind shadows and wgrad stashes, drop caches
        self.__dict__.pop("_pb2_params", None)
        self.__dict__.pop("_pb2_desc", None)
        self.__dict__["_pb2_epoch"] = self.__dict__.get("_pb2_epoch", 0) + 1
        self._rebind()
        return out

    def _rebind(self) -> None:
        pass


# ======================================================================================
# encoder layer
# ======================================================================================
# The launch that stacks x^T / dy^T for the deferred weight gradients runs on a stream of its own: nothing in
# backward reads its results, so on the main stream it was ~6.4 us plus a launch dependency between every layer's
# QKV dgrad and the next layer's LayerNorm backward.  PB2_SIDE_TRANSPOSE=0 puts it back in line.
_SIDE_TRANSPOSE = os.environ.get("PB2_SIDE_TRANSPOSE", "1") != "0"
_side_lock = threading.Lock()
_side_streams: Dict[int, "torch.cuda.Stream"] = {}


def _side_stream(device: torch.device) -> "torch.cuda.Stream":
    with _side_lock:
        st = _side_streams.get(device.index)
        if st is None:
            st = _side_streams[device.index] = torch.cuda.Stream(device)
        return st

def _uses_flash(S: int, D: int) -> bool:
    """The layer executor routes this (sequence length, head_dim) to the flash-attention kernels: extra buffers needed."""
    return bool(_lib.load().pb2_attention_uses_flash(ctypes.c_int(S), ctypes.c_int(D)))


def _flash_scratch(NB: int, S: int, H: int, D: int, device: torch.device) -> Tensor:
    n = int(_lib.load().pb2_flash_attention_bwd_scratch_bytes(ctypes.c_int(NB), ctypes.c_int(S), ctypes.c_int(H), ctypes.c_int(D)))
    return torch.empty(n, dtype=torch.uint8, device=device)


class _EncoderLayerFn(torch.autograd.Function):
    """Forward and backward of the whole layer, each one call into the C++ layer executor
    (``csrc/layer.cu``): 7 and 8 kernel launches respectively, no Python between them."""

    @staticmethod
    def forward(ctx, x: Tensor, layer: "FusedEncoderLayer", *params: Tensor) -> Tensor:  # type: ignore[override]
        S, NB, E = x.shape
        T = S * NB
        FF = layer.linear1.weight.size(0)
        x2 = x.contiguous().view(T, E)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        p = layer.dropout if layer.training else 0.0
        seed = F2.draw_seed() if p > 0 else 0
        dev = x2.device

        # one allocation for every saved activation: qkv | a | o | x1 | hmid | f
        widths = (3 * E, E, E, E, FF, E)
        act = torch.empty(T * sum(widths), dtype=torch.bfloat16, device=dev)
        flash = _uses_flash(S, E // layer.nhead)
        stats = torch.empty(4 * T + (layer.nhead * T if flash else 0), dtype=torch.float32, device=dev)
        y = torch.empty((S, NB, E), dtype=torch.bfloat16, device=dev)

        d = layer._desc(dev)
        d.T, d.NB, d.S = T, NB, S
        d.drop_p, d.seed = p, seed
        d.x, d.y = x2.data_ptr(), y.data_ptr()
        base, at = act.data_ptr(), 0
        for name, w in zip(("qkv", "a", "o", "x1", "hmid", "f"), widths):
            setattr(d, name, base + at * 2)
            at += T * w
        sb = stats.data_ptr()
        d.mean1, d.rstd1, d.mean2, d.rstd2 = sb, sb + 4 * T, sb + 8 * T, sb + 12 * T
        if flash:
            d.attn_lse = sb + 16 * T
        stream = _lib.raw_stream(dev)
        ws = workspace(dev, stream)
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
        with _lib.on_device(dev):
            rc = _lib.load().pb2_encoder_layer_fwd(ctypes.byref(d), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_encoder_layer_fwd")

        ctx.layer = layer
        ctx.meta = (S, NB, E, FF, p, seed)
        ctx.save_for_backward(x2, act, stats)
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        layer: FusedEncoderLayer = ctx.layer
        S, NB, E, FF, p, seed = ctx.meta
        T = S * NB
        x2, act, stats = ctx.saved_tensors
        dev = x2.device
        dy2 = dy.contiguous().view(T, E)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        at_, l1, l2, n1, n2 = layer.self_attn, layer.linear1, layer.linear2, layer.norm1, layer.norm2

        stream = _lib.raw_stream(dev)
        # scratch: dr2 | df | dz1 | dx1 | dr1 | do | da | dqkv
        widths = (E, E, FF, E, E, E, E, 3 * E)
        tmp = torch.empty(T * sum(widths), dtype=torch.bfloat16, device=dev)
        dx = torch.empty((S, NB, E), dtype=torch.bfloat16, device=dev)

        d = layer._desc(dev)
        d.T, d.NB, d.S = T, NB, S
        d.drop_p, d.seed = p, seed
        d.x, d.dy, d.dx = x2.data_ptr(), dy2.data_ptr(), dx.data_ptr()
        base, at = act.data_ptr(), 0
        for name, w in zip(("qkv", "a", "o", "x1", "hmid", "f"), (3 * E, E, E, E, FF, E)):
            setattr(d, name, base + at * 2)
            at += T * w
        sb = stats.data_ptr()
        d.mean1, d.rstd1, d.mean2, d.rstd2 = sb, sb + 4 * T, sb + 8 * T, sb + 12 * T
        scratch = None
        if _uses_flash(S, E // layer.nhead):
            scratch = _flash_scratch(NB, S, layer.nhead, E // layer.nhead, dev)
            d.attn_lse, d.attn_scratch, d.attn_scratch_bytes = sb + 16 * T, scratch.data_ptr(), scratch.numel()
        base, at = tmp.data_ptr(), 0
        for name, w in zip(("dr2", "df", "dz1", "dx1", "dr1", "do_", "da", "dqkv"), widths):
            setattr(d, name, base + at * 2)
            at += T * w
        d.dg1, d.dbe1 = ensure_grad(n1.weight).data_ptr(), ensure_grad(n1.bias).data_ptr()
        d.dg2, d.dbe2 = ensure_grad(n2.weight).data_ptr(), ensure_grad(n2.bias).data_ptr()
        for key, bias in (("qkv", at_.in_proj_bias), ("o", at_.out_proj.bias), ("1", l1.bias), ("2", l2.bias)):
            lg = layer._lg({"qkv": "qkv", "o": "o", "1": "l1", "2": "l2"}[key])
            off = lg.reserve(T, dev, stream)
            setattr(d, "xt_" + key, lg.xt.data_ptr())
            setattr(d, "dyt_" + key, lg.dyt.data_ptr())
            setattr(d, "ld_stash_" + key, lg.xt.stride(0))
            setattr(d, "stash_off_" + key, off)
            setattr(d, "db_" + key, bias.grad.data_ptr())
        deltas = layer._lg("qkv").peer_deltas       # data-parallel replicas' stashes (hybrid), same for all four
        d.n_peers = len(deltas)
        for q, delta in enumerate(deltas):
            d.peer_delta[q] = delta
        ws = workspace(dev, stream)
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
        side = _side_stream(dev) if _SIDE_TRANSPOSE else None
        d.side_stream = side.cuda_stream if side is not None else None
        with _lib.on_device(dev):
            rc = _lib.load().pb2_encoder_layer_bwd(ctypes.byref(d), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_encoder_layer_bwd")
        if side is not None:
            # the side stream reads the layer input, the saved activations and this call's scratch: keep the
            # allocator from recycling them before it has; consumers of the stacks wait in wgrad.flush_pending
            x2.record_stream(side)
            act.record_stream(side)
            tmp.record_stream(side)
            from ..ops import wgrad as _wgrad
            _wgrad.note_side_stream(dev, stream, side)
        return (dx, None) + (None,) * (len(ctx.needs_input_grad) - 2)


class FusedEncoderLayer(_FusedBase):
    """Post-norm Transformer encoder layer (ReLU FFN, sequence-first input
    ``(S, N, E)``) — ``nn.TransformerEncoderLayer(d_model, nhead, dim_feedforward,
    dropout)`` on hand-written kernels.  Extra positional inputs are passed
    through untouched (used to carry targets to a fused loss on the last stage)."""

    def __init__(self, d_model: int, nhead: int, dim_feedforward: int = 2048, dropout: float = 0.1,
                 causal: bool = False, layer_norm_eps: float = 1e-5):
        super().__init__()
        if d_model % nhead:
            raise ValueError("d_model must be divisible by nhead")
        self.d_model, self.nhead, self.dropout, self.causal = d_model, nhead, float(dropout), bool(causal)
        self.self_attn = _Attention(d_model)
        self.linear1 = _Linear(d_model, dim_feedforward)
        self.linear2 = _Linear(dim_feedforward, d_model)
        self.norm1 = _Norm(d_model, layer_norm_eps)
        self.norm2 = _Norm(d_model, layer_norm_eps)
        self._rebind()

    def _rebind(self) -> None:
        if not hasattr(self, "norm2"):
            return
        self._register_linear("qkv", self.self_attn.in_proj_weight, self.self_attn.in_proj_bias)
        self._register_linear("o", self.self_attn.out_proj.weight, self.self_attn.out_proj.bias)
        self._register_linear("l1", self.linear1.weight, self.linear1.bias)
        self._register_linear("l2", self.linear2.weight, self.linear2.bias)

    def _desc(self, device: torch.device) -> "_lib.EncoderLayerDesc":
        """A fresh launch descriptor with everything that does not change between calls
        (shapes, weight shadows, biases, LayerNorm parameters) filled in."""
        sh = {k: self._sh(k) for k in ("qkv", "o", "l1", "l2")}
        key = tuple(t.data_ptr() for pair in sh.values() for t in pair if t is not None)
        cached = self.__dict__.get("_pb2_desc")
        if cached is None or cached[0] != key:
            d = _lib.EncoderLayerDesc()
            d.E, d.heads, d.FF = self.d_model, self.nhead, self.linear1.weight.size(0)
            d.causal, d.eps = int(self.causal), self.norm1.eps
            for name, k in (("qkv", "qkv"), ("o", "o"), ("1", "l1"), ("2", "l2")):
                w16, w16t = sh[k]
                setattr(d, "w_" + name, w16.data_ptr())
                if w16t is not None:
                    setattr(d, "wt_" + name, w16t.data_ptr())
                    setattr(d, "ldt_" + name, w16t.stride(0))
            at = self.self_attn
            d.b_qkv, d.b_o = at.in_proj_bias.data_ptr(), at.out_proj.bias.data_ptr()
            d.b_1, d.b_2 = self.linear1.bias.data_ptr(), self.linear2.bias.data_ptr()
            d.g1, d.be1 = self.norm1.weight.data_ptr(), self.norm1.bias.data_ptr()
            d.g2, d.be2 = self.norm2.weight.data_ptr(), self.norm2.bias.data_ptr()
            cached = (key, bytes(d))
            self.__dict__["_pb2_desc"] = cached
        return _lib.EncoderLayerDesc.from_buffer_copy(cached[1])

    def reference_forward(self, x: Tensor) -> Tensor:
        """The same layer with the same parameters in plain PyTorch (fp32): the kernels' oracle and the path
        taken on non-CUDA devices.  Dropout draws from torch's generator instead of the kernels' counter RNG,
        so outputs agree with the fused path exactly only at ``dropout = 0`` / in eval mode."""
        S, N, E = x.shape
        H, D = self.nhead, E // self.nhead
        p = self.dropout if self.training else 0.0
        xf = x.float()
        qkv = F.linear(xf, self.self_attn.in_proj_weight, self.self_attn.in_proj_bias)
        q, k, v = (t.reshape(S, N, H, D).permute(1, 2, 0, 3) for t in qkv.chunk(3, dim=-1))
        a = F.scaled_dot_product_attention(q, k, v, dropout_p=p, is_causal=self.causal)
        a = a.permute(2, 0, 1, 3).reshape(S, N, E)
        o = F.linear(a, self.self_attn.out_proj.weight, self.self_attn.out_proj.bias)
        x1 = F.layer_norm(xf + F.dropout(o, p, True), (E,), self.norm1.weight, self.norm1.bias, self.norm1.eps)
        hidden = F.dropout(F.relu(F.linear(x1, self.linear1.weight, self.linear1.bias)), p, True)
        f = F.linear(hidden, self.linear2.weight, self.linear2.bias)
        y = F.layer_norm(x1 + F.dropout(f, p, True), (E,), self.norm2.weight, self.norm2.bias, self.norm2.eps)
        return y.to(x.dtype)

    @staticmethod
    def pb2_make_chain(layers: "List[FusedEncoderLayer]") -> "EncoderStack":
        """Called by ``PipeSequential`` for a run of consecutive layers of this class."""
        return EncoderStack(layers)

    def forward(self, x: Tensor, *extra: Any):
        if not x.is_cuda:
            y = self.reference_forward(x)
            return (y, *extra) if extra else y
        params = self.__dict__.get("_pb2_params")
        if params is None:
            params = self.__dict__["_pb2_params"] = tuple(self.parameters())
        y = _EncoderLayerFn.apply(x, self, *params)
        return (y, *extra) if extra else y


# ======================================================================================
# a run of encoder layers: one native call per (stage, micro-batch) cell and pass
# ======================================================================================
_STACK = os.environ.get("PB2_LAYER_STACK", "1") != "0"


class EncoderStack:
    """Executes consecutive :class:`FusedEncoderLayer` modules of one stage through
    ``pb2_encoder_stack_fwd / _bwd`` (``csrc/layer.cu``): the per-layer constants live in a cached
    array of launch descriptors, each pass allocates one slab per kind of buffer for the whole run,
    and the Python cost of a cell no longer grows with the number of layers in the stage — the
    single-process ``Pipe`` shares one interpreter between all stages, and a one-rank-per-stage
    pipeline stretches every tick by whatever its slowest host needs (reference: the worker thread
    running ``partition(*inputs)``, ``pipeline.py:225-230``).

    Not a module: the layers stay where they are (state-dict keys unchanged); this object only
    caches descriptors.  ``PipeSequential`` builds one per run of chainable layers."""

    def __init__(self, layers: "List[FusedEncoderLayer]") -> None:
        from ..ops.wgrad import StackStash
        self.layers = list(layers)
        self.n = len(self.layers)
        first = self.layers[0]
        if any(l.d_model != first.d_model or l.nhead != first.nhead for l in self.layers):
            raise ValueError("layers of a stack must share d_model and nhead")
        self.E = first.d_model
        self.anchor = first.self_attn.in_proj_weight
        self.stash = StackStash([l._lg(k) for l in self.layers for k in ("qkv", "o", "l1", "l2")])
        self._epoch_seen = self._layer_epoch()
        self._static: Optional[Dict[str, Any]] = None
        self._sizes: Dict[int, Tuple[int, int]] = {}

    # -- eligibility ---------------------------------------------------------------
    def usable(self, x: Tensor) -> bool:
        if not (_STACK and x.is_cuda and self.n <= 64):
            return False
        t0, p0 = self.layers[0].training, self.layers[0].dropout
        return all(l.training == t0 and l.dropout == p0 for l in self.layers)

    # -- cached per-layer constants ---------------------------------------------------
    def _layer_epoch(self) -> int:
        return sum(l.__dict__.get("_pb2_epoch", 0) for l in self.layers)

    def _grad_params(self, layer: "FusedEncoderLayer"):
        at = layer.self_attn
        return (layer.norm1.weight, layer.norm1.bias, layer.norm2.weight, layer.norm2.bias,
                at.in_proj_bias, at.out_proj.bias, layer.linear1.bias, layer.linear2.bias)

    def refresh(self) -> None:
        """Layers that were moved / cast re-create their shadows and wgrad stashes: follow them."""
        from ..ops.wgrad import StackStash
        e = self._layer_epoch()
        if e != self._epoch_seen:
            self._epoch_seen = e
            self.stash = StackStash([l._lg(k) for l in self.layers for k in ("qkv", "o", "l1", "l2")])
            self._static = None
        for sh in self._shadows():
            if sh.param._version != sh.version:
                sh.ensure()

    def _shadows(self):
        got = self.__dict__.get("_shadow_list")
        if got is None or got[0] != self._epoch_seen:
            got = (self._epoch_seen, [sh for l in self.layers for sh in l.pb2_shadows()])
            self.__dict__["_shadow_list"] = got
        return got[1]

    def static(self, device: torch.device, with_grads: bool):
        """The descriptor array, rebuilt when a layer was moved, a stash was re-allocated or a
        gradient buffer was replaced (``zero_grad(set_to_none=True)`` of a stock optimizer)."""
        from ..ops.wgrad import stash_epoch
        c = self._static
        if (c is not None and c["device"] == device.index and (c["with_grads"] or not with_grads)
                and (not c["with_grads"] or (c["stash_epoch"] == stash_epoch()
                                             and all(p.grad is g for p, g in c["pairs"])))):
            return c["arr"]
        arr = (_lib.EncoderLayerDesc * self.n)()
        pairs = []
        for k, layer in enumerate(self.layers):
            d = layer._desc(device)
            if with_grads:
                n1w, n1b, n2w, n2b, bq, bo, b1, b2 = self._grad_params(layer)
                d.dg1, d.dbe1 = ensure_grad(n1w).data_ptr(), ensure_grad(n1b).data_ptr()
                d.dg2, d.dbe2 = ensure_grad(n2w).data_ptr(), ensure_grad(n2b).data_ptr()
                for name, bias, lk in (("qkv", bq, "qkv"), ("o", bo, "o"), ("1", b1, "l1"), ("2", b2, "l2")):
                    lg = layer._lg(lk)
                    setattr(d, "db_" + name, ensure_grad(bias).data_ptr())
                    setattr(d, "xt_" + name, lg.xt.data_ptr())
                    setattr(d, "dyt_" + name, lg.dyt.data_ptr())
                    setattr(d, "ld_stash_" + name, lg.xt.stride(0))
                deltas = layer._lg("qkv").peer_deltas
                d.n_peers = len(deltas)
                for q, delta in enumerate(deltas):
                    d.peer_delta[q] = delta
                pairs += [(p, p.grad) for p in (n1w, n1b, n2w, n2b, bq, bo, b1, b2)]
            arr[k] = d
        self._static = {"device": device.index, "with_grads": with_grads, "stash_epoch": stash_epoch(),
                        "arr": arr, "pairs": pairs}
        self._sizes.clear()
        return arr

    def sizes(self, arr, T: int) -> Tuple[int, int]:
        got = self._sizes.get(T)
        if got is None:
            lib = _lib.load()
            got = (int(lib.pb2_encoder_stack_act_elems(arr, ctypes.c_int(self.n), ctypes.c_int(T))),
                   int(lib.pb2_encoder_stack_tmp_elems(arr, ctypes.c_int(self.n), ctypes.c_int(T))))
            self._sizes[T] = got
        return got

    def __call__(self, x: Tensor) -> Tensor:
        return _StackFn.apply(x, self, self.anchor)


class _StackFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor, stack: EncoderStack, anchor: Tensor) -> Tensor:  # type: ignore[override]
        S, NB, E = x.shape
        T = S * NB
        n = stack.n
        x2 = x.contiguous().view(T, E)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        first = stack.layers[0]
        p = first.dropout if first.training else 0.0
        seeds = (ctypes.c_ulonglong * n)(*[F2.draw_seed() for _ in range(n)]) if p > 0 else None
        dev = x2.device
        stack.refresh()
        arr = stack.static(dev, with_grads=False)
        act_elems, _ = stack.sizes(arr, T)
        act = torch.empty(act_elems, dtype=torch.bfloat16, device=dev)
        n_stats = int(_lib.load().pb2_encoder_stack_stats_floats(arr, ctypes.c_int(n), ctypes.c_int(S), ctypes.c_int(T)))
        stats = torch.empty(n_stats, dtype=torch.float32, device=dev)
        ys = torch.empty((n, S, NB, E), dtype=torch.bfloat16, device=dev)
        stream = _lib.raw_stream(dev)
        ws = workspace(dev, stream)
        c = _lib.EncoderStackDesc()
        c.n_layers, c.layers = n, ctypes.addressof(arr)
        c.T, c.NB, c.S, c.drop_p = T, NB, S, p
        c.seeds = ctypes.addressof(seeds) if seeds is not None else None
        c.x, c.y, c.act, c.stats = x2.data_ptr(), ys.data_ptr(), act.data_ptr(), stats.data_ptr()
        c.workspace, c.workspace_bytes = ws.data_ptr(), ws.numel()
        with _lib.on_device(dev):
            rc = _lib.load().pb2_encoder_stack_fwd(ctypes.byref(c), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_encoder_stack_fwd")
        ctx.stack = stack
        ctx.meta = (S, NB, E, p, seeds)
        ctx.save_for_backward(x2, ys, act, stats)
        return ys[n - 1]

    @staticmethod
    def backward(ctx, dy: Tensor):  # type: ignore[override]
        stack: EncoderStack = ctx.stack
        S, NB, E, p, seeds = ctx.meta
        T = S * NB
        n = stack.n
        x2, ys, act, stats = ctx.saved_tensors
        dev = x2.device
        dy2 = dy.contiguous().view(T, E)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        stream = _lib.raw_stream(dev)
        stack.refresh()
        off = stack.stash.reserve(T, dev, stream)          # may (re)allocate the stacks: before static()
        arr = stack.static(dev, with_grads=True)
        _, tmp_elems = stack.sizes(arr, T)
        tmp = torch.empty(tmp_elems, dtype=torch.bfloat16, device=dev)
        dx = torch.empty((S, NB, E), dtype=torch.bfloat16, device=dev)
        ws = workspace(dev, stream)
        side = _side_stream(dev) if _SIDE_TRANSPOSE else None
        c = _lib.EncoderStackDesc()
        c.n_layers, c.layers = n, ctypes.addressof(arr)
        c.T, c.NB, c.S, c.drop_p = T, NB, S, p
        c.seeds = ctypes.addressof(seeds) if seeds is not None else None
        c.x, c.y, c.act, c.stats = x2.data_ptr(), ys.data_ptr(), act.data_ptr(), stats.data_ptr()
        c.dy, c.dx, c.tmp, c.stash_off = dy2.data_ptr(), dx.data_ptr(), tmp.data_ptr(), off
        first = stack.layers[0]
        scratch = None
        if _uses_flash(S, E // first.nhead):
            scratch = _flash_scratch(NB, S, first.nhead, E // first.nhead, dev)
            c.attn_scratch, c.attn_scratch_bytes = scratch.data_ptr(), scratch.numel()
        c.workspace, c.workspace_bytes = ws.data_ptr(), ws.numel()
        c.side_stream = side.cuda_stream if side is not None else None
        with _lib.on_device(dev):
            rc = _lib.load().pb2_encoder_stack_bwd(ctypes.byref(c), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_encoder_stack_bwd")
        if side is not None:
            for t in (x2, ys, act, tmp):
                t.record_stream(side)
            from ..ops import wgrad as _wgrad
            _wgrad.note_side_stream(dev, stream, side)
        return dx, None, None


# ======================================================================================
# embedding + positions + dropout
# ======================================================================================
class _EmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens: Tensor, module: "FusedEncoder", weight: Tensor) -> Tensor:  # type: ignore[override]
        p = module.dropout if module.training else 0.0
        seed = F2.draw_seed() if p > 0 else 0
        tokens = tokens.contiguous()
        out = F2.embed_fwd(tokens, weight, module.pos_encoder.pe, math.sqrt(module.ninp), p, seed, 0)
        ctx.module = module
        ctx.meta = (p, seed)
        ctx.save_for_backward(tokens)
        return out

    @staticmethod
    def backward(ctx, dout: Tensor):  # type: ignore[override]
        (tokens,) = ctx.saved_tensors
        p, seed = ctx.meta
        module: FusedEncoder = ctx.module
        d = dout.contiguous()
        if d.dtype != torch.bfloat16:
            d = d.to(torch.bfloat16)
        stash = module.__dict__.get("_pb2_embed_stash")
        if stash is not None:
            # data-parallel hybrid: the rows go to every replica's stash; the table gradient is formed at flush time
            stash.pack(tokens, d, math.sqrt(module.ninp), p, seed, 0)
        else:
            F2.embed_bwd(tokens, d, ensure_grad(module.encoder.weight), math.sqrt(module.ninp), p, seed, 0)
        return None, None, None


class _PE(nn.Module):
    def __init__(self, d_model: int, max_len: int = 5000):
        super().__init__()
        position = torch.arange(max_len, dtype=torch.float32).unsqueeze(1)
        freq = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * (-math.log(10000.0) / d_model))
        pe = torch.zeros(max_len, 1, d_model)
        pe[:, 0, 0::2] = torch.sin(position * freq)
        pe[:, 0, 1::2] = torch.cos(position * freq)
        self.register_buffer("pe", pe)


class FusedEncoder(nn.Module):
    """``(N, S)`` int64 tokens -> ``(S, N, E)`` bf16: embedding * sqrt(E) + sinusoidal
    positions + dropout in one kernel (reference ``main.py:24-40,57-73``).  A second
    input (e.g. ``(N, S)`` targets) is passed through."""

    def __init__(self, ntoken: int, ninp: int, dropout: float = 0.5):
        super().__init__()
        self.pos_encoder = _PE(ninp)
        self.encoder = nn.Embedding(ntoken, ninp)
        self.ninp = ninp
        self.dropout = float(dropout)
        nn.init.uniform_(self.encoder.weight, -0.1, 0.1)

    def pb2_embed_weight(self) -> nn.Parameter:
        """The table whose gradient a data-parallel hybrid may exchange row-wise (parallel/replica_stash.py)."""
        return self.encoder.weight

    def forward(self, src: Tensor, *extra: Any):
        if src.size(1) > self.pos_encoder.pe.size(0):
            raise ValueError("sequence longer than the positional table")
        if not src.is_cuda:
            out = self.reference_forward(src)
            return (out, *extra) if extra else out
        out = _EmbedFn.apply(src, self, self.encoder.weight)
        return (out, *extra) if extra else out

    def reference_forward(self, src: Tensor) -> Tensor:
        """Plain PyTorch of the same module (fp32 out): oracle and non-CUDA path."""
        S = src.size(1)
        emb = F.embedding(src, self.encoder.weight).transpose(0, 1) * math.sqrt(self.ninp)
        return F.dropout(emb + self.pos_encoder.pe[:S], self.dropout, self.training)


# ======================================================================================
# vocabulary projection (+ fused cross-entropy)
# ======================================================================================
def _decoder_dgrad(module: "FusedDecoder", dl: Tensor, V: int) -> Tensor:
    """dX[T, E] = dlogits[T, V] W[V, E]: the forward shadow read as an MN-major operand (or the transposed shadow)."""
    w16, w16t = module._sh("dec")
    if w16t is not None:
        return gemm_nt(w16t[:, :V], dl, transpose_out=True)
    return gemm_nt(w16, dl, a_mn=True, transpose_out=True, a_static=True)


class _DecoderFn(torch.autograd.Function):
    """logits = x W^T + b, returned batch-first like the reference's Decoder.

    The GEMM runs on batch-first rows, so its output *is* the ``(N, S, V)`` tensor (row pitch padded to 64
    columns).  Inside a pipeline every micro-batch writes into its slice of ONE mini-batch-wide buffer
    (``microbatch.GatherBuffer``, found through ``microbatch.current_cell()``), which makes the final
    ``gather`` a view instead of a 2 x 59 MB copy per micro-batch at the reference's shape
    (reference ``pipe.py:490``, ``README.md:382``)."""

    @staticmethod
    def forward(ctx, x: Tensor, module: "FusedDecoder", *params: Tensor) -> Tensor:  # type: ignore[override]
        from .. import microbatch
        S, NB, E = x.shape
        T = S * NB
        xb = x.transpose(0, 1).contiguous().view(T, E)              # batch-first rows: n * S + s
        if xb.dtype != torch.bfloat16:
            xb = xb.to(torch.bfloat16)
        V = module.ntoken
        ld = _round_up(V, 64)
        w16, _ = module._sh("dec")
        cell = microbatch.current_cell()
        buf = None
        if cell is not None and cell.index < len(cell.sizes) and cell.sizes[cell.index] == NB:
            key = ("logits", id(module), S, ld)
            gb = cell.shared.get(key)
            if gb is None:
                gb = microbatch.GatherBuffer(sum(cell.sizes), (S, ld), dtype=torch.bfloat16, device=x.device)
                gb.plan(cell.sizes)
                cell.shared[key] = gb
            buf = gb.slot(cell.index)                               # [NB, S, ld] alias of the shared buffer
        if buf is None:
            buf = torch.empty((NB, S, ld), dtype=torch.bfloat16, device=x.device)
        gemm_nt(w16, xb, out=buf.view(T, ld)[:, :V], transpose_out=True, bias=module.decoder.bias, bias_along_m=True)
        ctx.module = module
        ctx.meta = (S, NB, E, V, ld)
        ctx.save_for_backward(xb)
        return buf[:, :, :V]

    @staticmethod
    def backward(ctx, dlogits: Tensor):  # type: ignore[override]
        module: FusedDecoder = ctx.module
        S, NB, E, V, ld = ctx.meta
        T = S * NB
        (xb,) = ctx.saved_tensors
        buf = torch.zeros((T, ld), dtype=torch.bfloat16, device=xb.device)
        buf[:, :V].view(NB, S, V).copy_(dlogits)
        dl = buf[:, :V]
        dx = _decoder_dgrad(module, dl, V)                           # batch-first rows
        lg = module._lg("dec")
        F2.transpose_multi(lg.jobs(xb, dl), lg.peer_deltas)
        return (dx.view(NB, S, E).transpose(0, 1), None) + (None,) * (len(ctx.needs_input_grad) - 2)


class _DecoderLossFn(torch.autograd.Function):
    """Vocabulary projection fused with softmax cross-entropy: returns the summed
    loss of each sequence, ``[N]`` fp32.  The bf16 logits live only in a scratch
    buffer that the backward turns into dlogits in place."""

    @staticmethod
    def forward(ctx, x: Tensor, targets: Tensor, module: "FusedDecoder", *params: Tensor) -> Tensor:  # type: ignore[override]
        S, NB, E = x.shape
        T = S * NB
        x2 = x.contiguous().view(T, E)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        V = module.ntoken
        ld = _round_up(V, 64)
        w16, _ = module._sh("dec")
        buf = torch.empty((T, ld), dtype=torch.bfloat16, device=x.device)
        gemm_nt(w16, x2, out=buf[:, :V], transpose_out=True, bias=module.decoder.bias, bias_along_m=True)
        tgt_rows = targets.t().contiguous().view(-1)            # row = s * NB + n
        loss_rows = F2.cross_entropy(buf, tgt_rows, V, write_grad=False)
        ctx.module = module
        ctx.meta = (S, NB, E, V)
        ctx.save_for_backward(x2, buf, tgt_rows)
        return loss_rows.view(S, NB).sum(0)

    @staticmethod
    def backward(ctx, g: Tensor):  # type: ignore[override]
        module: FusedDecoder = ctx.module
        S, NB, E, V = ctx.meta
        x2, buf, tgt_rows = ctx.saved_tensors
        row_scale = g.float().contiguous().repeat(S)
        F2.cross_entropy(buf, tgt_rows, V, write_grad=True, row_scale=row_scale, want_loss=False)
        dl = buf[:, :V]
        dx = _decoder_dgrad(module, dl, V)
        lg = module._lg("dec")
        F2.transpose_multi(lg.jobs(x2, dl), lg.peer_deltas)
        return (dx.view(S, NB, E), None, None) + (None,) * (len(ctx.needs_input_grad) - 3)


class FusedDecoder(_FusedBase):
    """``(S, N, E)`` -> batch-first logits ``(N, S, V)`` (reference ``main.py:42-55``).
    Called with a second input ``targets (N, S)`` it instead returns each sequence's
    summed cross-entropy ``[N]`` with the loss fused behind the projection (the
    logits are never handed to PyTorch)."""

    def __init__(self, ntoken: int, ninp: int):
        super().__init__()
        self.ntoken = ntoken
        self.decoder = _Linear(ninp, ntoken)
        nn.init.zeros_(self.decoder.bias)
        nn.init.uniform_(self.decoder.weight, -0.1, 0.1)
        self._rebind()

    def _rebind(self) -> None:
        if not hasattr(self, "decoder"):
            return
        self._register_linear("dec", self.decoder.weight, self.decoder.bias)

    def reference_forward(self, inp: Tensor, targets: Optional[Tensor] = None) -> Tensor:
        """Plain PyTorch of the same module (fp32): oracle and non-CUDA path."""
        logits = F.linear(inp.float(), self.decoder.weight, self.decoder.bias).permute(1, 0, 2)      # (N, S, V)
        if targets is None:
            return logits
        N, S, V = logits.shape
        per_token = F.cross_entropy(logits.reshape(N * S, V), targets.reshape(-1), reduction="none")
        return per_token.view(N, S).sum(dim=1)

    def forward(self, inp: Tensor, targets: Optional[Tensor] = None):
        if not inp.is_cuda:
            return self.reference_forward(inp, targets)
        if targets is None:
            return _DecoderFn.apply(inp, self, *self.parameters())
        return _DecoderLossFn.apply(inp, targets, self, *self.parameters())
