"""The weight-gradient "all-reduce" of a Pipe-inside-DDP hybrid, done inside the wgrad GEMM.

Classic data parallelism all-reduces ``dW`` (``in x out`` fp32 per Linear: 64 MiB for a 4096 x 4096 weight).
But this framework computes ``dW`` only once per mini-batch, from the stacked operands of all micro-batches:
``dW = dY^T X`` with ``X^T [in, tokens]`` and ``dY^T [out, tokens]`` in bf16 (``ops/wgrad.py``).  The sum over
``k`` replicas is just a longer reduction dimension::

    dW_mean = 1/k * [dY_0 ; dY_1 ; ...]^T [X_0 ; X_1 ; ...]

so instead of exchanging results, the replicas exchange *operands*: ``(in + out) * tokens`` bf16 per Linear
(8 MiB at 500 tokens, an eighth of the result), and they do it for free — the transpose kernels that stack
``X^T`` / ``dY^T`` during backward (off the critical path, on a side stream) store every element into all
replicas' stashes over NVLink (``csrc/norm_embed.cu`` ``transpose_multi_kernel``, peer-mapped symmetric
memory).  After one device-side barrier each replica holds every replica's operands at the same columns and
runs the same tcgen05 GEMM with K = ``k * cap`` tokens and ``alpha = 1/k``: the reduction over replicas happens
in the fp32 accumulator in TMEM, every replica gets bit-identical averaged gradients, and no ``dW`` ever
crosses a link.  A second barrier after the GEMMs keeps the next step's stores from overwriting operands a
slower peer is still reading.

Only the small gradients (biases, LayerNorm, embedding rows) still go through the all-reduce kernel
(``csrc/allreduce.cu``), launched on a side stream when backward ends so that it runs next to the wgrad GEMMs.

Reference counterpart: none in code — DDP around ``Pipe`` is a docstring remark (``pipe.py:290-293``);
BASELINE.json names this path the product.  SURVEY.md K18 / W11.
"""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist

from ..ops import wgrad
from ..ops.wgrad import LinearGrad
from .symmetric import SymmetricBuffer

__all__ = ["ReplicaStash", "EmbedStash"]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class EmbedStash:
    """The same idea for an embedding table: its dense gradient is ``V x E`` fp32 of which at most ``tokens`` rows are
    non-zero, so the replicas exchange the rows (already scaled and dropout-masked, fp32) and their token ids; after the
    barrier every replica sums them per token in a fixed order (``pb2_embed_scatter``: one writer per token, no
    atomics), which keeps the replicas bit-identical.  472 MB of all-reduce become ``2 x tokens x E x 4`` bytes."""

    def __init__(self, weight: torch.nn.Parameter, group: dist.ProcessGroup, device: torch.device, max_tokens: int) -> None:
        self.weight = weight
        self.replicas = dist.get_world_size(group)
        self.replica = dist.get_rank(group)
        self.cap = _round_up(max(int(max_tokens), 8), 8)
        self.E = weight.size(1)
        slots = self.replicas * self.cap
        self.buf = SymmetricBuffer(slots * self.E + 2 * slots, group, torch.device(device))
        if self.buf.mode == "torch":
            raise RuntimeError("peer-mapped memory is not available for this replica group")
        self.rows = self.buf.data[: slots * self.E].view(slots, self.E)
        self.toks = self.buf.data[slots * self.E: slots * self.E + 2 * slots].view(torch.int64)
        self.toks.fill_(-1)
        me = self.buf.peer_data[self.replica]
        self.deltas = [p - me for q, p in enumerate(self.buf.peer_data) if q != self.replica]
        self.used = 0
        torch.cuda.synchronize(device)
        dist.barrier(group=group)

    def pack(self, tokens: torch.Tensor, dout: torch.Tensor, scale: float, drop_p: float, seed: int, rng_stream: int) -> None:
        """Called by the embedding's backward for one micro-batch (``tokens [N, S]``, ``dout [S, N, E]`` bf16)."""
        import ctypes

        from ..ops import _lib
        N, S = tokens.shape
        T = N * S
        if self.used + T > self.cap:
            raise RuntimeError(f"embedding stash holds {self.cap} tokens per step and replica: raise max_tokens")
        off = self.replica * self.cap + self.used
        self.used += T
        arr = (ctypes.c_longlong * max(len(self.deltas), 1))(*self.deltas)
        with _lib.on_device(dout.device):
            rc = _lib.load().pb2_embed_pack(
                ctypes.c_void_p(tokens.data_ptr()), ctypes.c_void_p(dout.data_ptr()),
                ctypes.c_void_p(self.rows[off].data_ptr()), ctypes.c_void_p(self.toks[off:].data_ptr()),
                ctypes.c_int(N), ctypes.c_int(S), ctypes.c_int(self.E), ctypes.c_float(scale), ctypes.c_float(drop_p),
                ctypes.c_ulonglong(seed), ctypes.c_ulonglong(rng_stream), ctypes.c_int(len(self.deltas)), arr,
                _lib.stream_ptr(dout.device))
        _lib.check(rc, "pb2_embed_pack")

    def scatter(self, stream: int) -> None:
        """After the group barrier: every replica's rows are here; add them up into ``weight.grad`` (mean over replicas)."""
        import ctypes

        from ..ops import _lib
        from ..ops.wgrad import ensure_grad
        grad = ensure_grad(self.weight)
        dev = self.rows.device
        with _lib.on_device(dev):
            # reset_slots: slots a shorter next step leaves unused must read "empty"; done by the library on the same
            # stream, right behind the kernel
            rc = _lib.load().pb2_embed_scatter(ctypes.c_void_p(self.toks.data_ptr()), ctypes.c_void_p(self.rows.data_ptr()),
                                               ctypes.c_void_p(grad.data_ptr()), ctypes.c_int(self.toks.numel()),
                                               ctypes.c_int(self.E), ctypes.c_float(1.0 / self.replicas), ctypes.c_int(1),
                                               ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_embed_scatter")
        self.used = 0


class ReplicaStash:
    """One symmetric slab holding the ``X^T`` / ``dY^T`` stacks of every Linear of a stage, ``replicas * cap``
    columns each, bound to the stage's :class:`~pipe_b200.ops.wgrad.LinearGrad` objects.

    Args:
        lgs: the stage's deferred-wgrad objects (``module.pb2_linear_grads()`` of every fused module).
        group: the replica group (ranks that hold the same stage).
        max_tokens: most tokens a replica stacks per step (rows x sequence length of its mini-batch shard).
    """

    def __init__(self, lgs: Sequence[LinearGrad], group: dist.ProcessGroup, device: torch.device,
                 max_tokens: int, embeddings: Sequence[torch.nn.Module] = ()) -> None:
        """``embeddings``: modules with ``pb2_embed_weight()`` (the fused encoder) whose table gradient is exchanged
        row-wise (:class:`EmbedStash`) instead of all-reduced."""
        self.lgs: List[LinearGrad] = list(lgs)
        self.group = group
        self.device = torch.device(device)
        self.replicas = dist.get_world_size(group)
        self.replica = dist.get_rank(group)
        self.cap = _round_up(max(int(max_tokens), 64), 64)          # whole 64-element K blocks per replica
        cols = self.replicas * self.cap
        elems = sum((lg.weight.size(0) + lg.weight.size(1)) * cols for lg in self.lgs)     # bf16 elements
        self.buf = SymmetricBuffer((elems + 1) // 2, group, self.device)                   # fp32 units
        if self.buf.mode == "torch":
            raise RuntimeError("peer-mapped memory is not available for this replica group")
        slab = self.buf.data.view(torch.bfloat16)
        me = self.buf.peer_data[self.replica]
        deltas = tuple(p - me for q, p in enumerate(self.buf.peer_data) if q != self.replica)
        at = 0
        for lg in self.lgs:
            out_f, in_f = lg.weight.shape
            xt = slab[at:at + in_f * cols].view(in_f, cols)
            at += in_f * cols
            dyt = slab[at:at + out_f * cols].view(out_f, cols)
            at += out_f * cols
            lg.bind_replicated(xt, dyt, self.replica, self.cap, self.replicas, deltas)
        self.bytes_pushed_per_token = sum((lg.weight.size(0) + lg.weight.size(1)) * 2 for lg in self.lgs) \
            * (self.replicas - 1)
        self._mine = {id(lg) for lg in self.lgs}
        self.embed_stashes: List[EmbedStash] = []
        for mod in embeddings:
            es = EmbedStash(mod.pb2_embed_weight(), group, self.device, max_tokens)
            mod.__dict__["_pb2_embed_stash"] = es
            self.embed_stashes.append(es)
        self._embed_modules = list(embeddings)
        wgrad.pre_flush_hooks.append(self._before_gemms)
        wgrad.post_flush_hooks.append(self._after_gemms)

    def _stream_of(self, todo) -> int:
        for lg in todo:
            if id(lg) in self._mine and lg.stream is not None:
                return lg.stream
        from ..ops import _lib
        return _lib.raw_stream(self.device)

    def _before_gemms(self, todo) -> None:
        """All replicas' operand stores have landed before any wgrad GEMM reads them."""
        if any(id(lg) in self._mine for lg in todo):
            stream = self._stream_of(todo)
            self.buf.barrier(phase=0, stream=stream)
            for es in self.embed_stashes:
                es.scatter(stream)

    def _after_gemms(self, todo) -> None:
        """Nobody starts overwriting the stash (next step's backward) while a peer's GEMMs still read it."""
        if any(id(lg) in self._mine for lg in todo):
            self.buf.barrier(phase=1, stream=self._stream_of(todo))

    def close(self) -> None:
        for mod in self._embed_modules:
            mod.__dict__.pop("_pb2_embed_stash", None)
        for hooks, fn in ((wgrad.pre_flush_hooks, self._before_gemms), (wgrad.post_flush_hooks, self._after_gemms)):
            if fn in hooks:
                hooks.remove(fn)
