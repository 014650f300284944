"""Deferred weight gradients: one wgrad GEMM per Linear per mini-batch.

A pipeline runs every Linear ``m`` times per step, once per micro-batch, with
25-128 tokens each.  Computing ``dW += dY_i^T X_i`` per micro-batch would
read-modify-write the whole fp32 ``dW`` ``m`` times (8 bytes/param each) — more
HBM traffic than the forward and dgrad weight reads put together.  Instead each
backward only *stacks* its operands: ``X_i`` and ``dY_i`` are transposed into
column blocks of two persistent K-major buffers ``X^T [in, tokens]`` and
``dY^T [out, tokens]`` (one fused multi-transpose launch per layer, which also
column-sums ``dY`` into the bias gradient).  When the backward pass ends, a single
tcgen05 GEMM with K = all tokens of the mini-batch writes ``dW`` once (and folds
its squared norm into the clip scalar).  Same math up to fp32 reassociation
(SURVEY.md §7.4 item 1).

Flush points: automatically at the end of the autograd pass that stacked
something (engine callback), or — inside ``with deferred():`` — when the block
exits (``DistPipe.backward`` runs one engine pass per micro-batch and flushes
once at its end).
"""
from __future__ import annotations

import ctypes
import threading
from contextlib import contextmanager
from typing import Iterator, List, Optional, Tuple

import torch
from torch import Tensor, nn

from . import _lib
from .gemm import gemm_nt

__all__ = ["LinearGrad", "StackStash", "deferred", "outer_backward", "flush_pending", "ensure_grad", "stash_epoch"]

_lock = threading.RLock()
_pending: "Dict[int, LinearGrad]" = {}
# Side streams that stacked operands since the last flush: {(device index, main stream handle): side stream}.
# The transposes that fill ``xt`` / ``dyt`` (and add up the bias gradients) may run on a stream of their own
# (models/fused.py); whoever consumes their results on the main stream waits for that stream first.
_side_streams: "Dict[Tuple[int, int], torch.cuda.Stream]" = {}


def note_side_stream(device: torch.device, main_stream: int, side: "torch.cuda.Stream") -> None:
    with _lock:
        _side_streams[(device.index, main_stream)] = side


def wait_side_streams(device: Optional[torch.device] = None) -> None:
    """Make every main stream that has side-stream work outstanding wait for it (device-side wait)."""
    with _lock:
        items = [(k, v) for k, v in _side_streams.items() if device is None or k[0] == device.index]
        for k, _ in items:
            del _side_streams[k]
    for (index, main), side in items:
        with torch.cuda.device(index):
            rc = _lib.load().pb2_stream_wait_stream(ctypes.c_void_p(main), ctypes.c_void_p(side.cuda_stream))
        _lib.check(rc, "pb2_stream_wait_stream")

# bumped whenever any LinearGrad (re)allocates its stacks: cached descriptors that hold stash pointers
# compare it to know when to rebuild
_stash_epoch = 0


def stash_epoch() -> int:
    return _stash_epoch


#: number of wgrad GEMMs that *accumulated* into an existing gradient (gradient accumulation over several
#: backward passes); :class:`~pipe_b200.ops.optim.FusedClipAdam` re-derives the norm when it moved
accumulated_flushes = 0


_defer_depth = 0
_nested_depth = 0
_callback_queued = False


def ensure_grad(p: nn.Parameter) -> Tensor:
    """``p.grad`` as a zero-initialised fp32 buffer if it does not exist yet."""
    if p.grad is None:
        p.grad = torch.zeros_like(p, memory_format=torch.contiguous_format)
    return p.grad


class LinearGrad:
    """Stacked operands and the deferred wgrad GEMM of one weight ``[out, in]``.

    *Replicated* mode (Pipe-inside-DDP hybrid, :class:`pipe_b200.parallel.replica_stash.ReplicaStash`): the two
    stacks are views into a symmetric slab that every data-parallel replica maps, ``k * cap`` columns wide;
    replica ``r`` stacks its micro-batches at columns ``r * cap + ...`` of its own copy *and of every peer's*
    (the transposes store to all of them), so after a barrier each replica holds all replicas' operands and
    ONE GEMM with K = all replicas' tokens, scaled by ``1/k``, writes the averaged gradient: the
    all-reduce of the weight matrices happens in the tensor core's accumulator, and what crosses NVLink is
    ``(in + out) * tokens`` bf16 instead of ``in * out`` fp32."""

    def __init__(self, weight: nn.Parameter, bias: Optional[nn.Parameter]) -> None:
        self.weight = weight
        self.bias = bias
        self.xt: Optional[Tensor] = None      # [in, cap]  bf16
        self.dyt: Optional[Tensor] = None     # [out, cap] bf16
        self.used = 0
        self.stream: Optional[int] = None       # raw handle of the stream the operands were stacked on
        self.col0 = 0                           # first column of this replica's region
        self.rep: Optional[Tuple[int, int]] = None      # (columns per replica, replicas) in replicated mode
        self.peer_deltas: Tuple[int, ...] = ()  # byte offsets from this replica's slab to each peer's
        self.high = 0                           # most columns any flush has used (replicated: tail zeroing)
        self.on_flushed = None                  # optional callable(self) after the GEMM was enqueued

    def bind_replicated(self, xt: Tensor, dyt: Tensor, replica: int, cap: int, replicas: int,
                        peer_deltas: Tuple[int, ...]) -> None:
        global _stash_epoch
        out_f, in_f = self.weight.shape
        assert xt.shape == (in_f, replicas * cap) and dyt.shape == (out_f, replicas * cap) and not self.used
        with _lock:
            self.xt, self.dyt = xt, dyt
            self.col0, self.rep, self.peer_deltas = replica * cap, (cap, replicas), tuple(peer_deltas)
            self.high = 0
            _stash_epoch += 1

    def capacity(self) -> int:
        """Columns one step may stack."""
        if self.rep is not None:
            return self.rep[0]
        return 0 if self.xt is None else self.xt.size(1)

    def _reserve(self, tokens: int) -> int:
        out_f, in_f = self.weight.shape
        need = self.used + tokens
        if self.rep is not None:
            if need > self.rep[0]:
                raise RuntimeError(f"replicated wgrad stash holds {self.rep[0]} tokens per step and replica, "
                                   f"{need} were stacked: raise max_tokens of the hybrid optimizer")
            off = self.col0 + self.used
            self.used = need
            return off
        cap = 0 if self.xt is None else self.xt.size(1)
        if need > cap or self.xt.device != self.weight.device:
            new_cap = max(64, cap)
            while new_cap < need:
                new_cap *= 2
            dev = self.weight.device
            if self.used and dev.type == "cuda":
                wait_side_streams(dev)          # the columns copied below may still be in flight on a side stream
            global _stash_epoch
            _stash_epoch += 1
            xt = torch.zeros((in_f, new_cap), dtype=torch.bfloat16, device=dev)
            dyt = torch.zeros((out_f, new_cap), dtype=torch.bfloat16, device=dev)
            if self.used:
                xt[:, : self.used].copy_(self.xt[:, : self.used])
                dyt[:, : self.used].copy_(self.dyt[:, : self.used])
            self.xt, self.dyt = xt, dyt
        off = self.used
        self.used = need
        return off

    def reserve(self, tokens: int, device: torch.device, stream: Optional[int] = None) -> int:
        """Reserve a column block of ``tokens`` columns for one micro-batch, register this
        Linear for the end-of-backward flush, make sure the bias gradient exists; returns the
        column offset.  (The caller transposes its operands into ``xt`` / ``dyt`` there.)"""
        with _lock:
            off = self._reserve(tokens)
            self.stream = _lib.raw_stream(device) if stream is None else stream
            _pending[id(self)] = self
            _arm()
        if self.bias is not None:
            ensure_grad(self.bias)
        return off

    def jobs(self, x: Tensor, dy: Tensor) -> List[Tuple[Tensor, Tensor, int, Optional[Tensor], float]]:
        """Reserve a column block for one micro-batch and return the two transpose
        jobs that fill it (``dy``'s job also accumulates the bias gradient).  Pass ``self.peer_deltas``
        to ``transpose_multi`` along with them."""
        tokens = x.size(0)
        assert dy.size(0) == tokens
        with _lock:
            off = self._reserve(tokens)
            self.stream = _lib.raw_stream(x.device)
            _pending[id(self)] = self
            _arm()
        bias_grad = ensure_grad(self.bias) if self.bias is not None else None
        return [(x, self.xt, off, None, 1.0), (dy, self.dyt, off, bias_grad, 1.0)]

    def flush(self) -> None:
        if not self.used:
            return
        w = self.weight
        grad = ensure_grad(w)
        fresh = getattr(w, "_pb2_fresh", False)
        normsq = getattr(w, "_pb2_normsq", None) if fresh else None
        if self.rep is None:
            gemm_nt(self.dyt[:, : self.used], self.xt[:, : self.used], out=grad, accumulate=not fresh,
                    sumsq=normsq, stream=self.stream)
        else:
            cap, k = self.rep
            if self.used < self.high:
                # a shorter step than before: stale columns of every replica's region would be summed in
                lib = _lib.load()
                with _lib.on_device(self.xt.device):
                    for t in (self.xt, self.dyt):
                        for r in range(k):
                            part = t[:, r * cap + self.used: r * cap + self.high]
                            rc = lib.pb2_memset2d(ctypes.c_void_p(part.data_ptr()), ctypes.c_size_t(t.stride(0) * 2), ctypes.c_int(0),
                                                  ctypes.c_size_t(part.size(1) * 2), ctypes.c_size_t(part.size(0)),
                                                  ctypes.c_void_p(self.stream))
                            _lib.check(rc, "pb2_memset2d")
            self.high = self.used
            gemm_nt(self.dyt, self.xt, out=grad, accumulate=not fresh, alpha=1.0 / k, sumsq=normsq, stream=self.stream)
        if fresh:
            w._pb2_fresh = False
            w._pb2_sumsq_done = normsq is not None
        else:
            # a second backward pass before the optimizer step: the squared norm summed by the first pass's
            # epilogue no longer describes this gradient — the optimizer has to measure it again
            global accumulated_flushes
            accumulated_flushes += 1
            w._pb2_sumsq_done = False
        self.used = 0
        if self.on_flushed is not None:
            self.on_flushed(self)


class StackStash:
    """One column reservation per micro-batch for ALL linears of a run of layers.

    Every linear of a stage's layer run sees the same tokens in the same order, so their stacks
    fill in lock-step; reserving per linear (4 per layer, lock + bookkeeping each) was a visible
    share of the host time of a backward cell.  The fast path is a handful of attribute writes."""

    def __init__(self, lgs: "List[LinearGrad]") -> None:
        self.lgs = list(lgs)
        self._items = {id(lg): lg for lg in self.lgs}
        self.cap = 0
        self.col0 = 0
        self.epoch = -1

    def reserve(self, tokens: int, device: torch.device, stream: int) -> int:
        """Returns the column offset shared by every linear of the run."""
        with _lock:
            lgs = self.lgs
            used = lgs[0].used
            need = used + tokens
            if self.epoch != _stash_epoch or need > self.cap:
                used = lgs[0].used
                offs = {lg._reserve(tokens) for lg in lgs}
                if len(offs) != 1:
                    raise RuntimeError("the linears of a layer run were used unevenly within one step "
                                       "(some through the per-layer path): cannot share a stash offset")
                self.cap = min(lg.capacity() for lg in lgs)
                self.col0 = lgs[0].col0
                self.epoch = _stash_epoch
                for lg in lgs:
                    lg.stream = stream
                    if lg.bias is not None:
                        ensure_grad(lg.bias)
            else:
                if lgs[-1].used != used:
                    raise RuntimeError("the linears of a layer run were used unevenly within one step")
                for lg in lgs:
                    lg.used = need
                    lg.stream = stream
            _pending.update(self._items)
            _arm()
        return self.col0 + used


def _arm() -> None:
    """Make sure something will flush: queue an engine callback unless an outer
    ``deferred()`` block owns the flush."""
    global _callback_queued
    if _defer_depth > 0 or _nested_depth > 0 or _callback_queued:
        return
    try:
        torch.autograd.Variable._execution_engine.queue_callback(_engine_done)
        _callback_queued = True
    except RuntimeError:
        # not inside a backward pass (direct call): the caller flushes explicitly
        pass


def _engine_done() -> None:
    global _callback_queued
    with _lock:
        _callback_queued = False
    flush_pending()


#: ``hook(todo)`` callables run by :func:`flush_pending` right before / after the wgrad GEMMs of a non-empty
#: flush (after the side streams were joined).  The hybrid layout hangs its replica barriers and the launch of
#: the small-gradient reduction here.
pre_flush_hooks: List = []
post_flush_hooks: List = []


def flush_pending() -> None:
    """Run the deferred wgrad GEMM of every Linear that stacked operands."""
    with _lock:
        todo = list(_pending.values())
        _pending.clear()
    if not todo:
        return
    wait_side_streams()
    for hook in list(pre_flush_hooks):
        hook(todo)
    for lg in todo:
        lg.flush()
    for hook in list(post_flush_hooks):
        hook(todo)


@contextmanager
def outer_backward() -> Iterator[None]:
    """Wrap a *re-entrant* ``torch.autograd.backward`` call (checkpoint recompute).

    An engine callback queued from inside the nested pass would fire when the nested pass ends,
    i.e. once per checkpointed cell, and the "one wgrad GEMM per mini-batch" deferral would
    degenerate into one per micro-batch.  So the callback is queued here, in the context of the
    *outer* pass, and arming is suppressed while the nested pass runs."""
    global _nested_depth
    with _lock:
        _arm()
        _nested_depth += 1
    try:
        yield
    finally:
        with _lock:
            _nested_depth -= 1


@contextmanager
def deferred() -> Iterator[None]:
    """Hold all wgrad flushes until the block exits (nestable)."""
    global _defer_depth
    with _lock:
        _defer_depth += 1
    try:
        yield
    finally:
        with _lock:
            _defer_depth -= 1
            last = _defer_depth == 0
        if last:
            flush_pending()
