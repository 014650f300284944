"""``Copy`` and ``Wait``: the two autograd Functions that carry a micro-batch
across a stage boundary.

* ``Copy(prev_stream, next_stream, *values)`` moves every tensor to the device
  of ``next_stream`` with an asynchronous peer copy that is ordered on *both*
  copy streams, and tells the caching allocator about both uses.  Its backward
  ships the gradients the other way over the same stream pair.
* ``Wait(prev_stream, next_stream, *values)`` inserts "next waits for prev" in
  forward and "prev waits for next" in backward; it moves no data.

On B200 the copy is a ``cudaMemcpyPeerAsync`` over NVLink 5 (every peer is one
NVSwitch hop away, so skip routes cost the same as neighbour routes).

Parity: reference ``pipeline.py:51-60`` (call sites), ``README.md:185-237``
(Copy) and ``README.md:324-369`` (Wait); SURVEY.md A3.
"""
from __future__ import annotations

from collections import deque
from typing import Any, Deque, List, Optional, Tuple

import torch
from torch import Tensor

from .stream import (AbstractStream, current_stream, get_device, record_stream,
                     use_stream, wait_stream)

__all__ = ["Copy", "Wait", "Context"]


class Context:
    """Bare stand-in for an autograd ctx, so ``Copy.forward``/``backward`` can be
    driven by hand (the skip portals do that)."""

    prev_stream: AbstractStream
    next_stream: AbstractStream


def _move(tensor: Tensor, src_stream: AbstractStream, dst_stream: AbstractStream,
          consumer: AbstractStream) -> Tensor:
    """Asynchronous copy of one tensor from src_stream's device to dst_stream's.

    Must be called with both copy streams current.  ``consumer`` is the stream that
    was current on the destination device *before* the copy streams were entered
    — the compute stream whose kernels will read the result.  The source block
    stays reserved for ``src_stream``; the destination block is allocated on
    ``dst_stream`` (its pool) and additionally recorded on ``consumer``, so that
    dropping the Python reference while compute kernels still read it cannot hand
    the block to the next allocation on the copy stream (reference README.md:197-208,
    218-231: ``output_stream`` / ``input_stream`` are captured outside ``use_stream``).
    """
    moved = tensor.to(get_device(dst_stream), non_blocking=True)
    record_stream(tensor, src_stream)
    record_stream(moved, consumer)
    return moved


class Copy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                *values: Any) -> Tuple[Any, ...]:
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream

        out: List[Any] = []
        consumer = current_stream(get_device(next_stream))      # captured before the copy streams are entered
        with use_stream(prev_stream), use_stream(next_stream):
            for v in values:
                out.append(_move(v, prev_stream, next_stream, consumer) if torch.is_tensor(v) else v)
        return tuple(out)

    @staticmethod
    def backward(ctx, *grads: Optional[Tensor]) -> Tuple[Optional[Tensor], ...]:  # type: ignore[override]
        prev_stream, next_stream = ctx.prev_stream, ctx.next_stream

        # Walk the gradients back-to-front and drop our reference to each as soon
        # as its copy is queued, so the sender-side blocks free up early.
        moved: Deque[Optional[Tensor]] = deque(maxlen=len(grads))
        pending = list(grads)
        consumer = current_stream(get_device(prev_stream))
        with use_stream(prev_stream), use_stream(next_stream):
            while pending:
                g = pending.pop()
                moved.appendleft(None if g is None else _move(g, next_stream, prev_stream, consumer))
        return (None, None) + tuple(moved)


class Wait(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                *values: Any) -> Tuple[Any, ...]:
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream
        wait_stream(next_stream, prev_stream)
        return tuple(v.detach() if torch.is_tensor(v) else v for v in values)

    @staticmethod
    def backward(ctx, *grads: Optional[Tensor]) -> Tuple[Optional[Tensor], ...]:  # type: ignore[override]
        wait_stream(ctx.prev_stream, ctx.next_stream)
        return (None, None) + grads
