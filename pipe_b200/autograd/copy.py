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
from typing import Any, Deque, List, Optional, Sequence, Tuple

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


def _move(tensor: Tensor, src_stream: AbstractStream, dst_stream: AbstractStream) -> Tensor:
    """Asynchronous copy of one tensor from src_stream's device to dst_stream's.

    Must be called with both streams current.  The source block stays reserved
    for ``src_stream`` and the destination block for the stream that is current
    on the destination device, until the copy has drained.
    """
    moved = tensor.to(get_device(dst_stream), non_blocking=True)
    record_stream(tensor, src_stream)
    # `moved` was allocated on dst device's *current* stream, which is dst_stream
    # here; recording makes that explicit if an outer context differs.
    record_stream(moved, current_stream(get_device(dst_stream)))
    return moved


class Copy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                *values: Any) -> Tuple[Any, ...]:
        ctx.prev_stream = prev_stream
        ctx.next_stream = next_stream

        out: List[Any] = []
        with use_stream(prev_stream), use_stream(next_stream):
            for v in values:
                out.append(_move(v, prev_stream, next_stream) if torch.is_tensor(v) else v)
        return tuple(out)

    @staticmethod
    def backward(ctx, *grads: Optional[Tensor]) -> Tuple[Optional[Tensor], ...]:  # type: ignore[override]
        prev_stream, next_stream = ctx.prev_stream, ctx.next_stream

        # Walk the gradients back-to-front and drop our reference to each as soon
        # as its copy is queued, so the sender-side blocks free up early.
        moved: Deque[Optional[Tensor]] = deque(maxlen=len(grads))
        pending = list(grads)
        with use_stream(prev_stream), use_stream(next_stream):
            while pending:
                g = pending.pop()
                moved.appendleft(None if g is None else _move(g, next_stream, prev_stream))
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
