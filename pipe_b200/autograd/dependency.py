"""``fork`` / ``join``: data-free edges in the autograd graph.

``fork(x)`` returns ``(x, phony)``; ``join(y, phony)`` returns ``y``.  Forward
does nothing.  In backward the node behind ``y`` cannot hand its gradient on
until the *Join* ran, and the *Fork* on ``x`` cannot run until the phony's
(empty) gradient has arrived from the Join — so everything upstream of ``x``
is ordered after everything downstream of ``y``.  The scheduler uses this to
make micro-batch ``i-1`` run its stage-``j`` backward after micro-batch ``i``
(reference ``pipeline.py:43-48,129-132``), and checkpointing uses it to run
Recompute before Checkpoint's backward.

Parity: reference ``README.md:106-183`` (excerpts of the absent ``dependency``
module); SURVEY.md Appendix A "dependency".
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from .phony import get_phony

__all__ = ["fork", "join", "Fork", "Join"]


class Fork(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tensor: Tensor) -> Tuple[Tensor, Tensor]:  # type: ignore[override]
        phony = get_phony(tensor.device, requires_grad=False)
        return tensor.detach(), phony.detach()

    @staticmethod
    def backward(ctx, grad_tensor: Tensor, grad_phony: Tensor) -> Tensor:  # type: ignore[override]
        return grad_tensor


class Join(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tensor: Tensor, phony: Tensor) -> Tensor:  # type: ignore[override]
        return tensor.detach()

    @staticmethod
    def backward(ctx, grad_tensor: Tensor):  # type: ignore[override]
        return grad_tensor, None


def fork(tensor: Tensor) -> Tuple[Tensor, Tensor]:
    """Branch a phony off ``tensor``.  Outside autograd (grad disabled or the
    tensor does not require grad) this degenerates to a cached non-grad phony."""
    if torch.is_grad_enabled() and tensor.requires_grad:
        tensor, phony = Fork.apply(tensor)
    else:
        phony = get_phony(tensor.device, requires_grad=False)
    return tensor, phony


def join(tensor: Tensor, phony: Tensor) -> Tensor:
    """Merge a phony into ``tensor``'s lineage."""
    if torch.is_grad_enabled() and (tensor.requires_grad or phony.requires_grad):
        tensor = Join.apply(tensor, phony)
    return tensor
