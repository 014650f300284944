"""Portals: how a stashed tensor jumps from stage ``a`` to stage ``b > a``
without riding through the stages in between.

A :class:`Portal` is a small mailbox that holds the skip tensor on the forward
way and its gradient on the way back.  Three data-free autograd nodes tie the
mailbox into the micro-batch's graph through phonies:

    stash side (stage a)   PortalBlue    tensor -> phony      (join into batch)
    fence  (main thread)   PortalCopy    phony  -> phony      (moves portal.tensor a -> b)
    pop side   (stage b)   PortalOrange  phony  -> tensor     (fork from batch)

Backward runs them in the opposite order: Orange parks the gradient in the
portal, Copy moves it b -> a over the same copy-stream pair, Blue hands it to
the layer that stashed.  The tensor is reference-counted by ``tensor_life`` so
it is dropped at its last planned use: 2 uses normally
(copy is free, blue + orange), 3 when the stash ran under checkpointing
(orange runs again during recompute), and 1 when re-stashed by a recompute.

On B200 every stage pair is one NVSwitch hop apart, so a far skip costs the
same as a neighbour copy.

Parity: reference ``pipeline.py:134-138`` (the only call site shipped);
SURVEY.md §3.4, A8.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor

from ..autograd.copy import Copy
from ..autograd.phony import get_phony
from ..autograd.stream import AbstractStream, get_device

__all__ = ["Portal", "PortalBlue", "PortalOrange", "PortalCopy"]


class Portal:
    def __init__(self, tensor: Optional[Tensor], tensor_life: int) -> None:
        self.tensor: Optional[Tensor] = None
        self.tensor_life = 0
        self.grad: Optional[Tensor] = None
        self.put_tensor(tensor, tensor_life)

    # -- graph hooks -----------------------------------------------------
    def blue(self) -> Tensor:
        """Phony that depends on the stashed tensor (stash side)."""
        tensor = self.use_tensor()
        if tensor is None:
            return get_phony(torch.device("cpu"), requires_grad=False)
        return PortalBlue.apply(self, tensor)

    def orange(self, phony: Tensor) -> Optional[Tensor]:
        """The skip tensor, depending on ``phony`` (pop side)."""
        self.check_tensor_life()
        if self.tensor is None:
            return self.use_tensor()
        return PortalOrange.apply(self, phony)

    def copy(self, prev_stream: AbstractStream, next_stream: AbstractStream, phony: Tensor) -> Tensor:
        """Move the held tensor to ``next_stream``'s device; returns a phony."""
        if self.tensor is None:
            return get_phony(torch.device("cpu"), requires_grad=False)
        return PortalCopy.apply(self, prev_stream, next_stream, phony)

    # -- tensor life -------------------------------------------------------
    def check_tensor_life(self) -> None:
        if self.tensor_life <= 0:
            raise RuntimeError("tensor in portal has been removed")

    def put_tensor(self, tensor: Optional[Tensor], tensor_life: int) -> None:
        self.tensor_life = tensor_life
        self.tensor = tensor if tensor_life > 0 else None

    def use_tensor(self) -> Optional[Tensor]:
        """Hand out the tensor and burn one life; the last use drops it."""
        self.check_tensor_life()
        tensor = self.tensor
        self.tensor_life -= 1
        if self.tensor_life <= 0:
            self.tensor = None
        return tensor

    # -- gradient mailbox --------------------------------------------------
    def put_grad(self, grad: Tensor) -> None:
        self.grad = grad

    def use_grad(self) -> Tensor:
        if self.grad is None:
            raise RuntimeError("grad in portal has been removed or never set")
        grad, self.grad = self.grad, None
        return grad


class PortalBlue(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, tensor: Tensor) -> Tensor:  # type: ignore[override]
        ctx.portal = portal
        return get_phony(tensor.device, requires_grad=False).detach()

    @staticmethod
    def backward(ctx, grad_phony: Tensor):  # type: ignore[override]
        return None, ctx.portal.use_grad()


class PortalOrange(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, phony: Tensor) -> Tensor:  # type: ignore[override]
        ctx.portal = portal
        tensor = portal.use_tensor()
        assert tensor is not None
        return tensor.detach()

    @staticmethod
    def backward(ctx, grad: Tensor):  # type: ignore[override]
        ctx.portal.put_grad(grad)
        return None, None


class PortalCopy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, portal: Portal, prev_stream: AbstractStream, next_stream: AbstractStream,  # type: ignore[override]
                phony: Tensor) -> Tensor:
        ctx.portal = portal
        assert portal.tensor is not None
        (portal.tensor,) = Copy.forward(ctx, prev_stream, next_stream, portal.tensor)
        return get_phony(get_device(next_stream), requires_grad=False).detach()

    @staticmethod
    def backward(ctx, grad_phony: Tensor):  # type: ignore[override]
        portal = ctx.portal
        assert portal.grad is not None
        _, _, portal.grad = Copy.backward(ctx, portal.grad)
        return None, None, None, None
