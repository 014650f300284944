"""Two-node activation checkpointing for pipeline cells.

``torch.utils.checkpoint`` recomputes *inside* the backward of the checkpointed
node, i.e. only once the gradient has already arrived.  In a pipeline the
gradient arrives over a copy stream, and waiting for it before starting the
recompute wastes exactly the time the copy takes.  So a cell is split into two
autograd nodes that share a one-slot mailbox:

* ``Checkpoint``  — forward: run the stage under ``no_grad`` and remember the
  RNG state.  backward: pop the recomputed graph from the mailbox, backprop
  through it, return the input gradients.
* ``Recompute``   — forward: nothing (returns a phony).  backward: restore the
  RNG state, re-run the stage under ``enable_grad`` on detached leaves, push
  ``(output, leaves)`` into the mailbox.

``Recompute`` hangs off a phony that is forked from the cell's output *after*
the trailing ``Wait`` and joined back in, so in backward it fires before the
``Wait`` that blocks on the incoming gradient copy: the recompute kernels are
queued first and overlap the transfer.  ``Fork`` has two outputs, so
``Checkpoint.backward`` (behind ``Fork``) can never run before ``Recompute``.

Dropout determinism: CPU and device RNG states are captured in the forward and
reinstated for the recompute.  The fused sm_100a dropout kernels of this
package take explicit seeds; inside a checkpointed cell those come from a base
owned by the cell (``autograd/seeds.py``), not from the process-global CPU
generator that other stage threads advance concurrently, so they replay
bit-identically on any thread.

Parity: reference ``pipeline.py:16,195-214`` (call sites),
``README.md:446-545`` (excerpts of the absent ``checkpoint`` module),
``pptx:slide2-3``; SURVEY.md A5 / Appendix A "checkpoint".
"""
from __future__ import annotations

import threading
from collections import deque
from contextlib import contextmanager
from typing import Any, Deque, Generator, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from ..microbatch import Batch
from .dependency import fork, join
from .phony import get_phony
from .seeds import SeedBase, seed_scope

__all__ = [
    "Checkpointing", "Checkpoint", "Recompute", "checkpoint",
    "is_checkpointing", "is_recomputing", "enable_checkpointing", "enable_recomputing",
    "save_rng_states", "restore_rng_states",
]

RNGStates = Tuple[Tensor, Optional[Tensor]]          # (cpu state, device state)
Recomputed = Tuple[Any, Tuple[Any, ...]]             # (output, detached input leaves)


# ----------------------------------------------------------------------------
# thread-local mode flags
# ----------------------------------------------------------------------------
class _Mode(threading.local):
    def __init__(self) -> None:
        self.checkpointing = False
        self.recomputing = False


_mode = _Mode()


@contextmanager
def enable_checkpointing() -> Generator[None, None, None]:
    prev, _mode.checkpointing = _mode.checkpointing, True
    try:
        yield
    finally:
        _mode.checkpointing = prev


@contextmanager
def enable_recomputing() -> Generator[None, None, None]:
    prev, _mode.recomputing = _mode.recomputing, True
    try:
        yield
    finally:
        _mode.recomputing = prev


def is_checkpointing() -> bool:
    """True while a stage forward runs as the (grad-free) first pass of a
    checkpointed cell.  Lets a layer skip side effects that must happen once."""
    return _mode.checkpointing


def is_recomputing() -> bool:
    """True while a stage forward is being re-run inside backward.  Layers with
    running statistics (DeferredBatchNorm) use it to avoid double counting."""
    return _mode.recomputing


# ----------------------------------------------------------------------------
# RNG capture / replay
# ----------------------------------------------------------------------------
class _RngSlot(deque):
    """The one-slot RNG mailbox of a cell; also carries the cell's fused-dropout seed base."""

    def __init__(self) -> None:
        super().__init__(maxlen=1)
        self.seed_base = SeedBase()


def _seed_base(rng_states) -> SeedBase:
    base = getattr(rng_states, "seed_base", None)
    return base if base is not None else SeedBase()


def save_rng_states(device: torch.device, rng_states: Deque[RNGStates]) -> None:
    cpu_state = torch.get_rng_state()
    dev_state: Optional[Tensor] = None
    if device.type == "cuda":
        dev_state = torch.cuda.get_rng_state(device)
    rng_states.append((cpu_state, dev_state))


@contextmanager
def restore_rng_states(device: torch.device, rng_states: Deque[RNGStates]) -> Generator[None, None, None]:
    """Reinstate the captured states inside a forked RNG scope, so the replay
    does not disturb the generators seen by the rest of backward."""
    cpu_state, dev_state = rng_states.pop()
    fork_devices: List[torch.device] = [device] if device.type == "cuda" else []
    with torch.random.fork_rng(fork_devices):
        torch.set_rng_state(cpu_state)
        if dev_state is not None:
            torch.cuda.set_rng_state(dev_state, device)
        yield


# ----------------------------------------------------------------------------
# the two nodes
# ----------------------------------------------------------------------------
def _detach_leaves(inputs: Sequence[Any]) -> Tuple[Any, ...]:
    leaves = []
    for x in inputs:
        if torch.is_tensor(x):
            leaves.append(x.detach().requires_grad_(x.requires_grad))
        else:
            leaves.append(x)
    return tuple(leaves)


class Checkpoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, phony: Tensor, recomputed: Deque[Recomputed], rng_states: Deque[RNGStates],  # type: ignore[override]
                function, input_atomic: bool, *inputs: Any):
        ctx.recomputed = recomputed
        ctx.rng_states = rng_states

        device = next(x for x in inputs if torch.is_tensor(x)).device
        save_rng_states(device, rng_states)

        # fused-dropout seeds of this cell come from a base only this cell owns (autograd/seeds.py)
        with torch.no_grad(), enable_checkpointing(), seed_scope(_seed_base(rng_states)):
            output = function(inputs[0]) if input_atomic else function(*inputs)
        return output

    @staticmethod
    def backward(ctx, *grad_output: Optional[Tensor]):  # type: ignore[override]
        output, leaves = ctx.recomputed.pop()

        outputs = (output,) if torch.is_tensor(output) else tuple(output)
        pairs = [(y, g) for y, g in zip(outputs, grad_output)
                 if torch.is_tensor(y) and y.requires_grad]
        if pairs:
            # The nested pass must not end the mini-batch-wide deferral of weight gradients
            # (pipe_b200.ops.wgrad): queue its flush on the outer pass instead.
            from ..ops import wgrad
            with wgrad.outer_backward():
                torch.autograd.backward([y for y, _ in pairs], [g for _, g in pairs])

        grads: List[Optional[Tensor]] = [None] * 5
        grads.extend(x.grad if torch.is_tensor(x) else None for x in leaves)
        return tuple(grads)


class Recompute(torch.autograd.Function):
    @staticmethod
    def forward(ctx, phony: Tensor, recomputed: Deque[Recomputed], rng_states: Deque[RNGStates],  # type: ignore[override]
                function, input_atomic: bool, *inputs: Any) -> Tensor:
        ctx.recomputed = recomputed
        ctx.rng_states = rng_states
        ctx.function = function
        ctx.input_atomic = input_atomic
        # Tensors go through save_for_backward so an in-place edit between the
        # forward and the replay is caught by the version counter.
        ctx.slots = [None if torch.is_tensor(x) else x for x in inputs]
        ctx.tensor_at = [k for k, x in enumerate(inputs) if torch.is_tensor(x)]
        ctx.save_for_backward(*(inputs[k] for k in ctx.tensor_at))
        return phony

    @staticmethod
    def backward(ctx, *grad_output: Tensor):  # type: ignore[override]
        inputs = list(ctx.slots)
        for k, t in zip(ctx.tensor_at, ctx.saved_tensors):
            inputs[k] = t
        leaves = _detach_leaves(inputs)
        device = inputs[ctx.tensor_at[0]].device

        base = _seed_base(ctx.rng_states)
        with restore_rng_states(device, ctx.rng_states):
            with torch.enable_grad(), enable_recomputing(), seed_scope(base):
                output = ctx.function(leaves[0]) if ctx.input_atomic else ctx.function(*leaves)

        ctx.recomputed.append((output, leaves))
        return (None,) * (5 + len(inputs))


class Checkpointing:
    """Builds the Checkpoint / Recompute pair for one ``(micro-batch, stage)``."""

    def __init__(self, function, batch: Batch) -> None:
        self.function = function
        self.batch = batch
        # One-slot mailboxes shared by the two nodes.
        self.recomputed: Deque[Recomputed] = deque(maxlen=1)
        self.rng_states: Deque[RNGStates] = _RngSlot()

    def _inputs(self) -> Tuple[bool, Tuple[Any, ...]]:
        if self.batch.atomic:
            return True, (self.batch.tensor,)
        return False, tuple(self.batch)

    def checkpoint(self) -> Batch:
        """First pass: run the stage without building a graph."""
        input_atomic, inputs = self._inputs()
        # A grad-requiring phony keeps the Checkpoint node in the graph even when
        # no real input requires grad (stage 0 fed by integer tokens).
        phony = get_phony(self.batch.get_device(), requires_grad=True)
        output = Checkpoint.apply(phony, self.recomputed, self.rng_states,
                                  self.function, input_atomic, *inputs)

        # Integer/bool outputs cannot carry gradient: cut them loose.
        if isinstance(output, tuple):
            output = tuple(y.detach() if torch.is_tensor(y) and not y.is_floating_point() else y
                           for y in output)
        return Batch(output)

    def recompute(self, batch: Batch) -> None:
        """Attach the Recompute node to ``batch`` (the cell's output), in place."""
        input_atomic, inputs = self._inputs()
        k = batch.find_tensor_idx()
        batch[k], phony = fork(batch[k])
        phony = Recompute.apply(phony, self.recomputed, self.rng_states,
                                self.function, input_atomic, *inputs)
        batch[k] = join(batch[k], phony)


def checkpoint(function, input):
    """Checkpoint a plain function call outside the pipeline (drop-in for
    ``torch.utils.checkpoint.checkpoint`` with the early-recompute property)."""
    batch = Batch(input)
    chk = Checkpointing(function, batch)
    out = chk.checkpoint()
    chk.recompute(out)
    return out.values
