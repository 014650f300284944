"""Per-cell seeds for the counter-based dropout of the fused kernels.

The reference replays dropout under checkpointing by saving and restoring the generators a stage
draws from: the CPU generator and the stage device's CUDA generator (``README.md:472-482, 525-537``).
That is race-free there because a device's CUDA generator is only touched by that device's worker /
autograd thread.  The fused kernels here take an explicit 63-bit seed per call instead of mutating
a device generator, and that seed must not come straight from the process-global CPU generator:
stage workers, the inline last cell on the calling thread and the per-device autograd threads all
advance it concurrently, so the value drawn in a recompute could differ from the forward's.

Rule: inside a :func:`seed_scope` (one per checkpointed cell, entered by ``Checkpoint.forward`` and
again by ``Recompute.backward`` on whatever thread runs it) :func:`draw_seed` serves the k-th call
from ``mix(base, k)``; the base is drawn from the CPU generator once — lazily, by the first call in
the forward, under a lock — and kept on the scope's owner, so the replay sees the same sequence
without touching any shared generator.  Outside a scope a draw is a locked read of the CPU
generator (``torch.manual_seed`` still decides every mask).
"""
from __future__ import annotations

import threading
from contextlib import contextmanager
from typing import Iterator, List, Optional

import torch

__all__ = ["draw_seed", "seed_scope", "SeedBase"]

_MASK63 = (1 << 63) - 1
_MASK64 = (1 << 64) - 1
_local = threading.local()
_generator_lock = threading.Lock()


def _fresh() -> int:
    with _generator_lock:
        return int(torch.empty((), dtype=torch.int64).random_().item())


def _mix(base: int, k: int) -> int:
    """SplitMix64 finaliser of ``base + k * golden``: distinct, well-spread seeds per call index."""
    z = (base + k * 0x9E3779B97F4A7C15) & _MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK64
    return (z ^ (z >> 31)) & _MASK63


class SeedBase:
    """Holder of one cell's base seed (``None`` until the cell draws for the first time)."""

    __slots__ = ("value",)

    def __init__(self, value: Optional[int] = None) -> None:
        self.value = value


@contextmanager
def seed_scope(owner: SeedBase) -> Iterator[None]:
    """Serve this thread's :func:`draw_seed` calls from ``owner``'s base, call index starting at 0."""
    stack: List[list] = _local.__dict__.setdefault("stack", [])
    stack.append([owner, 0])
    try:
        yield
    finally:
        stack.pop()


def draw_seed() -> int:
    stack = _local.__dict__.get("stack")
    if not stack:
        return _fresh()
    top = stack[-1]
    owner: SeedBase = top[0]
    if owner.value is None:
        owner.value = _fresh()
    top[1] += 1
    return _mix(owner.value, top[1])
