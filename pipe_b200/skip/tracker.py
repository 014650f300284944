"""Per-micro-batch bookkeeping of stashed skip tensors.

A plain :class:`SkipTracker` keeps stashes in a dict — enough when a skippable
model runs un-piped or when stash and pop live in the same stage.
:class:`SkipTrackerThroughPotals` (the spelling is public API of the reference,
``pipeline.py:21``) additionally routes stage-crossing skips through
:class:`~.portal.Portal` objects so the tensor travels directly between the two
stages on the micro-batch's copy streams.

The active tracker is thread-local: the scheduler installs micro-batch ``i``'s
tracker around each task with :func:`use_skip_tracker`
(reference ``pipeline.py:208,228``).

Parity: reference ``pipeline.py:21,113,136-138,208,228``; SURVEY.md §3.4, A8.
"""
from __future__ import annotations

import threading
from contextlib import contextmanager
from typing import Dict, Generator, Optional, Tuple

from torch import Tensor

from ..autograd.checkpoint import is_checkpointing
from ..autograd.dependency import fork, join
from ..autograd.stream import AbstractStream
from ..microbatch import Batch
from .layout import SkipLayout
from .namespace import Namespace
from .portal import Portal

__all__ = ["SkipTracker", "SkipTrackerThroughPotals", "use_skip_tracker", "current_skip_tracker"]


class SkipTracker:
    """Dict-backed tracker for skips that never leave their stage."""

    def __init__(self) -> None:
        self.tensors: Dict[Tuple[Namespace, str], Optional[Tensor]] = {}

    def save(self, batch: Batch, ns: Namespace, name: str, tensor: Optional[Tensor]) -> None:
        self.tensors[(ns, name)] = tensor

    def load(self, batch: Batch, ns: Namespace, name: str) -> Optional[Tensor]:
        return self.tensors.pop((ns, name))

    def copy(self, batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream,
             ns: Namespace, name: str) -> None:
        raise TypeError("copy is not supported for non-portal skip tensors")


class SkipTrackerThroughPotals(SkipTracker):
    """Tracker that sends stage-crossing skips through portals."""

    def __init__(self, skip_layout: SkipLayout) -> None:
        super().__init__()
        self.skip_layout = skip_layout
        self.portals: Dict[Tuple[Namespace, str], Portal] = {}

    def save(self, batch: Batch, ns: Namespace, name: str, tensor: Optional[Tensor]) -> None:
        if not self.skip_layout.requires_copy(ns, name):
            super().save(batch, ns, name, tensor)
            return

        key = (ns, name)
        # ``tensor`` may be None (a layer may stash "nothing"); the portal then
        # only hands out CPU phonies and None.
        portal = self.portals.get(key)
        if portal is None:
            # First stash.  Planned uses: PortalBlue now, PortalOrange at the pop;
            # under checkpointing the pop side runs a second time in recompute.
            life = 3 if is_checkpointing() else 2
            portal = self.portals[key] = Portal(tensor, life)
        else:
            # Re-stash by a recompute: only PortalBlue (for its backward edge)
            # will look at it again.
            portal.put_tensor(tensor, 1)

        phony = portal.blue()
        k = batch.find_tensor_idx()
        batch[k] = join(batch[k], phony)

    def load(self, batch: Batch, ns: Namespace, name: str) -> Optional[Tensor]:
        if not self.skip_layout.requires_copy(ns, name):
            return super().load(batch, ns, name)

        portal = self.portals[(ns, name)]
        k = batch.find_tensor_idx()
        batch[k], phony = fork(batch[k])
        return portal.orange(phony)

    def copy(self, batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream,
             ns: Namespace, name: str) -> None:
        assert self.skip_layout.requires_copy(ns, name)

        k = batch.find_tensor_idx()
        batch[k], phony = fork(batch[k])
        phony = self.portals[(ns, name)].copy(prev_stream, next_stream, phony)
        batch[k] = join(batch[k], phony)


class _Local(threading.local):
    def __init__(self) -> None:
        self.tracker: Optional[SkipTracker] = None


_local = _Local()


@contextmanager
def use_skip_tracker(tracker: SkipTracker) -> Generator[None, None, None]:
    prev, _local.tracker = _local.tracker, tracker
    try:
        yield
    finally:
        _local.tracker = prev


def current_skip_tracker() -> SkipTracker:
    """The tracker of the running micro-batch; a fresh plain tracker when a
    skippable model runs outside ``Pipe``."""
    if _local.tracker is None:
        _local.tracker = SkipTracker()
    return _local.tracker
