"""Static routing table of skip connections across stages.

``inspect_skip_layout(partitions)`` walks the stages once at ``Pipe``
construction and records, for every ``(namespace, name)`` that is stashed in
stage ``a`` and popped in stage ``b``, the route ``a -> b``.  The scheduler
asks ``copy_policy(b)`` at every fence to learn which portals must be moved to
stage ``b`` before it runs.

Parity: reference ``pipe.py:20,348`` and ``pipeline.py:20,136`` (call sites; the
module itself is absent); SURVEY.md A8.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Tuple

from torch import nn

from .namespace import Namespace

__all__ = ["SkipLayout", "inspect_skip_layout"]


class SkipLayout:
    """Routes of skip tensors, indexed by the stage that pops them."""

    def __init__(self, num_partitions: int,
                 skip_routes: Dict[Tuple[Namespace, str], Tuple[int, int]]) -> None:
        self.by_ns_name = skip_routes
        self.by_partition: List[List[Tuple[int, Namespace, str]]] = [[] for _ in range(num_partitions)]
        for (ns, name), (prev_j, next_j) in skip_routes.items():
            self.by_partition[next_j].append((prev_j, ns, name))
        for routes in self.by_partition:
            routes.sort(key=lambda r: (r[0], r[2]))

    def copy_policy(self, next_j: int) -> Iterable[Tuple[int, Namespace, str]]:
        """Yield ``(prev_j, ns, name)`` for each skip that ends at ``next_j`` and
        starts on a *different* stage (same-stage skips never leave the device)."""
        for prev_j, ns, name in self.by_partition[next_j]:
            if prev_j != next_j:
                yield (prev_j, ns, name)

    def requires_copy(self, ns: Namespace, name: str) -> bool:
        """True when the skip crosses a stage boundary and so needs a portal."""
        prev_j, next_j = self.by_ns_name.get((ns, name), (-1, -1))
        return prev_j != next_j


def inspect_skip_layout(partitions: Iterable[nn.Sequential]) -> SkipLayout:
    """Build the :class:`SkipLayout` of already-verified partitions."""
    from .skippable import Skippable  # cycle guard

    partitions = list(partitions)
    routes: Dict[Tuple[Namespace, str], Tuple[int, int]] = {}
    stashed_at: Dict[Tuple[Namespace, str], int] = {}

    for j, partition in enumerate(partitions):
        for layer in partition:
            if not isinstance(layer, Skippable):
                continue
            for ns, name in layer.stashable():
                stashed_at[(ns, name)] = j
            for ns, name in layer.poppable():
                routes[(ns, name)] = (stashed_at.pop((ns, name)), j)

    return SkipLayout(len(partitions), routes)
