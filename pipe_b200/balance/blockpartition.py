"""Contiguous block partition of a cost sequence.

``solve(costs, k)`` cuts ``costs`` into ``k`` non-empty contiguous blocks so that
the heaviest block is as light as possible (and, among those, the sum of squared
block weights is smallest — i.e. the rest is as even as possible).  It is an
exact O(k * n^2) dynamic programme; n is a layer count, so this is instant.

Parity: the reference only advertises the balancers (``pipe.py:42-58``); the
upstream helper was a Barany-Grinberg heuristic.  SURVEY.md A11.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

__all__ = ["solve"]


def solve(sequence: Sequence[float], partitions: int = 1) -> List[List[float]]:
    """Split ``sequence`` into ``partitions`` contiguous, non-empty blocks with a
    minimal maximum block sum.  Returns the blocks themselves."""
    if partitions < 1:
        raise ValueError(f"partitions must be a positive integer ({partitions} < 1)")
    n = len(sequence)
    if n < partitions:
        raise ValueError(f"sequence is shorter than intended partitions ({n} < {partitions})")

    prefix = [0.0]
    for c in sequence:
        prefix.append(prefix[-1] + float(c))

    def weight(a: int, b: int) -> float:
        return prefix[b] - prefix[a]

    INF = (float("inf"), float("inf"))
    # best[p][e] = (max block, sum of squares) for the first e items in p blocks
    best: List[List[Tuple[float, float]]] = [[INF] * (n + 1) for _ in range(partitions + 1)]
    cut: List[List[int]] = [[-1] * (n + 1) for _ in range(partitions + 1)]
    best[0][0] = (0.0, 0.0)
    for p in range(1, partitions + 1):
        for e in range(p, n - (partitions - p) + 1):
            for s in range(p - 1, e):
                prev = best[p - 1][s]
                if prev == INF:
                    continue
                w = weight(s, e)
                cand = (max(prev[0], w), prev[1] + w * w)
                if cand < best[p][e]:
                    best[p][e] = cand
                    cut[p][e] = s

    bounds = [n]
    e = n
    for p in range(partitions, 0, -1):
        e = cut[p][e]
        bounds.append(e)
    bounds.reverse()
    return [list(sequence[a:b]) for a, b in zip(bounds, bounds[1:])]
