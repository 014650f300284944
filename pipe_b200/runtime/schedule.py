"""The fill-drain (GPipe) clock schedule.

With ``m`` micro-batches and ``n`` stages there are ``m + n - 1`` clock ticks;
tick ``k`` runs every cell ``(i, j)`` with ``i + j == k``.  Cells of one tick
touch distinct stages, so they run concurrently on distinct devices.

              tick:   0      1      2      3      4
    m=3, n=3        (0,0)  (1,0)  (2,0)
                           (0,1)  (1,1)  (2,1)
                                  (0,2)  (1,2)  (2,2)

Parity: reference ``pipeline.py:63-79``; worked table ``README.md:72-84``.
"""
from __future__ import annotations

from typing import Iterator, List, Tuple

__all__ = ["clock_cycles", "bubble_fraction", "Cell"]

Cell = Tuple[int, int]  # (micro-batch index, stage index)


def clock_cycles(m: int, n: int) -> Iterator[List[Cell]]:
    """Yield, tick by tick, the list of ``(micro_batch, stage)`` cells to run."""
    for tick in range(m + n - 1):
        first_stage = max(tick - m + 1, 0)
        last_stage = min(tick, n - 1)
        yield [(tick - j, j) for j in range(first_stage, last_stage + 1)]


def bubble_fraction(m: int, n: int) -> float:
    """Idle share of the ideal fill-drain schedule with equal-cost cells:
    ``(n - 1) / (m + n - 1)``.  Holds for forward and backward alike."""
    if m <= 0 or n <= 0:
        return 0.0
    return (n - 1) / (m + n - 1)
