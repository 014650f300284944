"""Automatic balancing: how many layers each stage should take.

>>> balance = balance_by_time(torch.cuda.device_count(), model, sample)
>>> pipe = Pipe(model, balance=balance, chunks=8)

Parity: advertised by the reference at ``pipe.py:42-58`` (the package itself is
absent); SURVEY.md A11.
"""
from __future__ import annotations

from typing import Any, List, Sequence, Union

import torch
from torch import nn

from . import blockpartition
from .profile import profile_sizes, profile_times

__all__ = ["balance_by_time", "balance_by_size", "balance_cost"]

Device = Union[torch.device, int, str]


def balance_cost(cost: Sequence[float], partitions: int) -> List[int]:
    """Layer counts per stage from per-layer costs."""
    return [len(block) for block in blockpartition.solve(cost, partitions)]


def _default_device() -> torch.device:
    return torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")


def balance_by_time(partitions: int, module: nn.Sequential, sample: Any, *,
                    timeout: float = 1.0, device: Device = None) -> List[int]:
    """Balance by measured forward+backward time per layer."""
    dev = torch.device(device) if device is not None else _default_device()
    return balance_cost(profile_times(module, sample, timeout, dev), partitions)


def balance_by_size(partitions: int, module: nn.Sequential, input: Any, *, chunks: int = 1,
                    param_scale: float = 2.0, device: Device = None) -> List[int]:
    """Balance by CUDA memory per layer.  ``param_scale`` accounts for what the
    optimizer keeps per parameter (SGD 2-3, Adam 4-5, ...)."""
    dev = torch.device(device) if device is not None else torch.device("cuda")
    return balance_cost(profile_sizes(module, input, chunks, param_scale, dev), partitions)
