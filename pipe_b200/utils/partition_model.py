"""Re-nest a flat ``nn.Sequential`` into one ``nn.Sequential`` per device.

The result is the "one nested Sequential per device" shape that device-inferred
``Pipe`` splitting expects, with each group already moved to its device.

Parity: SURVEY.md A11 (upstream ``utils.partition_model``; absent from the
reference tree).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Union

import torch
from torch import nn

__all__ = ["partition_model"]


def partition_model(module: nn.Sequential, balance: Sequence[int],
                    devices: Optional[Sequence[Union[int, str, torch.device]]] = None) -> nn.Sequential:
    """Group ``module``'s children by ``balance`` and place group ``k`` on
    ``devices[k]`` (default ``cuda:k``; ``cpu`` when CUDA is unavailable)."""
    layers = list(module.children())
    if sum(balance) != len(layers):
        raise ValueError(f"sum of balance ({sum(balance)}) must equal the number of layers ({len(layers)})")
    if devices is None:
        if torch.cuda.is_available():
            devices = list(range(len(balance)))
        else:
            devices = ["cpu"] * len(balance)
    if len(devices) < len(balance):
        raise IndexError(f"too few devices ({len(devices)}) for {len(balance)} partitions")

    groups: List[nn.Module] = []
    at = 0
    for count, device in zip(balance, devices):
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        group = nn.Sequential(*layers[at:at + count])
        group.to(dev)
        groups.append(group)
        at += count
    return nn.Sequential(*groups)
