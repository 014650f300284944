"""State-dict key translation between ``Pipe`` and the wrapped ``nn.Sequential``.

``Pipe`` registers only ``partitions`` (an ``nn.ModuleList`` of flattened
stages), so its keys read ``partitions.<stage>.<index in stage>.<param path>``
(reference ``pipe.py:181-188,216,344``).  With the usual "one
nested ``nn.Sequential`` per device" idiom those are the wrapped module's own
keys plus the ``partitions.`` prefix; with a flat ``nn.Sequential`` (several
children per device, CPU children splitting stages) they differ.  The helpers
below translate either way through the identity of the parameter/buffer
tensors, so a checkpoint written from the plain model loads into the pipe and
vice versa.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Mapping

import torch
from torch import nn

__all__ = ["key_map", "to_sequential_state_dict", "from_sequential_state_dict"]


def _flat_layers(module: nn.Sequential):
    """(sequential-key-prefix, layer) in execution order, opening nested
    Sequentials one level the way stage assembly does."""
    from .partition import WithDevice
    for name, child in module.named_children():
        if isinstance(child, WithDevice):
            yield f"{name}._module", child.module
        elif isinstance(child, nn.Sequential):
            for sub_name, sub in child.named_children():
                yield f"{name}.{sub_name}", sub
        else:
            yield name, child


def key_map(pipe, module: nn.Sequential) -> "OrderedDict[str, str]":
    """Map each key of ``module.state_dict()`` to its key in ``pipe.state_dict()``.

    ``module`` must be the Sequential the pipe was built from (the layer objects
    are matched by identity)."""
    where: Dict[int, str] = {}
    for j, stage in enumerate(pipe.partitions):
        for k, layer in enumerate(stage):
            where[id(layer)] = f"partitions.{j}.{k}"

    mapping: "OrderedDict[str, str]" = OrderedDict()
    for prefix, layer in _flat_layers(module):
        target = where.get(id(layer))
        if target is None:
            raise KeyError(f"layer '{prefix}' of the given module is not part of this Pipe")
        for key in layer.state_dict().keys():
            mapping[f"{prefix}.{key}"] = f"{target}.{key}"
    return mapping


def to_sequential_state_dict(pipe, module: nn.Sequential) -> "OrderedDict[str, torch.Tensor]":
    """``pipe.state_dict()`` re-keyed like ``module.state_dict()``."""
    sd = pipe.state_dict()
    return OrderedDict((seq_key, sd[pipe_key]) for seq_key, pipe_key in key_map(pipe, module).items())


def from_sequential_state_dict(pipe, module: nn.Sequential,
                               state_dict: Mapping[str, torch.Tensor]) -> "OrderedDict[str, torch.Tensor]":
    """A state dict keyed like ``module`` re-keyed for ``pipe.load_state_dict``."""
    mapping = key_map(pipe, module)
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for key, value in state_dict.items():
        if key not in mapping:
            raise KeyError(f"unexpected key '{key}' for the wrapped Sequential")
        out[mapping[key]] = value
    return out
