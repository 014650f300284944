"""Turning an ``nn.Sequential`` into pipeline stages.

Two ways are offered, and both end in the same shape — an
``nn.ModuleList`` of :class:`PipeSequential` plus a parallel list of devices:

* :func:`split_by_device` — *device-inferred*: every child already lives on a
  device (its parameters', or the one named by :class:`WithDevice`, or CPU when
  it has no parameters); a new stage starts whenever the device changes, and
  every CPU child is a stage of its own.  Children that are themselves
  ``nn.Sequential`` are flattened into their stage.  This is what the reference
  ships (``pipe.py:94-118,181-218``).
* :func:`split_by_balance` — torchgpipe-style ``balance=[3, 2, ...]`` with
  ``devices=[...]``: the first ``balance[0]`` children go to ``devices[0]`` and
  so on; layers are *moved* to their device.  The reference keeps only vestiges
  of it (``pipe.py:42-58,90-91``) but BASELINE.json lists it as preserved surface.

Parity: reference ``pipe.py:61-218``; SURVEY.md C2-C4, C6.
"""
from __future__ import annotations

from typing import Any, Iterable, List, Optional, Sequence, Tuple, Union

import torch
from torch import nn

__all__ = ["PipeSequential", "WithDevice", "BalanceError", "split_by_device", "split_by_balance",
           "verify_module", "verify_splitting", "retrieve_device", "recommend_auto_balance"]

Device = Union[torch.device, int, str]


class BalanceError(ValueError):
    """The requested ``balance`` does not fit the module or the devices."""


def recommend_auto_balance(message: str) -> str:
    """Append a pointer to the automatic balancers to an error message."""
    return (f"{message}\n\n"
            "If the model is still changing, its best balance changes with it.  The automatic\n"
            "balancers give a reasonable starting point:\n\n"
            "  from pipe_b200 import Pipe\n"
            "  from pipe_b200.balance import balance_by_time\n\n"
            "  stages = torch.cuda.device_count()\n"
            "  sample = torch.empty(...)\n"
            "  balance = balance_by_time(stages, model, sample)\n\n"
            "  model = Pipe(model, balance=balance, chunks=...)\n")


class PipeSequential(nn.Sequential):
    """``nn.Sequential`` for stages: a tuple returned by one layer is splatted
    into the next layer's positional arguments; anything else (a tensor, a list)
    is passed as one argument."""

    def forward(self, *inputs: Any):  # type: ignore[override]
        for step in self._plan():
            if type(step) is tuple:
                inputs = _run_chain(step, inputs)
            else:
                inputs = step(*inputs) if isinstance(inputs, tuple) else step(inputs)
        return inputs

    def _plan(self):
        """The stage's layers, with every run of two or more consecutive *chainable* layers (a class
        that offers ``pb2_make_chain``: the fused encoder layer) replaced by ``(chain, layers)`` so the
        run executes as one native call per pass.  Rebuilt when the children change."""
        mods = tuple(self._modules.values())
        cached = self.__dict__.get("_pb2_plan")
        if cached is not None and cached[0] == tuple(map(id, mods)):
            return cached[1]
        plan: List[Any] = []
        run: List[nn.Module] = []

        def close() -> None:
            chain = None
            if len(run) >= 2:
                try:
                    chain = type(run[0]).pb2_make_chain(list(run))
                except ValueError:
                    chain = None               # layers of different widths: run them one by one
            if chain is not None:
                plan.append((chain, tuple(run)))
            else:
                plan.extend(run)
            run.clear()

        for m in mods:
            if hasattr(type(m), "pb2_make_chain") and (not run or type(run[0]) is type(m)):
                run.append(m)
            else:
                close()
                if hasattr(type(m), "pb2_make_chain"):
                    run.append(m)
                else:
                    plan.append(m)
        close()
        self.__dict__["_pb2_plan"] = (tuple(map(id, mods)), plan)
        return plan


def _run_chain(step, inputs):
    chain, layers = step
    x, extra = (inputs[0], inputs[1:]) if isinstance(inputs, tuple) else (inputs, ())
    if torch.is_tensor(x) and chain.usable(x):
        y = chain(x)
        return (y, *extra) if extra else y
    for layer in layers:
        inputs = layer(*inputs) if isinstance(inputs, tuple) else layer(inputs)
    return inputs


class WithDevice(nn.Module):
    """Pins a layer to a device for device-inferred partitioning.

    Needed for layers without parameters (Dropout, ReLU, reshapes), which would
    otherwise be taken for CPU layers and become CPU stages.  ``Pipe`` unwraps
    the marker and moves the layer to ``device``.

    >>> Pipe(nn.Sequential(fc1, WithDevice(nn.Dropout(0.1), 'cuda:1'), fc2))
    """

    def __init__(self, module: nn.Module, device: Device):
        super().__init__()
        self._module = module
        self._device = torch.device(device)

    def forward(self, *args: Any, **kwargs: Any):
        return self._module(*args, **kwargs)

    @property
    def module(self) -> nn.Module:
        return self._module

    @property
    def device(self) -> torch.device:
        return self._device


def verify_module(module: nn.Sequential) -> None:
    if not isinstance(module, nn.Sequential):
        raise TypeError("module must be nn.Sequential to be partitioned")
    if len(list(module.named_children())) != len(module):
        raise ValueError("module with duplicate children is not supported")


def verify_splitting(module: nn.Sequential, partitions: Sequence[nn.Module],
                     devices: Sequence[torch.device]) -> None:
    """Reject parameters shared between stages on *different* devices (sharing
    within one device is fine)."""
    n_all = len(list(module.parameters()))
    n_children = sum(len(list(c.parameters())) for c in module.children())
    if n_all == n_children:
        return  # nothing is shared anywhere

    owner = {}
    for j, part in enumerate(partitions):
        for p in part.parameters():
            prev = owner.setdefault(id(p), j)
            if prev != j and devices[prev] != devices[j]:
                raise ValueError("module with duplicate parameters on distinct devices is not supported")


def retrieve_device(module: nn.Module) -> torch.device:
    """The single device of ``module``'s parameters (CPU if it has none)."""
    found: Optional[torch.device] = None
    for p in module.parameters():
        if found is None:
            found = p.device
        elif found != p.device:
            raise ValueError(f"nn.Module: {module}, should have all parameters on a single device,"
                             " please use .to() to place the module on a single device")
    return found if found is not None else torch.device("cpu")


def _assemble(layers: List[nn.Module]) -> PipeSequential:
    """Build one stage.  A child that is itself an ``nn.Sequential`` is opened up
    ONE level (deeper nesting stays nested), and children are renumbered from 0 —
    which is what fixes the ``partitions.<j>.<k>.*`` state-dict keys."""
    flat: List[nn.Module] = []
    for layer in layers:
        if isinstance(layer, nn.Sequential):
            flat.extend(layer.children())
        else:
            flat.append(layer)
    return PipeSequential(*flat)


def split_by_device(module: nn.Sequential) -> Tuple[nn.ModuleList, List[torch.device]]:
    partitions: List[PipeSequential] = []
    devices: List[torch.device] = []

    pending: List[nn.Module] = []
    pending_device: Optional[torch.device] = None

    for _name, child in module.named_children():
        if isinstance(child, WithDevice):
            device = child.device
            child = child.module.to(device)
        else:
            device = retrieve_device(child)

        # Cut on a device change; CPU layers never share a stage.
        if pending and (pending_device != device or device.type == "cpu"):
            partitions.append(_assemble(pending))
            devices.append(pending_device)  # type: ignore[arg-type]
            pending = []
        pending.append(child)
        pending_device = device

    if pending:
        partitions.append(_assemble(pending))
        devices.append(pending_device)  # type: ignore[arg-type]

    return nn.ModuleList(partitions), devices


def split_by_balance(module: nn.Sequential, balance: Iterable[int],
                     devices: Optional[Iterable[Device]] = None) -> Tuple[nn.ModuleList, List[torch.device]]:
    balance = [int(b) for b in balance]
    if len(module) != sum(balance):
        raise BalanceError(recommend_auto_balance(
            f"module and sum of balance have different length "
            f"(module: {len(module)}, sum of balance: {sum(balance)})"))
    if any(b <= 0 for b in balance):
        raise BalanceError(recommend_auto_balance(f"all balance numbers must be positive integer (balance: {balance})"))

    if devices is None:
        if torch.cuda.is_available():
            devices = range(torch.cuda.device_count())
        else:
            devices = ["cpu"] * len(balance)
    devs = [torch.device("cuda", d) if isinstance(d, int) else torch.device(d) for d in devices]
    if len(balance) > len(devs):
        raise IndexError(f"too few devices to hold given partitions (devices: {len(devs)}, partitions: {len(balance)})")
    devs = devs[:len(balance)]

    children = [c.module if isinstance(c, WithDevice) else c for c in module.children()]
    partitions: List[PipeSequential] = []
    at = 0
    for b, device in zip(balance, devs):
        stage = _assemble(children[at:at + b])
        stage.to(device)
        partitions.append(stage)
        at += b
    return nn.ModuleList(partitions), devs
