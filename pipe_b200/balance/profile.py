"""Per-layer cost probes that feed the balancers.

``profile_times`` measures forward+backward time of each layer on a sample,
``profile_sizes`` the CUDA memory each layer pins (activations for the
micro-batch plus ``param_scale`` x parameter bytes).  On CUDA the timing uses
events on the device, not wall clock.

Parity: SURVEY.md A11 (upstream ``_balance/profile.py``; absent from the
reference, advertised at ``pipe.py:42-58``).
"""
from __future__ import annotations

import copy
import time
from typing import Any, Generator, List, Union

import torch
from torch import nn

from ..microbatch import Batch

__all__ = ["profile_times", "profile_sizes"]

Device = Union[torch.device, int, str]


def _layerwise_sandbox(module: nn.Sequential, device: torch.device) -> Generator[nn.Module, None, None]:
    """Yield a training-mode deep copy of each layer on ``device``, one at a time,
    so profiling never touches the user's weights or BN statistics."""
    for layer in module:
        clone = copy.deepcopy(layer)
        clone.to(device)
        clone.train()
        yield clone


def _detach(batch: Batch) -> None:
    for k, x in enumerate(batch):
        if torch.is_tensor(x):
            batch[k] = x.detach().requires_grad_(x.requires_grad)


def _sync(device: torch.device) -> None:
    if device.type == "cuda":
        torch.cuda.synchronize(device)


def profile_times(module: nn.Sequential, sample: Any, timeout: float, device: torch.device) -> List[int]:
    """Accumulated forward+backward microseconds per layer, sampled repeatedly
    until ``timeout`` seconds have passed (at least once)."""
    if any(p.grad is not None for p in module.parameters()):
        raise ValueError("some parameter already has gradient")

    batch0 = Batch(sample)
    for k, x in enumerate(batch0):
        if torch.is_tensor(x):
            batch0[k] = x.detach().to(device).requires_grad_(x.requires_grad)

    totals: List[float] = [0.0] * len(module)
    warm = True     # first pass pays one-off costs (engine threads, cuBLAS handles): not counted
    begun = time.time()
    while True:
        batch = batch0
        for idx, layer in enumerate(_layerwise_sandbox(module, device)):
            _detach(batch)
            if device.type == "cuda":
                t0 = torch.cuda.Event(enable_timing=True)
                t1 = torch.cuda.Event(enable_timing=True)
                t0.record(torch.cuda.current_stream(device))
            else:
                w0 = time.perf_counter()

            batch = batch.call(layer)
            outs = tuple(y for y in batch if torch.is_tensor(y) and y.requires_grad)
            if outs:
                torch.autograd.backward(outs, [torch.ones_like(y) for y in outs])

            if device.type == "cuda":
                t1.record(torch.cuda.current_stream(device))
                _sync(device)
                spent = t0.elapsed_time(t1) * 1e3
            else:
                spent = (time.perf_counter() - w0) * 1e6
            if not warm:
                totals[idx] += spent
        if warm:
            warm = False
            begun = time.time()
            continue
        if time.time() - begun >= timeout:
            break
    return [int(t) for t in totals]


def profile_sizes(module: nn.Sequential, input: Any, chunks: int, param_scale: float,
                  device: torch.device) -> List[int]:
    """Bytes each layer occupies on ``device`` for one micro-batch: latent
    activation memory + ``param_scale`` x parameter bytes.  CUDA only."""
    if device.type != "cuda":
        raise ValueError("size profiler supports only CUDA device")

    batch = Batch(input)
    sizes: List[int] = []
    latent_scale = batch[batch.find_tensor_idx()].size(0) / chunks
    for k, x in enumerate(batch):
        if torch.is_tensor(x):
            batch[k] = x[:1].detach().to(device).requires_grad_(x.requires_grad)

    for layer in _layerwise_sandbox(module, device):
        _detach(batch)
        before = torch.cuda.memory_allocated(device)
        batch = batch.call(layer)
        after = torch.cuda.memory_allocated(device)
        latent = after - before

        param_bytes = sum(p.numel() * p.element_size() for p in layer.parameters())
        sizes.append(int(latent * latent_scale + param_bytes * param_scale))
    return sizes
