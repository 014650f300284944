"""Uniform stream layer over CUDA streams and a CPU stand-in.

Everything above this module (Copy/Wait, the scheduler, checkpointing, portals)
talks to "a stream on a device" without caring whether that device is a B200 or
the host.  On the host the stand-in :data:`CPUStream` makes every operation a
no-op (or a plain synchronise), which is what lets the whole scheduler run on a
CPU-only machine and is how the CPU test tier exercises multi-stage pipelines.

Parity: the reference imports this layer as ``.stream`` (``pipe.py:22``,
``pipeline.py:22``); the only excerpt it ships is ``wait_stream``
(``README.md:349-357``).  Behavioural contract: SURVEY.md Appendix A "stream".
"""
from __future__ import annotations

from contextlib import contextmanager
from typing import Generator, Union

import torch

__all__ = [
    "CPUStreamType", "CPUStream", "AbstractStream", "is_cuda", "new_stream",
    "current_stream", "default_stream", "use_device", "use_stream",
    "get_device", "wait_stream", "record_stream", "as_cuda",
]


class CPUStreamType:
    """Singleton type of the host-side stream stand-in."""

    _instance = None

    def __new__(cls):
        if cls._instance is None:
            cls._instance = super().__new__(cls)
        return cls._instance

    def __repr__(self) -> str:
        return "CPUStream"


CPUStream = CPUStreamType()

AbstractStream = Union[torch.cuda.Stream, CPUStreamType]


def is_cuda(stream: AbstractStream) -> bool:
    return stream is not CPUStream


def as_cuda(stream: AbstractStream) -> torch.cuda.Stream:
    if stream is CPUStream:
        raise TypeError("CPUStream cannot be used as a CUDA stream")
    return stream  # type: ignore[return-value]


def new_stream(device: torch.device) -> AbstractStream:
    """A fresh side stream on ``device`` (the sentinel on non-CUDA devices)."""
    if device.type != "cuda":
        return CPUStream
    return torch.cuda.Stream(device)


def current_stream(device: torch.device) -> AbstractStream:
    if device.type != "cuda":
        return CPUStream
    return torch.cuda.current_stream(device)


def default_stream(device: torch.device) -> AbstractStream:
    if device.type != "cuda":
        return CPUStream
    return torch.cuda.default_stream(device)


@contextmanager
def use_device(device: torch.device) -> Generator[None, None, None]:
    """Make ``device`` current for the duration of the block (no-op on CPU)."""
    if device.type != "cuda":
        yield
        return
    with torch.cuda.device(device):
        yield


@contextmanager
def use_stream(stream: AbstractStream) -> Generator[None, None, None]:
    """Make ``stream`` (and its device) current; no-op for the CPU sentinel."""
    if stream is CPUStream:
        yield
        return
    with torch.cuda.stream(as_cuda(stream)):
        yield


def get_device(stream: AbstractStream) -> torch.device:
    if stream is CPUStream:
        return torch.device("cpu")
    return as_cuda(stream).device


def wait_stream(source: AbstractStream, target: AbstractStream) -> None:
    """Make work queued on ``source`` *after* this call wait for the work queued
    on ``target`` *before* this call ("source waits for target").

    CUDA/CUDA is an event edge between the streams.  A host source has to block
    until the device target drains.  A host target needs nothing: host work is
    already complete when it returns.
    """
    if target is CPUStream:
        return
    if source is CPUStream:
        as_cuda(target).synchronize()
    else:
        as_cuda(source).wait_stream(as_cuda(target))


def record_stream(tensor: torch.Tensor, stream: AbstractStream) -> None:
    """Tell the caching allocator that ``tensor``'s storage is in use on
    ``stream`` so it is not recycled while a cross-stream copy is in flight.

    ``Tensor.record_stream`` keys on the data pointer of the tensor it is given;
    a view that starts part-way into the storage would record the wrong block
    start, so we record on a zero-offset alias of the whole storage instead.
    """
    if stream is CPUStream or not tensor.is_cuda:
        return
    storage = tensor.untyped_storage()
    if storage.size() == 0:
        return
    alias = torch.empty(0, dtype=torch.uint8, device=tensor.device).set_(storage)
    alias.record_stream(as_cuda(stream))
