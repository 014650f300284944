"""``Pipe``: wrap an ``nn.Sequential`` for synchronous pipeline parallelism.

    model = nn.Sequential(stage0_layers.cuda(0), stage1_layers.cuda(1))
    model = Pipe(model, chunks=8, checkpoint="except_last")
    out = model(tokens).local_value()

The wrapped module is cut into stages (by the devices its layers already live
on, or by an explicit ``balance=`` / ``devices=``), each mini-batch is cut into
``chunks`` micro-batches, and the ``(m + n - 1)``-tick fill-drain schedule runs
the stages concurrently on their devices.  Activations cross stage boundaries as
asynchronous peer copies on dedicated per-(stage, micro-batch) copy streams —
over NVLink 5 / NVSwitch on a B200 box every stage is one hop from every other.

Parity: reference ``pipe.py:224-494``; SURVEY.md C1, C5 and Appendix B.
"""
from __future__ import annotations

from typing import Any, Iterable, Iterator, List, Optional, Union

import torch
from torch import Tensor, nn

from . import microbatch
from .autograd.stream import AbstractStream, new_stream
from .batchnorm import DeferredBatchNorm
from .partition import (BalanceError, PipeSequential, WithDevice, split_by_balance,
                        split_by_device, verify_module, verify_splitting)
from .rref import make_rref
from .runtime.executor import Pipeline
from .runtime.timeline import Timeline
from .skip.layout import inspect_skip_layout
from .skip.skippable import verify_skippables

__all__ = ["Pipe", "BalanceError", "PipeSequential", "WithDevice", "MOVING_DENIED"]

Device = Union[torch.device, int, str]

MOVING_DENIED = TypeError("denied to move parameters and buffers, because Pipe should manage device placement")

_CHECKPOINT_MODES = ("always", "except_last", "never")


class Pipe(nn.Module):
    """Synchronous (GPipe-style) pipeline parallelism over an ``nn.Sequential``.

    Args:
        module: the model; an ``nn.Sequential`` whose children are already on
            their devices (use :class:`WithDevice` for parameter-less layers),
            unless ``balance`` is given.
        chunks: number of micro-batches per mini-batch (default 1).
        checkpoint: ``'always'``, ``'except_last'`` (default) or ``'never'`` —
            which micro-batches recompute their activations in backward.
        deferred_batch_norm: update BatchNorm running statistics once per
            mini-batch instead of once per micro-batch.
        balance: optional layer counts per stage (torchgpipe style); layers are
            then *moved* to ``devices`` (default: ``cuda:0..``, or CPU stages).
        devices: devices for ``balance``.
        strict_except_last: the reference computes "all but the last
            micro-batch" from the *requested* ``chunks`` (``pipe.py:354``), so when
            the batch yields fewer micro-batches (20 rows / 8 chunks -> 7)
            ``except_last`` silently checkpoints every one of them
            (``README.md:398``).  ``False`` (default) keeps that behaviour;
            ``True`` exempts the actual last micro-batch.

    Raises:
        TypeError: ``module`` is not an ``nn.Sequential``, or skippables mismatch.
        ValueError: bad ``chunks`` / ``checkpoint``; duplicate children; parameters
            shared across devices.
        BalanceError / IndexError: ``balance`` does not fit the module / devices.
    """

    def __init__(self, module: nn.Sequential, chunks: int = 1, checkpoint: str = "except_last",
                 deferred_batch_norm: bool = False, *, balance: Optional[Iterable[int]] = None,
                 devices: Optional[Iterable[Device]] = None, strict_except_last: bool = False) -> None:
        super().__init__()

        chunks = int(chunks)
        checkpoint = str(checkpoint)
        if chunks <= 0:
            raise ValueError("number of chunks must be positive integer")
        if checkpoint not in _CHECKPOINT_MODES:
            raise ValueError("checkpoint is not one of 'always', 'except_last', or 'never'")

        verify_module(module)
        # Stash/pop pairing is static, so it is checked before anything runs.
        verify_skippables(module)

        self.chunks = chunks
        self.checkpoint = checkpoint
        self.strict_except_last = bool(strict_except_last)

        if deferred_batch_norm:
            module = DeferredBatchNorm.convert_deferred_batch_norm(module, chunks)

        if balance is not None:
            self.partitions, self.devices = split_by_balance(module, balance, devices)
        else:
            if devices is not None:
                raise ValueError("'devices' is only meaningful together with 'balance'")
            self.partitions, self.devices = split_by_device(module)
        verify_splitting(module, self.partitions, self.devices)

        self._copy_streams: List[List[AbstractStream]] = []
        self._skip_layout = inspect_skip_layout(self.partitions)
        copy_streams = self._ensure_copy_streams()

        self.pipeline = Pipeline(self.partitions, self.devices, copy_streams, self._skip_layout,
                                 self._checkpoint_stop(chunks))

    # -- configuration helpers -----------------------------------------------
    def _checkpoint_stop(self, m: int) -> int:
        """Index of the first micro-batch that is *not* checkpointed, for ``m``
        micro-batches."""
        return {"always": m, "except_last": m - 1, "never": 0}[self.checkpoint]

    def _ensure_copy_streams(self) -> List[List[AbstractStream]]:
        """One copy stream per (stage, micro-batch slot); created lazily once.
        The grid is sized by the *requested* chunks even when a batch yields
        fewer micro-batches."""
        if not self._copy_streams:
            for device in self.devices:
                self._copy_streams.append([new_stream(device) for _ in range(self.chunks)])
        return self._copy_streams

    @property
    def timeline(self) -> Optional[Timeline]:
        return self.pipeline.timeline

    @timeline.setter
    def timeline(self, value: Optional[Timeline]) -> None:
        self.pipeline.timeline = value

    def trace_cells(self, enabled: bool = True) -> None:
        """Label every cell ``chunk{i}-part{j}`` for torch.profiler (the hook the
        reference's author removed, ``pipeline.py:206-209``)."""
        self.pipeline.trace_cells = enabled

    def close(self) -> None:
        """Stop the worker threads of this pipe."""
        self.pipeline.close()

    # -- container protocol ----------------------------------------------------
    def __len__(self) -> int:
        return sum(len(p) for p in self.partitions)

    def __getitem__(self, index: int) -> nn.Module:
        if index < 0:
            index += len(self)
        if index < 0:
            raise IndexError
        for stage in self.partitions:
            if index < len(stage):
                return stage[index]
            index -= len(stage)
        raise IndexError

    def __iter__(self) -> Iterator[nn.Module]:
        for stage in self.partitions:
            yield from stage

    # -- device movement is the pipe's business ----------------------------------
    def cuda(self, device: Optional[Device] = None) -> "Pipe":
        raise MOVING_DENIED

    def cpu(self) -> "Pipe":
        raise MOVING_DENIED

    def to(self, *args: Any, **kwargs: Any) -> "Pipe":
        # Anything that names a device (or a tensor to imitate) is refused;
        # dtype-only conversions pass through.
        if "device" in kwargs or "tensor" in kwargs:
            raise MOVING_DENIED
        if args and isinstance(args[0], (torch.device, int, str, Tensor)):
            raise MOVING_DENIED
        return super().to(*args, **kwargs)

    # -- forward -----------------------------------------------------------------
    def forward(self, *inputs: Any):
        """Run one mini-batch through the pipeline.

        Every tensor input is split along dim 0 into micro-batches (wrap one in
        :class:`~pipe_b200.NoChunk` to hand it whole to each); non-tensors are
        replicated.  At least one input must be a tensor, and tensors must be on
        the first stage's device.  Returns an RRef to the gathered output (a
        :class:`~pipe_b200.rref.LocalRRef` when RPC is not initialised).
        """
        first_device = self.devices[0] if self.devices else torch.device("cpu")
        microbatch.check(first_device, *inputs)

        if not self.devices:
            # Empty Sequential: identity.
            return make_rref(inputs[0] if len(inputs) == 1 else inputs)

        batches = microbatch.scatter(*inputs, chunks=self.chunks)
        if self.strict_except_last:
            self.pipeline.checkpoint_stop = self._checkpoint_stop(len(batches))

        self.pipeline.run(batches)

        return make_rref(microbatch.gather(batches))
