"""BatchNorm whose running statistics see the mini-batch, not the micro-batch.

Splitting a mini-batch into ``chunks`` micro-batches would make a stock
``BatchNorm`` update its running mean/var ``chunks`` times per step, each time
from a fraction of the data.  :class:`DeferredBatchNorm` accumulates ``sum`` and
``sum of squares`` over the micro-batches and commits one update when the last
one has passed; normalisation itself always uses the micro-batch's own
statistics (training) or the running ones (eval).  Nothing is tracked while a
checkpointed cell is being recomputed, so a step is never counted twice.

Enabled with ``Pipe(..., deferred_batch_norm=True)``.

Parity: reference ``pipe.py:18,261-265,313,341-342``, ``README.md:547-554`` (the
module itself is absent); SURVEY.md A6 and Appendix A "batchnorm".
"""
from __future__ import annotations

from typing import Optional, TypeVar, cast

import torch
import torch.nn.functional as F
from torch import Tensor, nn
from torch.nn.modules.batchnorm import _BatchNorm

from .autograd.checkpoint import is_recomputing

__all__ = ["DeferredBatchNorm"]

TModule = TypeVar("TModule", bound=nn.Module)


class DeferredBatchNorm(_BatchNorm):
    sum: Tensor
    sum_squares: Tensor
    running_mean: Tensor
    running_var: Tensor
    num_batches_tracked: Tensor

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: Optional[float] = 0.1,
                 affine: bool = True, chunks: int = 1) -> None:
        super().__init__(num_features, eps, momentum, affine, track_running_stats=True)
        self.register_buffer("sum", torch.zeros_like(self.running_mean))
        self.register_buffer("sum_squares", torch.zeros_like(self.running_var))
        self.counter = 0    # elements per channel accumulated so far
        self.tracked = 0    # micro-batches accumulated so far
        self.chunks = chunks

    def _check_input_dim(self, input: Tensor) -> None:
        if input.dim() <= 2:
            raise ValueError("expected at least 3D input (got %dD input)" % input.dim())

    def _track(self, input: Tensor) -> bool:
        """Fold one micro-batch into the accumulators; True when the mini-batch
        is complete."""
        reduce_dims = [0] + list(range(2, input.dim()))
        with torch.no_grad():
            self.sum += input.sum(reduce_dims)
            self.sum_squares += (input * input).sum(reduce_dims)
        self.counter += input.numel() // input.size(1)
        self.tracked += 1
        return self.tracked == self.chunks

    def _commit(self) -> None:
        """Fold the accumulated mini-batch statistics into the running ones."""
        self.num_batches_tracked += 1
        if self.momentum is None:
            factor = 1.0 / float(self.num_batches_tracked)
        else:
            factor = self.momentum

        mean = self.sum / self.counter
        var = self.sum_squares / self.counter - mean * mean
        self.running_mean += factor * (mean - self.running_mean)
        self.running_var += factor * (var - self.running_var)

        self.sum.zero_()
        self.sum_squares.zero_()
        self.counter = 0
        self.tracked = 0

    def forward(self, input: Tensor) -> Tensor:  # type: ignore[override]
        self._check_input_dim(input)
        if not self.training:
            return F.batch_norm(input, self.running_mean, self.running_var, self.weight, self.bias,
                                training=False, momentum=0.0, eps=self.eps)

        if not is_recomputing():
            if self._track(input):
                self._commit()

        # Normalise with this micro-batch's statistics; running stats untouched.
        return F.batch_norm(input, None, None, self.weight, self.bias,
                            training=True, momentum=0.0, eps=self.eps)

    @classmethod
    def convert_deferred_batch_norm(cls, module: TModule, chunks: int = 1) -> TModule:
        """Recursively replace every ``_BatchNorm`` in ``module`` by a
        :class:`DeferredBatchNorm` carrying the same parameters and buffers."""
        if isinstance(module, DeferredBatchNorm) and module.chunks == chunks:
            return cast(TModule, module)

        out: nn.Module = module
        if isinstance(module, _BatchNorm) and module.track_running_stats:
            bn = DeferredBatchNorm(module.num_features, module.eps, module.momentum, module.affine, chunks)
            if module.affine:
                bn.register_parameter("weight", module.weight)
                bn.register_parameter("bias", module.bias)
            bn.register_buffer("running_mean", module.running_mean)
            bn.register_buffer("running_var", module.running_var)
            bn.register_buffer("num_batches_tracked", module.num_batches_tracked)
            bn.sum = bn.sum.to(module.running_mean.device)
            bn.sum_squares = bn.sum_squares.to(module.running_var.device)
            out = bn

        for name, child in module.named_children():
            out.add_module(name, cls.convert_deferred_batch_norm(child, chunks))
        return cast(TModule, out)
