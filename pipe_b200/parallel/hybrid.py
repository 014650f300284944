"""Pipe-inside-DDP hybrid: ``k`` pipeline replicas x ``n`` stages on ``k * n`` GPUs.

Rank layout: ``rank = replica * n + stage``.  The ``n`` ranks of a replica form a pipeline group
(a :class:`~pipe_b200.parallel.DistPipe`); the ``k`` ranks that hold the same stage form a replica
group whose gradients are averaged between backward and the optimizer step.  Each replica is fed
its own shard of the global batch.

The reference only notes that DDP *may* wrap ``Pipe`` and only with ``checkpoint='never'``
(``pipe.py:290-293``): DDP's per-parameter hooks fire inside ``Checkpoint.backward``'s nested
autograd pass.  Here the reduction is not hook-driven.  Every gradient of a stage lives in one
autograd pass.  Here the reduction is not hook-driven, so all three checkpoint modes work.

Default (``reduce_backend="gather"``): the weight matrices are never all-reduced at all.  The transpose kernels
that stack the wgrad operands of every micro-batch during backward store them into every replica's stash over
NVLink; the one deferred wgrad GEMM per Linear then runs over all replicas' tokens and averages in its fp32
accumulator (``parallel/replica_stash.py``).  Small gradients use the all-reduce kernel on a side stream, under
the wgrad GEMMs.  ``"allreduce"`` keeps dense gradients and runs the NVLink all-reduce kernel
(``csrc/allreduce.cu``) per weight-matrix bucket as each wgrad GEMM is enqueued; ``"nccl"`` does the same with
asynchronous ``dist.all_reduce`` calls — the DDP-style baseline to beat.  Later stages finish their backward
first, so their reduction overlaps the backward of the earlier stages.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
from torch import nn

from .dist_pipe import DistPipe

__all__ = ["HybridPipe", "hybrid_groups"]


def hybrid_groups(stages: int):
    """Create (collectively) every pipeline group and every replica group; return this rank's
    ``(pipeline_group, replica_group, replica_index, stage_index, replicas)``."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if world % stages:
        raise ValueError(f"world size {world} is not a multiple of the stage count {stages}")
    replicas = world // stages
    my_replica, my_stage = divmod(rank, stages)
    pipe_group = rep_group = None
    for r in range(replicas):
        g = dist.new_group([r * stages + j for j in range(stages)])
        if r == my_replica:
            pipe_group = g
    for j in range(stages):
        g = dist.new_group([r * stages + j for r in range(replicas)])
        if j == my_stage:
            rep_group = g
    return pipe_group, rep_group, my_replica, my_stage, replicas


class HybridPipe(DistPipe):
    """A :class:`DistPipe` stage that also belongs to a data-parallel replica group.

    Args:
        stage: this rank's layers, on this rank's device (identical initial weights on every
            replica: build with a fixed seed).
        stages: pipeline depth ``n``; the world size must be ``k * n``.
        chunks, checkpoint: as for :class:`DistPipe` (all three checkpoint modes are allowed).
    """

    def __init__(self, stage: nn.Module, stages: int, chunks: int = 1, checkpoint: str = "except_last",
                 device: Optional[torch.device] = None, transport: str = "nccl", **dist_pipe_kwargs) -> None:
        pipe_group, rep_group, replica, stage_idx, replicas = hybrid_groups(stages)
        world = dist.get_world_size()
        pipelines = [[r * stages + j for j in range(stages)] for r in range(world // stages)]
        super().__init__(stage, chunks=chunks, checkpoint=checkpoint, group=pipe_group, device=device,
                         transport=transport, pipelines=pipelines, **dist_pipe_kwargs)
        self.replica_group = rep_group
        self.replica = replica
        self.replicas = replicas

    def make_optimizer(self, reduce_backend: str = "gather", max_tokens: int = 0, **adam_kwargs):
        """A :class:`~pipe_b200.ops.optim.FusedClipAdam` that averages this stage's gradients over the replica
        group: ``"gather"`` (operand exchange during backward, reduction inside the wgrad GEMM; needs
        ``max_tokens`` = tokens of one replica's mini-batch), ``"allreduce"`` (bucketed NVLink all-reduce kernel
        under the wgrad GEMMs) or ``"nccl"`` (bucketed asynchronous NCCL all-reduce, the baseline)."""
        from ..ops.optim import FusedClipAdam
        if reduce_backend == "none":
            # measurement only: no averaging at all (the replicas drift apart); what a step costs without the reduction
            return FusedClipAdam(self.stage, distributed=True, group=self.group, **adam_kwargs)
        return FusedClipAdam(self.stage, distributed=True, group=self.group, replica_group=self.replica_group,
                             replicas=self.replicas, reduce_backend=reduce_backend, max_tokens=max_tokens,
                             **adam_kwargs)
