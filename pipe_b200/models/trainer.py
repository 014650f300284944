"""The flagship training step: tutorial Transformer LM in a pipeline.

One object, :class:`LMTrainer`, covers the layouts the benchmarks need:

* one process, ``n`` local GPUs  -> :class:`pipe_b200.Pipe` (the reference's shape);
* ``torchrun`` with one rank per GPU -> :class:`pipe_b200.parallel.DistPipe`.

The step is the reference's (``main.py:203-220``): ``zero_grad`` -> forward ->
cross-entropy on the last stage -> backward -> ``clip_grad_norm_(0.5)`` (global
norm over all stages) -> Adam.  ``impl="eager"`` uses stock ``torch.nn`` bodies
(fp32 / tf32 / bf16-autocast) and is the on-box baseline; ``impl="fused"`` uses the
sm_100a kernels, bf16 compute with fp32 master weights.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import Tensor, nn

from ..api import Pipe
from ..parallel.dist_pipe import DistPipe
from .transformer import TransformerConfig, build_stage, build_stages

__all__ = ["LMTrainer", "AutocastStage", "clip_grad_norm_"]


class AutocastStage(nn.Module):
    """Runs a stage body under bf16 autocast.  Autocast state is thread-local and
    the pipeline runs stages on worker threads, so it has to be entered inside
    the stage."""

    def __init__(self, body: nn.Module, dtype: torch.dtype = torch.bfloat16):
        super().__init__()
        self.body = body
        self.dtype = dtype

    def forward(self, *inputs):
        with torch.autocast("cuda", dtype=self.dtype):
            out = self.body(*inputs)
        return out


def clip_grad_norm_(parameters: Iterable[nn.Parameter], max_norm: float,
                    group: Optional[dist.ProcessGroup] = None, distributed: bool = False) -> Tensor:
    """Global-norm clipping over parameters that may live on several devices
    and, with ``distributed=True``, on several ranks (one stage each): squared
    norms are summed per device, then across devices / ranks, and every gradient
    is scaled by ``min(1, max_norm / total)``.  Matches
    ``torch.nn.utils.clip_grad_norm_`` on a single device."""
    grads = [p.grad for p in parameters if p.grad is not None]
    if not grads:
        return torch.zeros(())
    by_dev: Dict[torch.device, List[Tensor]] = {}
    for g in grads:
        by_dev.setdefault(g.device, []).append(g)
    first = grads[0].device
    total_sq = torch.zeros((), dtype=torch.float32, device=first)
    for dev, gs in by_dev.items():
        norms = torch._foreach_norm(gs, 2.0)
        sq = torch.stack([n.float() for n in norms]).pow(2).sum()
        total_sq = total_sq + sq.to(first, non_blocking=True)
    if distributed:
        dist.all_reduce(total_sq, group=group)
    total = total_sq.sqrt()
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for dev, gs in by_dev.items():
        torch._foreach_mul_(gs, coef.to(dev, non_blocking=True))
    return total


class LMTrainer:
    """``replicas > 1`` (with ``torchrun``: world = ``replicas * n_stages``) builds the Pipe-inside-DDP hybrid:
    rank = ``replica * n_stages + stage``; every replica is fed its own mini-batch of ``cfg.batch`` rows and the
    stages' gradients are averaged over the replica group (``reduce_backend``: ``gather | allreduce | nccl``,
    see :class:`~pipe_b200.ops.optim.FusedClipAdam`)."""

    def __init__(self, cfg: TransformerConfig, *, n_stages: int = 1, impl: str = "eager",
                 precision: str = "bf16", checkpoint: str = "never", chunks: Optional[int] = None,
                 lr: float = 1e-4, max_norm: float = 0.5, seed: int = 0, replicas: int = 1,
                 reduce_backend: str = "gather", static_shapes: bool = False, timeline: bool = False,
                 single_process: bool = False) -> None:
        """``single_process=True`` builds the in-process :class:`~pipe_b200.Pipe` over ``n_stages`` local
        GPUs even when a process group exists (benchmarks compare the two launchers inside one job)."""
        self.cfg = cfg
        self.impl = impl
        self.precision = precision
        self.max_norm = max_norm
        self.chunks = chunks or cfg.chunks
        self.replicas = int(replicas)
        self.distributed = (not single_process and dist.is_available() and dist.is_initialized()
                            and dist.get_world_size() > 1)
        torch.manual_seed(seed)

        if precision == "tf32":
            torch.backends.cuda.matmul.allow_tf32 = True
            torch.backends.cudnn.allow_tf32 = True

        if self.distributed:
            self.rank = dist.get_rank()
            self.world = dist.get_world_size()
            if n_stages * self.replicas != self.world:
                raise ValueError("with torchrun, n_stages * replicas must equal the world size")
            self.device = torch.device("cuda", torch.cuda.current_device())
            stage_idx = self.rank % n_stages
            body: nn.Module = build_stage(cfg, stage_idx, n_stages, self.device, impl, seed)
            body = self._wrap(body)
            if self.replicas > 1:
                from ..parallel.hybrid import HybridPipe
                self.pipe = HybridPipe(body, stages=n_stages, chunks=self.chunks, checkpoint=checkpoint,
                                       device=self.device, static_shapes=static_shapes, timeline=timeline)
            else:
                self.pipe = DistPipe(body, chunks=self.chunks, checkpoint=checkpoint, device=self.device,
                                     static_shapes=static_shapes, timeline=timeline)
            self.first_device = self.device
            self.last_device = self.device
            self.is_first = self.pipe.is_first
            self.is_last = self.pipe.is_last
        else:
            if self.replicas != 1:
                raise ValueError("replicas > 1 needs one process per GPU (torchrun)")
            self.rank, self.world = 0, 1
            devices = list(range(n_stages))
            model = build_stages(cfg, devices, impl, seed=seed)
            model = nn.Sequential(*[self._wrap(stage) for stage in model])
            self.pipe = Pipe(model, chunks=self.chunks, checkpoint=checkpoint)
            self.first_device = self.pipe.devices[0]
            self.last_device = self.pipe.devices[-1]
            self.device = self.first_device
            self.is_first = self.is_last = True

        self.n_stages = n_stages
        params = list(self.pipe.parameters())
        self.params = params
        if impl == "fused":
            from ..ops.optim import FusedClipAdam
            if self.replicas > 1:
                self.opt = self.pipe.make_optimizer(reduce_backend=reduce_backend, max_tokens=cfg.batch * cfg.bptt,
                                                    lr=lr, max_norm=max_norm)
            else:
                self.opt = FusedClipAdam(self.pipe, lr=lr, max_norm=max_norm, distributed=self.distributed)
        else:
            if self.replicas > 1:
                raise ValueError("the hybrid layout is implemented for impl='fused'")
            self.opt = torch.optim.Adam(params, lr=lr, fused=True)

    def set_checkpoint(self, mode: str) -> None:
        """Switch the checkpoint mode of the running pipeline (benchmarks sweep the modes on one model)."""
        if mode not in ("always", "except_last", "never"):
            raise ValueError("checkpoint is not one of 'always', 'except_last', or 'never'")
        if isinstance(self.pipe, DistPipe):
            self.pipe.checkpoint = mode
        else:
            self.pipe.checkpoint = mode
            self.pipe.pipeline.checkpoint_stop = {"always": self.pipe.chunks, "except_last": self.pipe.chunks - 1,
                                                  "never": 0}[mode]

    def close(self) -> None:
        """Stop worker threads, detach optimizer hooks and drop the model so its memory can be reused."""
        if getattr(self, "opt", None) is not None and hasattr(self.opt, "close"):
            self.opt.close()
        if getattr(self, "pipe", None) is not None and hasattr(self.pipe, "close"):
            self.pipe.close()
        self.opt = None
        self.pipe = None
        self.params = []

    def _wrap(self, stage: nn.Module) -> nn.Module:
        if self.impl == "eager" and self.precision == "bf16":
            return AutocastStage(stage)
        return stage

    @property
    def bubble(self) -> float:
        from ..runtime.schedule import bubble_fraction
        m = len(torch.empty(self.cfg.batch).chunk(self.chunks))
        return bubble_fraction(m, self.n_stages)

    # -- one training step -------------------------------------------------------
    def step(self, tokens: Optional[Tensor], targets: Optional[Tensor]) -> Optional[Tensor]:
        """``tokens``: ``(N, S)`` int64 on the first stage's device (first rank).
        ``targets``: ``(N*S,)`` int64, batch-major like the gathered logits.  The eager
        arm wants them on the *last* stage's device (last rank), as in the reference
        (``main.py:217``); the fused arm feeds them through the pipe next to the
        tokens (first stage's device / first rank) so that the last stage can fuse the
        loss behind the vocabulary projection.  Returns the mean loss on the last
        stage (``None`` on other ranks)."""
        if self.impl == "fused":
            return self._step_fused(tokens, targets)
        V = self.cfg.ntokens
        self.opt.zero_grad(set_to_none=True)
        if self.distributed:
            out = self.pipe(tokens) if self.is_first else self.pipe()
            loss = None
            if self.is_last:
                logits = out.local_value()
                loss = F.cross_entropy(logits.reshape(-1, V).float(), targets)
            self.pipe.backward(loss)
        else:
            logits = self.pipe(tokens).local_value()
            loss = F.cross_entropy(logits.reshape(-1, V).float(), targets)
            loss.backward()
        clip_grad_norm_(self.params, self.max_norm, distributed=self.distributed)
        self.opt.step()
        return loss

    def _step_fused(self, tokens: Optional[Tensor], targets: Optional[Tensor]) -> Optional[Tensor]:
        cfg = self.cfg
        self.opt.zero_grad()
        loss = None
        if self.distributed:
            if self.is_first:
                out = self.pipe(tokens, targets.view(cfg.batch, cfg.bptt))
            else:
                out = self.pipe()
            if self.is_last:
                loss = out.local_value().sum() / float(cfg.batch * cfg.bptt)
            self.pipe.backward(loss)
        else:
            per_seq = self.pipe(tokens, targets.view(cfg.batch, cfg.bptt)).local_value()
            loss = per_seq.sum() / float(cfg.batch * cfg.bptt)
            loss.backward()
        self.opt.step()
        return loss
