"""``FusedClipAdam``: global-norm clipping + Adam for every stage, one launch per device.

What the reference does with three passes over the parameters
(``clip_grad_norm_`` = norm pass + scale pass, then ``Adam.step``;
``main.py:219-220``) is one kernel here (``csrc/optim.cu``): the squared gradient
norm is already on the device when backward ends — the wgrad GEMM epilogues add
their tiles' squares to a scalar, one small kernel covers biases / LayerNorm /
embedding — and the Adam kernel derives the clip factor from that scalar itself,
updates the fp32 master weights and moments, and rewrites the bf16 (and
transposed bf16) shadows the GEMMs read.  No host synchronisation anywhere in
the step.

Gradient storage: matrix gradients are persistent fp32 buffers that the deferred
wgrad GEMM *overwrites* once per step (no zeroing, no read-modify-write); all
small gradients (biases, LayerNorm, embedding rows) live in one flat buffer per
device that is cleared with one memset.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist
from torch import Tensor, nn

from . import _lib, wgrad
from . import functional as F2

__all__ = ["FusedClipAdam"]

_DESC = np.dtype([("w", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("w16", "<u8"), ("w16t", "<u8"),
                  ("rows", "<i4"), ("cols", "<i4"), ("ld_t", "<i4"), ("tile_start", "<i4")])
assert _DESC.itemsize == 64


class _DeviceState:
    def __init__(self, device: torch.device) -> None:
        self.device = device
        self.params: List[nn.Parameter] = []
        self.table: Optional[Tensor] = None
        self.n_tensors = 0
        self.total_tiles = 0
        self.flat_small: Optional[Tensor] = None
        self.normsq = torch.zeros((), dtype=torch.float32, device=device)
        self.total_normsq = torch.zeros((), dtype=torch.float32, device=device)
        self.keep: List[Tensor] = []


class FusedClipAdam:
    """Args mirror ``torch.optim.Adam`` plus ``max_norm`` (<= 0 disables clipping),
    ``grad_scale`` (multiplies every gradient) and ``distributed`` (sum the squared norm over the
    ranks of ``group`` — one stage per rank).

    ``replica_group`` / ``replicas`` switch on data parallelism across pipeline replicas (the Pipe-inside-DDP
    hybrid).  ``reduce_backend`` picks how the gradients are averaged over the replica group:

    ``"gather"`` (default; ``"fused"`` is an alias)
        weight matrices: the replicas exchange the stacked wgrad *operands* while backward runs (the transpose
        kernels store into every replica's stash over NVLink) and the averaging happens inside the wgrad GEMM
        (:mod:`pipe_b200.parallel.replica_stash`); needs ``max_tokens``.  Small gradients (biases, LayerNorm,
        embedding) go through the all-reduce kernel on a side stream, next to the wgrad GEMMs.
    ``"allreduce"``
        every gradient lives in ONE symmetric buffer; the hand-written NVLink all-reduce kernel
        (``csrc/allreduce.cu``: scale fused) runs per bucket — one bucket per weight matrix — on a side
        stream as soon as that matrix's wgrad GEMM has been enqueued, i.e. under the remaining GEMMs.
    ``"nccl"``
        the baseline: the same buckets, ``dist.all_reduce(async_op=True)`` per bucket as it becomes ready
        (what DDP does), ``1/k`` folded into the optimizer's gradient scale.
    """

    def __init__(self, module: nn.Module, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, max_norm: float = 0.0, grad_scale: float = 1.0,
                 distributed: bool = False, group=None, replica_group=None, replicas: int = 1,
                 reduce_backend: str = "gather", max_tokens: int = 0, reduce_ctas: int = 32) -> None:
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.max_norm, self.grad_scale = max_norm, grad_scale
        self.distributed, self.group = distributed, group
        if reduce_backend == "fused":
            reduce_backend = "gather"
        if reduce_backend not in ("gather", "allreduce", "nccl"):
            raise ValueError("reduce_backend is not one of 'gather', 'allreduce', or 'nccl'")
        self.replica_group, self.replicas, self.reduce_backend = replica_group, int(replicas), reduce_backend
        self.reduce_ctas = int(reduce_ctas)
        self.step_count = 0
        self.module = module
        self._cpu: Optional[torch.optim.Optimizer] = None
        self._armed = True             # zero_grad() ran since the last step()
        self._accumulated_seen = wgrad.accumulated_flushes
        self._works: List = []
        self._reduce_stream: Optional[torch.cuda.Stream] = None
        self._stash = None
        self.exposed_reduce_events = None
        if not any(p.is_cuda for p in module.parameters()):
            # Non-CUDA parameters (the fused modules' reference path): same surface, same semantics —
            # average over replicas, clip by the global norm, Adam — on stock PyTorch ops.
            self._cpu_params = [p for p in module.parameters() if p.requires_grad]
            self._cpu = torch.optim.Adam(self._cpu_params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
            self._cpu_norm = 0.0
            self.devs = {}
            self._shadows = {}
            self._gemm_written = set()
            return

        shadows = {}
        gemm_written = set()
        lgs = []
        for sub in module.modules():
            for sh in getattr(sub, "pb2_shadows", lambda: [])():
                shadows[id(sh.param)] = sh
            for lg in getattr(sub, "pb2_linear_grads", lambda: [])():
                gemm_written.add(id(lg.weight))
                lgs.append(lg)
        self._shadows = shadows
        self._lgs = lgs

        self.devs: Dict[torch.device, _DeviceState] = {}
        seen = set()
        for p in module.parameters():
            if id(p) in seen or not p.requires_grad:
                continue
            seen.add(id(p))
            if p.dtype != torch.float32 or not p.is_cuda:
                raise TypeError("FusedClipAdam needs fp32 CUDA parameters (masters)")
            st = self.devs.setdefault(p.device, _DeviceState(p.device))
            st.params.append(p)

        hybrid = self.replicas > 1
        if hybrid and len(self.devs) != 1:
            raise ValueError("replicated (hybrid) mode expects one stage = one device per rank")
        # gather backend: embedding tables exchange their gradient rows instead of all-reducing V x E floats
        embed_mods = [sub for sub in module.modules() if hasattr(sub, "pb2_embed_weight")] \
            if hybrid and self.reduce_backend == "gather" else []
        exchanged = {id(m.pb2_embed_weight()) for m in embed_mods}
        # dense: every gradient of the stage in one flat buffer [matrices | small], reduced bucket by bucket
        dense = hybrid and self.reduce_backend in ("allreduce", "nccl")
        pad = 32                       # bucket granularity: a multiple of 4 * world for every world <= 8
        for st in self.devs.values():
            small = [p for p in st.params if id(p) not in gemm_written and id(p) not in exchanged]
            local = [p for p in st.params if id(p) in exchanged]
            big = [p for p in st.params if id(p) in gemm_written]
            st.flat_local = None
            if local:
                # gradients this rank forms itself, already averaged (EmbedStash.scatter adds into them): zeroed per step
                st.flat_local = torch.zeros(sum((p.numel() + pad - 1) // pad * pad for p in local), dtype=torch.float32,
                                            device=st.device)
                at = 0
                for p in local:
                    p.grad = st.flat_local[at:at + p.numel()].view_as(p)
                    at += (p.numel() + pad - 1) // pad * pad
            n_small = max(sum((p.numel() + pad - 1) // pad * pad for p in small), pad)
            n_big = sum((p.numel() + pad - 1) // pad * pad for p in big) if dense else 0
            st.symm = None
            st.buckets = {}
            if hybrid:
                from ..parallel.symmetric import SymmetricBuffer
                if self.reduce_backend == "nccl":
                    flat = torch.zeros(n_big + n_small, dtype=torch.float32, device=st.device)
                else:
                    st.symm = SymmetricBuffer(n_big + n_small, self.replica_group, st.device)
                    flat = st.symm.data
                st.flat = flat
                st.flat_small = flat[n_big:n_big + n_small]
                st.small_range = (n_big, n_big + n_small)
                at = 0
                for p in big:
                    if dense:
                        p.grad = flat[at:at + p.numel()].view_as(p)
                        st.buckets[id(p)] = (at, at + (p.numel() + pad - 1) // pad * pad)
                        at += (p.numel() + pad - 1) // pad * pad
            else:
                st.flat = None
                st.flat_small = torch.zeros(n_small, dtype=torch.float32, device=st.device)
            at = 0
            for p in small:
                p.grad = st.flat_small[at:at + p.numel()].view_as(p)
                at += (p.numel() + pad - 1) // pad * pad
            rows_desc = np.zeros(len(st.params), dtype=_DESC)
            tile = 0
            for k, p in enumerate(st.params):
                if id(p) in gemm_written:
                    if not dense:
                        p.grad = torch.zeros_like(p)
                        p._pb2_normsq = st.normsq      # every replica writes the final gradient: fuse its norm into wgrad
                    elif hasattr(p, "_pb2_normsq"):
                        del p._pb2_normsq
                    p._pb2_fresh = True
                    p._pb2_sumsq_done = False
                m = torch.zeros_like(p)
                v = torch.zeros_like(p)
                st.keep += [m, v]
                rows, cols = (p.shape[0], p.numel() // p.shape[0]) if p.dim() >= 2 else (1, p.numel())
                d = rows_desc[k]
                d["w"], d["g"], d["m"], d["v"] = p.data_ptr(), p.grad.data_ptr(), m.data_ptr(), v.data_ptr()
                sh = shadows.get(id(p))
                if sh is not None:
                    sh.ensure()
                    d["w16"] = sh.w16.data_ptr()
                    if sh.w16t is not None:
                        d["w16t"], d["ld_t"] = sh.w16t.data_ptr(), sh.w16t.stride(0)
                d["rows"], d["cols"], d["tile_start"] = rows, cols, tile
                tile += ((rows + 31) // 32) * ((cols + 63) // 64)
            st.table = torch.from_numpy(rows_desc.view(np.uint8).copy()).to(st.device)
            st.n_tensors = len(st.params)
            st.total_tiles = tile
        self._gemm_written = gemm_written

        if hybrid:
            (st,) = self.devs.values()
            self._reduce_stream = torch.cuda.Stream(st.device)
            if self.reduce_backend == "gather":
                if max_tokens <= 0:
                    raise ValueError("reduce_backend='gather' needs max_tokens (tokens a replica stacks per step)")
                from ..parallel.replica_stash import ReplicaStash
                self._stash = ReplicaStash(lgs, self.replica_group, st.device, max_tokens, embeddings=embed_mods)
            else:
                for lg in lgs:
                    lg.on_flushed = self._bucket_ready
            wgrad.pre_flush_hooks.append(self._backward_done)

    # -- torch.optim-like surface --------------------------------------------------
    def zero_grad(self, set_to_none: bool = True) -> None:
        """Small gradients: one memset.  Matrix gradients: nothing — the next wgrad
        GEMM overwrites them (``set_to_none`` is accepted for API compatibility;
        buffers are always kept)."""
        if self._cpu is not None:
            self._cpu.zero_grad(set_to_none=set_to_none)
            return
        for st in self.devs.values():
            st.flat_small.zero_()
            if getattr(st, "flat_local", None) is not None:
                st.flat_local.zero_()
            st.normsq.zero_()
            for p in st.params:
                if id(p) in self._gemm_written:
                    p._pb2_fresh = True
                    p._pb2_sumsq_done = False
        self._armed = True

    # -- replica reduction (hybrid) ----------------------------------------------------
    def _side(self) -> "torch.cuda.Stream":
        """The reduce stream, made to wait for everything enqueued so far on the current stream."""
        side = self._reduce_stream
        side.wait_stream(torch.cuda.current_stream(side.device))
        return side

    def _backward_done(self, todo) -> None:
        """Pre-flush hook: backward has ended (the side-stream transposes were joined), the deferred wgrad
        GEMMs are about to be enqueued.  The small gradients are final: start their reduction on the reduce
        stream so that it runs next to the GEMMs."""
        if not any(id(lg.weight) in self._gemm_written for lg in todo):
            return
        (st,) = self.devs.values()
        lo, hi = st.small_range
        if self.reduce_backend == "nccl":
            self._works.append(dist.all_reduce(st.flat[lo:hi], group=self.replica_group, async_op=True))
            return
        side = self._side()
        st.flat.record_stream(side)
        st.symm.all_reduce_(scale=1.0 / self.replicas, ctas=self.reduce_ctas, stream=side.cuda_stream, lo=lo, hi=hi)

    def _bucket_ready(self, lg) -> None:
        """A weight matrix's wgrad GEMM has been enqueued (dense backends): reduce its bucket under the
        GEMMs that follow."""
        (st,) = self.devs.values()
        rng = st.buckets.get(id(lg.weight))
        if rng is None:
            return
        lo, hi = rng
        if self.reduce_backend == "nccl":
            self._works.append(dist.all_reduce(st.flat[lo:hi], group=self.replica_group, async_op=True))
            return
        side = self._side()
        st.symm.all_reduce_(scale=1.0 / self.replicas, ctas=self.reduce_ctas, stream=side.cuda_stream, lo=lo, hi=hi)

    def _join_reduction(self) -> None:
        """Make the current stream wait for every reduction launched since backward ended."""
        for w in self._works:
            w.wait()
        self._works = []
        if self._reduce_stream is not None:
            torch.cuda.current_stream(self._reduce_stream.device).wait_stream(self._reduce_stream)

    def _hybrid_normsq(self) -> None:
        """Squared norm of the averaged gradients of all stages, on every rank."""
        (st,) = self.devs.values()
        if self.reduce_backend == "gather":
            # matrices: summed into st.normsq by the wgrad epilogues (identical on every replica)
            F2.sumsq(st.flat_small, st.normsq)
            if st.flat_local is not None:
                F2.sumsq(st.flat_local, st.normsq)
            for p in st.params:
                if id(p) in self._gemm_written and not getattr(p, "_pb2_sumsq_done", False):
                    F2.sumsq(p.grad.view(-1), st.normsq)
        else:
            F2.sumsq(st.flat, st.normsq)
        st.total_normsq.copy_(st.normsq)
        # The local sums come from fp32 atomics, so two replicas holding identical gradients can still differ in
        # the last bits of their norm — and then in the clip factor and the weights.  Summing every rank's value
        # and dividing by the replica count gives all ranks the same number.
        stages = dist.get_world_size(self.group) if self.distributed else 1
        if self.distributed and dist.get_world_size() == stages * self.replicas:
            dist.all_reduce(st.total_normsq)                             # stages x replicas = the whole job
        else:
            dist.all_reduce(st.total_normsq, group=self.replica_group)
            if self.distributed:
                dist.all_reduce(st.total_normsq, group=self.group)      # over the stages of this pipeline
        st.total_normsq.mul_(1.0 / self.replicas)

    def _total_normsq(self) -> None:
        """Finish the squared norm on every device, then share the total."""
        for st in self.devs.values():
            with torch.cuda.device(st.device):
                F2.sumsq(st.flat_small, st.normsq)
                for p in st.params:
                    if id(p) in self._gemm_written and not getattr(p, "_pb2_sumsq_done", False):
                        F2.sumsq(p.grad.view(-1), st.normsq)
        states = list(self.devs.values())
        first = states[0]
        total = first.total_normsq
        total.copy_(first.normsq)
        for st in states[1:]:
            total.add_(st.normsq.to(first.device, non_blocking=True))
        if self.distributed:
            dist.all_reduce(total, group=self.group)
        for st in states[1:]:
            st.total_normsq.copy_(total, non_blocking=True)

    def _step_cpu(self) -> None:
        params = [p for p in self._cpu_params if p.grad is not None]
        if self.replicas > 1:
            for p in params:
                dist.all_reduce(p.grad, group=self.replica_group)
                p.grad.mul_(1.0 / self.replicas)
        if self.grad_scale != 1.0:
            for p in params:
                p.grad.mul_(self.grad_scale)
        total = torch.zeros((), dtype=torch.float32)
        for p in params:
            total += p.grad.float().pow(2).sum()
        if self.distributed:
            dist.all_reduce(total, group=self.group)
        self._cpu_norm = float(total.sqrt())
        if self.max_norm > 0:
            coef = min(1.0, self.max_norm / (self._cpu_norm + 1e-6))
            for p in params:
                p.grad.mul_(coef)
        self._cpu.step()

    @torch.no_grad()
    def step(self) -> None:
        wgrad.flush_pending()
        self.step_count += 1
        if self._cpu is not None:
            self._step_cpu()
            return
        if wgrad.accumulated_flushes != self._accumulated_seen:
            self._accumulated_seen = wgrad.accumulated_flushes
            self._armed = False            # gradients were accumulated over several backward passes
        if not self._armed:
            # step() without zero_grad() since the last one (gradient accumulation, a retried step): the norm
            # scalar still holds the previous step's sum and the matrix gradients were accumulated outside the
            # fused epilogue sums — start the norm over from the gradients as they are now
            for st in self.devs.values():
                st.normsq.zero_()
                for p in st.params:
                    if id(p) in self._gemm_written:
                        p._pb2_sumsq_done = False
        self._armed = False
        hybrid = self.replicas > 1
        if hybrid:
            self._join_reduction()
        scale = self.grad_scale
        if hybrid and self.reduce_backend == "nccl":
            scale = self.grad_scale / self.replicas           # the sum was reduced; the mean is folded into the update
        if self.max_norm > 0:
            if hybrid:
                self._hybrid_normsq()
            else:
                self._total_normsq()
        lib = _lib.load()
        for st in self.devs.values():
            with torch.cuda.device(st.device):
                rc = lib.pb2_adam_step(ctypes.c_void_p(st.table.data_ptr()), ctypes.c_int(st.n_tensors),
                                       ctypes.c_int(st.total_tiles), ctypes.c_void_p(st.total_normsq.data_ptr()),
                                       ctypes.c_float(self.lr), ctypes.c_float(self.betas[0]),
                                       ctypes.c_float(self.betas[1]), ctypes.c_float(self.eps),
                                       ctypes.c_float(self.weight_decay), ctypes.c_int(self.step_count),
                                       ctypes.c_float(self.max_norm), ctypes.c_float(scale),
                                       _lib.stream_ptr(st.device))
            _lib.check(rc, "pb2_adam_step")
        for sh in self._shadows.values():
            sh.mark_synced()

    def grad_norm(self) -> float:
        """The global gradient norm used by the last ``step`` (synchronises)."""
        if self._cpu is not None:
            return self._cpu_norm
        first = next(iter(self.devs.values()))
        scale = self.grad_scale / self.replicas if (self.replicas > 1 and self.reduce_backend == "nccl") else self.grad_scale
        return float(first.total_normsq.sqrt().item()) * scale

    def close(self) -> None:
        """Detach from the wgrad flush hooks (a second optimizer on the same process would otherwise be
        notified too)."""
        if self._backward_done in wgrad.pre_flush_hooks:
            wgrad.pre_flush_hooks.remove(self._backward_done)
        for lg in getattr(self, "_lgs", []):
            if lg.on_flushed == self._bucket_ready:
                lg.on_flushed = None
        if self._stash is not None:
            self._stash.close()
