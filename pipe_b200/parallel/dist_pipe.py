"""``DistPipe``: the multi-process variant — one rank per stage, one GPU per rank.

The in-process :class:`~pipe_b200.Pipe` drives every device from one Python
process (worker thread per device), which is the reference's design.  On an
8 x B200 box the better fit is one process per GPU: no GIL shared between
stages, each rank launches only its own kernels, and stage boundaries become
NCCL point-to-point transfers over NVLink 5 / NVSwitch (BASELINE.json: "NCCL p2p
in a multi-process variant"; the reference defers this as future work,
``pipe.py:295-302``).

Semantics are the same synchronous GPipe: ``m`` micro-batches flow forward
through ``n`` stages in ``m + n - 1`` ticks, then backward in reverse
micro-batch order; checkpoint modes ``always | except_last | never`` decide
which micro-batches keep their graph and which recompute it (with RNG replay)
just before their backward.  Every rank calls the same two methods::

    out = pipe(tokens)               # first rank feeds inputs; others pass nothing
    if pipe.is_last:
        loss = criterion(out.local_value().view(-1, V), targets)
    pipe.backward(loss if pipe.is_last else None)

Protocol per step:

* *Header.*  The first stage broadcasts ``[m, signature]`` over the pipeline group: the number
  of micro-batches ``Tensor.chunk`` produced and a hash of the input shapes / dtypes / grad
  mode.  Every rank keeps one cache of boundary shapes per signature, so a short last batch or
  an evaluation pass under ``no_grad`` simply selects (or creates, with a fresh shape handshake)
  another cache — all ranks take the same branch by construction (the reference handles a short
  last batch naturally, ``pipe.py:484``).  ``static_shapes=True`` skips the header after the
  first step for jobs whose input never changes; the first stage then refuses a changed input.
* *Boundaries.*  Each direction of each stage boundary has its own two-rank communicator, so a
  rank's pending receive never queues ahead of its own sends on one NCCL stream.  All tensors of
  a micro-batch cross in ONE batched point-to-point operation (``dist.batch_isend_irecv``).  The
  receive for micro-batch ``i + 1`` (forward) / ``i - 1`` (backward) is posted before cell ``i``
  runs, so the transfer — and under checkpointing the recompute — overlaps it (the property the
  reference gets from its two-node Checkpoint/Recompute split).
* *Skips.*  ``@skippable`` stash / pop pairs that span ranks travel **directly** from the
  stashing rank to the popping rank over the pipeline group, not through the stages in between
  (reference ``pipeline.py:134-138``); gradients take the same route back.
* *Deferred batch norm.*  ``deferred_batch_norm=True`` converts the stage's BatchNorm layers
  (reference ``pipe.py:341-342``).

With ``transport="mailbox"`` (or ``PB2_DIST_TRANSPORT=mailbox``) the process group only carries the
first step (it doubles as the shape handshake); from the second step on every boundary talks through an
*NVLink mailbox* (``parallel/mailbox.py``, ``csrc/mailbox.cu``): the sender's kernel stores the activation
straight into the receiver's memory and raises a flag, the receiver's kernel waits for the flag on its
own stream — no NCCL call, no host wait, no early posting.  The default is ``transport="nccl"``.

``timeline=True`` brackets every forward and backward cell with CUDA events (host clocks on CPU);
:meth:`DistPipe.timeline_summary` gathers the per-rank busy time and measured bubble to every rank.

Works with ``gloo`` on CPU (tests) and ``nccl`` on GPUs.
"""
from __future__ import annotations

import os
import time
import zlib
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor, nn

from .. import microbatch
from ..autograd.checkpoint import enable_checkpointing, enable_recomputing
from ..autograd.seeds import SeedBase, seed_scope
from ..microbatch import Batch
from ..rref import LocalRRef
from ..runtime.schedule import bubble_fraction
from ..skip.tracker import SkipTracker, use_skip_tracker

__all__ = ["DistPipe"]

_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32,
           torch.int16, torch.int8, torch.uint8, torch.bool]
_META_LEN = 16  # [n_tensors] then per tensor: dtype code, requires_grad, ndim, dims...

SkipKey = Tuple[Optional[str], str]      # (namespace key or None, name): identical on every rank


class _Spec:
    """Shape / dtype / grad flag of one tensor crossing a boundary.  Two specs are *equal* when
    shape and dtype agree: ``requires_grad`` is per-step state (an evaluation pass after a training
    step sends the same shapes without gradients) and is kept in the cache of its own signature."""

    __slots__ = ("shape", "dtype", "requires_grad")

    def __init__(self, shape: Sequence[int], dtype: torch.dtype, requires_grad: bool):
        self.shape = tuple(shape)
        self.dtype = dtype
        self.requires_grad = requires_grad

    @classmethod
    def of(cls, t: Tensor, force_grad: bool = False) -> "_Spec":
        return cls(t.shape, t.dtype, t.is_floating_point() and (t.requires_grad or force_grad))

    def __eq__(self, other: object) -> bool:
        return isinstance(other, _Spec) and self.shape == other.shape and self.dtype == other.dtype

    def __repr__(self) -> str:
        return f"_Spec({self.shape}, {self.dtype}, grad={self.requires_grad})"


class _Shapes:
    """Everything learned about one input signature: micro-batch count and boundary shapes."""

    def __init__(self, sig: int, m: int) -> None:
        self.sig = sig
        self.m = m
        self.recv: Dict[int, List[_Spec]] = {}          # micro-batch -> tensors from the previous stage
        self.sent: Dict[int, List[_Spec]] = {}
        self.skip_recv: Dict[Tuple[int, SkipKey], _Spec] = {}
        self.skip_sent: Dict[Tuple[int, SkipKey], _Spec] = {}
        self.steps = 0


class _Cell:
    """What one micro-batch leaves behind on this stage between forward and backward."""

    __slots__ = ("inputs", "outputs", "checkpointed", "cpu_rng", "dev_rng", "leaves", "seed_base",
                 "skip_in", "skip_leaves", "skip_out")

    def __init__(self) -> None:
        self.inputs: Tuple[Any, ...] = ()
        self.outputs: Tuple[Any, ...] = ()
        self.checkpointed = False
        self.cpu_rng: Optional[Tensor] = None
        self.dev_rng: Optional[Tensor] = None
        self.leaves: Tuple[Any, ...] = ()
        self.seed_base = SeedBase()
        self.skip_in: Dict[SkipKey, Tensor] = {}         # received skip tensors (detached)
        self.skip_leaves: Dict[SkipKey, Tensor] = {}     # the leaves the stage popped (carry .grad)
        self.skip_out: Dict[SkipKey, Optional[Tensor]] = {}   # stashed tensors, attached to the graph


def _skip_key(ns, name: str) -> SkipKey:
    if ns is None:
        return (None, name)
    key = getattr(ns, "key", None)
    if key is None:
        # a process-local (random) namespace: fine while stash and pop share a rank
        return (f"local:{id(ns)}", name)
    return (str(key), name)


class _RankSkipTracker(SkipTracker):
    """Skip tracker of one micro-batch on one rank: stashes whose pop lives on another rank go to
    an outbox (``DistPipe`` ships them), pops whose stash lives on another rank are served from
    an inbox of received leaves; everything else stays in the dict of the plain tracker."""

    def __init__(self, remote_out: Dict[SkipKey, int], remote_in: Dict[SkipKey, int]) -> None:
        super().__init__()
        self.remote_out = remote_out
        self.remote_in = remote_in
        self.outbox: Dict[SkipKey, Optional[Tensor]] = {}
        self.inbox: Dict[SkipKey, Optional[Tensor]] = {}

    def save(self, batch: Batch, ns, name: str, tensor: Optional[Tensor]) -> None:
        key = _skip_key(ns, name)
        if key in self.remote_out:
            self.outbox[key] = tensor
        else:
            super().save(batch, ns, name, tensor)

    def load(self, batch: Batch, ns, name: str) -> Optional[Tensor]:
        key = _skip_key(ns, name)
        if key in self.remote_in:
            return self.inbox[key]
        return super().load(batch, ns, name)


class DistPipe(nn.Module):
    """One pipeline stage per rank of ``group``.

    Args:
        stage: this rank's layers (already on this rank's device).
        chunks: micro-batches per mini-batch.
        checkpoint: ``'always' | 'except_last' | 'never'``.
        group: process group whose ranks are the stages, in order (default: world).
        device: this rank's device (default: the stage's parameter device).
        transport: ``"nccl"`` (process-group send / recv) or ``"mailbox"`` / ``"auto"`` (NVLink mailboxes).
        pipelines: rank lists of *all* pipelines of the job when there are several (``HybridPipe`` passes them);
            needed because process groups must be created by every rank of the job in the same order.
        deferred_batch_norm: convert the stage's BatchNorm layers to :class:`~pipe_b200.batchnorm.DeferredBatchNorm`
            (statistics committed once per mini-batch; reference ``pipe.py:261-265,341-342``).
        static_shapes: the mini-batch shape never changes: skip the per-step header after the first step.
        timeline: record per-cell device timestamps (see :meth:`timeline_summary`).
    """

    def __init__(self, stage: nn.Module, chunks: int = 1, checkpoint: str = "except_last",
                 group: Optional[dist.ProcessGroup] = None, device: Optional[torch.device] = None,
                 strict_except_last: bool = False, transport: str = "nccl",
                 pipelines: Optional[Sequence[Sequence[int]]] = None, deferred_batch_norm: bool = False,
                 static_shapes: bool = False, timeline: bool = False) -> None:
        super().__init__()
        if transport not in ("auto", "nccl", "mailbox"):
            raise ValueError("transport is not one of 'auto', 'nccl', or 'mailbox'")
        chunks = int(chunks)
        if chunks <= 0:
            raise ValueError("number of chunks must be positive integer")
        if checkpoint not in ("always", "except_last", "never"):
            raise ValueError("checkpoint is not one of 'always', 'except_last', or 'never'")
        if not dist.is_initialized():
            raise RuntimeError("DistPipe needs torch.distributed to be initialised")

        if type(stage) is nn.Sequential:
            # same calling convention as an in-process stage: tuples are splatted between layers
            from ..partition import PipeSequential
            stage = PipeSequential(*stage.children())
        if deferred_batch_norm:
            from ..batchnorm import DeferredBatchNorm
            stage = DeferredBatchNorm.convert_deferred_batch_norm(stage, chunks)
        self.stage = stage
        self.chunks = chunks
        self.checkpoint = checkpoint
        self.strict_except_last = strict_except_last
        self.static_shapes = bool(static_shapes)
        self.group = group
        self.ranks = dist.get_process_group_ranks(group) if group is not None else list(range(dist.get_world_size()))
        self.index = self.ranks.index(dist.get_rank())
        self.n_stages = len(self.ranks)
        if device is None:
            p = next(stage.parameters(), None)
            device = p.device if p is not None else torch.device("cpu")
        self.device = torch.device(device)

        # A communicator per boundary and direction (see module docstring).  ``dist.new_group`` is a
        # collective over the whole job: every rank has to create every group, in the same order, members
        # or not — so when several pipelines run side by side (``pipelines``: the rank lists of all of
        # them, this one included) each rank walks all of them and keeps the groups of its own.
        self._fwd_groups: List[dist.ProcessGroup] = []
        self._bwd_groups: List[dist.ProcessGroup] = []
        all_pipes = [list(p) for p in pipelines] if pipelines is not None else [list(self.ranks)]
        if list(self.ranks) not in all_pipes:
            raise ValueError("pipelines must contain this pipeline's own rank list")
        for ranks in all_pipes:
            mine = ranks == list(self.ranks)
            for j in range(len(ranks) - 1):
                pair = [ranks[j], ranks[j + 1]]
                f, b = dist.new_group(pair), dist.new_group(pair)
                if mine:
                    self._fwd_groups.append(f)
                    self._bwd_groups.append(b)

        self._caches: Dict[int, _Shapes] = {}
        self._cur: Optional[_Shapes] = None
        self._cells: List[_Cell] = []
        self._out_leaves: List[Tuple[Any, ...]] = []
        self._sends: List[Any] = []
        self._early: Dict[int, Any] = {}                  # receives posted ahead of their micro-batch
        self._early_grads: Dict[int, Any] = {}
        self.last_bubble = 0.0

        # Boundary transport: "nccl" (default: process-group send / recv, receives pre-posted on their own
        # communicators), "mailbox" / "auto" (NVLink mailboxes, parallel/mailbox.py: the first step still goes
        # through the process group because it carries the shape handshake).  PB2_DIST_TRANSPORT overrides.
        self.transport = os.environ.get("PB2_DIST_TRANSPORT", transport)
        self._mb: Optional[Dict[str, Any]] = None         # {"fwd_in", "fwd_out", "bwd_in", "bwd_out"} once built
        self._mb_shapes: Optional[_Shapes] = None         # the signature the mailboxes were laid out for
        self._mb_failed = False
        self._fwd_seq = 0
        self._bwd_seq = 0

        self.timeline = bool(timeline)
        self._t_spans: List[Tuple[str, Any, Any]] = []
        #: called as ``hook(i)`` right after micro-batch ``i``'s backward has been enqueued on this stage
        #: (``i == 0`` is the last one: every gradient of the stage is final except the deferred wgrad GEMMs)
        self.cell_backward_hooks: List[Any] = []
        self._plan_skips()

    # -- topology --------------------------------------------------------------
    @property
    def is_first(self) -> bool:
        return self.index == 0

    @property
    def is_last(self) -> bool:
        return self.index == self.n_stages - 1

    def _prev(self) -> int:
        return self.ranks[self.index - 1]

    def _next(self) -> int:
        return self.ranks[self.index + 1]

    def reset(self) -> None:
        """Forget every cached boundary shape.  Not needed when only the *input* changes (the
        per-step header notices); call it on every rank when a stage starts producing different
        shapes for the same input, or after ``static_shapes=True`` refused a new input."""
        self._caches.clear()
        self._cur = None
        self._mb = None            # back to the process group until the new shapes are known
        self._mb_shapes = None
        self._mb_failed = False
        self._early.clear()
        self._early_grads.clear()

    # views of the current signature's caches (tests and the mailbox builder look at them)
    @property
    def _recv_specs(self) -> Dict[int, List[_Spec]]:
        return self._cur.recv if self._cur is not None else {}

    @property
    def _sent_specs(self) -> Dict[int, List[_Spec]]:
        return self._cur.sent if self._cur is not None else {}

    @property
    def _steps(self) -> int:
        return self._cur.steps if self._cur is not None else 0

    # -- skip routes across ranks ------------------------------------------------
    def _plan_skips(self) -> None:
        """Collective over the pipeline: every rank announces what its stage stashes and pops, in
        layer order; every rank derives the same routes ``key -> (stash stage, pop stage)``.
        Reference: ``inspect_skip_layout`` / ``copy_policy`` (``pipe.py:348``, ``pipeline.py:134-138``)."""
        from ..skip.skippable import Skippable
        events: List[Tuple[str, SkipKey]] = []
        self._skip_out: Dict[SkipKey, int] = {}     # key -> stage that pops it
        self._skip_in: Dict[SkipKey, int] = {}      # key -> stage that stashed it
        self._skip_out_order: List[SkipKey] = []
        self._skip_in_order: List[SkipKey] = []
        walk = list(self.stage) if isinstance(self.stage, nn.Sequential) else list(self.stage.modules())
        for layer in walk:
            if not isinstance(layer, Skippable):
                continue
            for ns, name in layer.poppable():
                events.append(("pop", _skip_key(ns, name)))
            for ns, name in layer.stashable():
                events.append(("stash", _skip_key(ns, name)))
        if self.n_stages == 1:
            return
        gathered: List[Any] = [None] * self.n_stages
        dist.all_gather_object(gathered, events, group=self.group)
        if not any(gathered):
            return
        open_at: Dict[SkipKey, int] = {}
        problems: List[str] = []
        for j, evs in enumerate(gathered):
            for kind, key in evs:
                key = tuple(key)
                if kind == "stash":
                    open_at[key] = j
                    continue
                src = open_at.pop(key, None)
                if src is None:
                    hint = (" (a namespace created without a key is process-local: use Namespace(key=...) "
                            "for skips that cross ranks)") if key[0] and str(key[0]).startswith("local:") else ""
                    problems.append(f"stage {j} pops '{key[1]}' which no earlier stage stashes{hint}")
                elif src != j:
                    if src == self.index:
                        self._skip_out[key] = j
                    if j == self.index:
                        self._skip_in[key] = src
        for key, j in open_at.items():
            problems.append(f"stage {j} stashes '{key[1]}' which no later stage pops")
        if problems:
            raise TypeError("skip connections do not match across stages: " + "; ".join(problems))
        self._skip_out_order = sorted(self._skip_out, key=lambda k: (self._skip_out[k], str(k)))
        self._skip_in_order = sorted(self._skip_in, key=lambda k: (self._skip_in[k], str(k)))

    @property
    def _has_skips(self) -> bool:
        return bool(self._skip_out) or bool(self._skip_in)

    # -- NVLink mailboxes ----------------------------------------------------------
    def _mailboxes_wanted(self) -> bool:
        return (self.transport in ("auto", "mailbox") and self.device.type == "cuda" and self._mb is None
                and not self._mb_failed and self._cur is not None and self._cur.steps >= 1
                and torch.is_grad_enabled() and dist.get_backend() == "nccl")

    def _build_mailboxes(self) -> None:
        """Collective over the pipeline: every rank builds the mailboxes of its one or two boundaries,
        boundary by boundary in stage order (so the pairwise rendezvous cannot cycle), and all ranks
        then agree on whether everybody succeeded."""
        from .mailbox import PeerMailbox, mailbox_supported
        mb: Dict[str, Any] = {"fwd_in": None, "fwd_out": None, "bwd_in": None, "bwd_out": None}
        ok = True
        cur = self._cur
        try:
            for b in range(self.n_stages - 1):
                if self.index not in (b, b + 1):
                    continue
                sending = self.index == b
                specs = cur.sent if sending else cur.recv
                fwd = {i: [(tuple(s.shape), s.dtype) for s in v] for i, v in specs.items()}
                bwd = {i: [(tuple(s.shape), s.dtype) for s in v if s.requires_grad] for i, v in specs.items()}
                flat = [t for v in fwd.values() for t in v]
                if not fwd or not mailbox_supported(self.device, flat):
                    raise RuntimeError("boundary tensors are not mailbox-compatible")
                f = PeerMailbox(self._fwd_groups[b], self.device, self.ranks[b], fwd)
                g = PeerMailbox(self._bwd_groups[b], self.device, self.ranks[b + 1], bwd) if any(bwd.values()) else None
                if sending:
                    mb["fwd_out"], mb["bwd_in"] = f, g
                else:
                    mb["fwd_in"], mb["bwd_out"] = f, g
        except Exception as e:  # noqa: BLE001 - any failure means "stay on the process group"
            ok = False
            self._mb_error = f"{type(e).__name__}: {e}"
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 1:
            self._mb = mb
            self._mb_shapes = cur
            self._fwd_seq = 0
            self._bwd_seq = 0
        else:
            self._mb_failed = True
            if self.transport == "mailbox":
                raise RuntimeError("transport='mailbox' requested but not available: "
                                   + getattr(self, "_mb_error", "a peer could not build its mailbox"))

    def _box(self, name: str):
        """The mailbox of one boundary direction, if mailboxes exist for the current signature."""
        if self._mb is None or self._mb_shapes is not self._cur:
            return None
        return self._mb[name]

    def _mb_slot(self, box, i: int, specs: Optional[List[_Spec]]) -> bool:
        return box is not None and specs is not None and (i, 0) in box.nbytes

    # -- per-step header -----------------------------------------------------------
    def _signature(self, inputs: Sequence[Any], m: int) -> int:
        desc: List[Any] = [m, int(torch.is_grad_enabled()), int(self.stage.training)]
        for v in inputs:
            t = v.tensor if isinstance(v, microbatch.NoChunk) else v
            if torch.is_tensor(t):
                desc.append((tuple(t.shape), str(t.dtype), bool(t.requires_grad)))
            else:
                desc.append(type(t).__name__)
        return zlib.crc32(repr(desc).encode()) & 0x7FFFFFFF

    def _begin_step(self, inputs: Sequence[Any], m_first: int) -> int:
        """Agree on (micro-batch count, signature) and select that signature's shape cache."""
        sig = self._signature(inputs, m_first) if self.is_first else -1
        if self.static_shapes and self._cur is not None:
            if self.is_first and sig != self._cur.sig:
                raise RuntimeError("static_shapes=True but the mini-batch changed (shape, dtype or grad mode); "
                                   "build the pipe with static_shapes=False, or call reset() on every rank")
            return self._cur.m
        if self.n_stages > 1:
            t = torch.tensor([m_first, sig], dtype=torch.int64, device=self.device)
            dist.broadcast(t, self.ranks[0], group=self.group)
            if not self.is_first:
                m_first, sig = (int(v) for v in t.tolist())
        cur = self._caches.get(sig)
        if cur is None or cur.m != m_first:
            cur = self._caches[sig] = _Shapes(sig, m_first)
        if cur is not self._cur:
            self._early.clear()
            self._early_grads.clear()
        self._cur = cur
        return cur.m

    # -- boundary transfers ------------------------------------------------------
    def _p2p(self, ops: List[Any]) -> List[Any]:
        """One batched point-to-point operation for all tensors of a micro-batch (NCCL coalesces
        them into one group launch; gloo issues them one by one)."""
        if not ops:
            return []
        return list(dist.batch_isend_irecv(ops))

    def _send_meta(self, specs: List[_Spec], dst: int, group) -> None:
        meta = [len(specs)]
        for s in specs:
            meta += [_DTYPES.index(s.dtype), int(s.requires_grad), len(s.shape), *s.shape]
        if len(meta) > _META_LEN * 8:
            raise ValueError("too many / too high-rank tensors cross the stage boundary")
        meta += [0] * (_META_LEN * 8 - len(meta))
        t = torch.tensor(meta, dtype=torch.int64, device=self.device)
        for w in self._p2p([dist.P2POp(dist.isend, t, dst, group)]):
            w.wait()

    def _recv_meta(self, src: int, group) -> List[_Spec]:
        buf = torch.empty(_META_LEN * 8, dtype=torch.int64, device=self.device)
        for w in self._p2p([dist.P2POp(dist.irecv, buf, src, group)]):
            w.wait()
        meta = buf.tolist()
        specs, at = [], 1
        for _ in range(meta[0]):
            code, rg, nd = meta[at:at + 3]
            shape = meta[at + 3:at + 3 + nd]
            specs.append(_Spec(shape, _DTYPES[code], bool(rg)))
            at += 3 + nd
        return specs

    def _send_forward(self, i: int, values: Tuple[Any, ...], force_grad: bool) -> None:
        tensors = [v for v in values if torch.is_tensor(v)]
        if len(tensors) != len(values):
            raise TypeError("only tensors can cross a DistPipe stage boundary")
        specs = [_Spec.of(t, force_grad) for t in tensors]
        group = self._fwd_groups[self.index]
        known = self._cur.sent.get(i)
        if known is None:
            self._send_meta(specs, self._next(), group)
            self._cur.sent[i] = specs
        elif known != specs:
            raise RuntimeError(f"stage {self.index}: output shapes of micro-batch {i} changed for an unchanged input "
                               f"({known} -> {specs}); call DistPipe.reset() on every rank")
        box = self._box("fwd_out")
        if known is not None and self._mb_slot(box, i, known):
            for k, t in enumerate(tensors):
                box.send(i, k, t.detach().contiguous(), self._fwd_seq)
            return
        ops = [dist.P2POp(dist.isend, t.detach().contiguous(), self._next(), group) for t in tensors]
        self._sends += self._p2p(ops)

    def _post_recv_forward(self, i: int) -> None:
        """Queue the receives of micro-batch ``i`` now (its shapes must be known), so the transfer
        is already in flight — or done — when the stage gets to it."""
        specs = self._cur.recv.get(i)
        if i in self._early or specs is None:
            return
        if self._mb_slot(self._box("fwd_in"), i, specs):
            return              # the sender pushes into the mailbox on its own; nothing to post
        group = self._fwd_groups[self.index - 1]
        out = [torch.empty(s.shape, dtype=s.dtype, device=self.device) for s in specs]
        works = self._p2p([dist.P2POp(dist.irecv, t, self._prev(), group) for t in out])
        self._early[i] = (out, works)

    def _recv_forward(self, i: int) -> Tuple[Tensor, ...]:
        group = self._fwd_groups[self.index - 1]
        specs = self._cur.recv.get(i)
        grad_on = torch.is_grad_enabled()
        box = self._box("fwd_in")
        if self._mb_slot(box, i, specs):
            out = [torch.empty(s.shape, dtype=s.dtype, device=self.device) for s in specs]
            for k, t in enumerate(out):
                box.recv(i, k, t, self._fwd_seq)
            return tuple(t.requires_grad_(s.requires_grad and grad_on) for t, s in zip(out, specs))
        if specs is None:
            specs = self._cur.recv[i] = self._recv_meta(self._prev(), group)
        if i in self._early:
            out, works = self._early.pop(i)
        else:
            out = [torch.empty(s.shape, dtype=s.dtype, device=self.device) for s in specs]
            works = self._p2p([dist.P2POp(dist.irecv, t, self._prev(), group) for t in out])
        for w in works:
            w.wait()
        return tuple(t.requires_grad_(s.requires_grad and grad_on) for t, s in zip(out, specs))

    def _grads_by_mailbox(self, i: int) -> bool:
        return self._mb_slot(self._box("bwd_in"), i, self._cur.sent.get(i))

    def _recv_grads_mailbox(self, i: int):
        box = self._box("bwd_in")
        bufs, k = [], 0
        for s in self._cur.sent[i]:
            if not s.requires_grad:
                bufs.append(None)
                continue
            g = torch.empty(s.shape, dtype=s.dtype, device=self.device)
            box.recv(i, k, g, self._bwd_seq)
            bufs.append(g)
            k += 1
        return bufs

    def _post_recv_grads(self, i: int):
        """Post receives for the gradients of micro-batch ``i``'s outputs."""
        group = self._bwd_groups[self.index]
        bufs: List[Optional[Tensor]] = []
        for s in self._cur.sent[i]:
            bufs.append(torch.empty(s.shape, dtype=s.dtype, device=self.device) if s.requires_grad else None)
        works = self._p2p([dist.P2POp(dist.irecv, g, self._next(), group) for g in bufs if g is not None])
        return bufs, works

    def _send_grads(self, i: int, inputs: Tuple[Any, ...]) -> None:
        group = self._bwd_groups[self.index - 1]
        box = self._box("bwd_out")
        by_mailbox = self._mb_slot(box, i, self._cur.recv.get(i))
        k = 0
        ops = []
        for t, s in zip(inputs, self._cur.recv[i]):
            if not s.requires_grad:
                continue
            g = t.grad if t.grad is not None else torch.zeros_like(t)
            if by_mailbox:
                box.send(i, k, g.contiguous(), self._bwd_seq)
                k += 1
            else:
                ops.append(dist.P2POp(dist.isend, g.contiguous(), self._prev(), group))
        self._sends += self._p2p(ops)

    def _drain_sends(self) -> None:
        for w in self._sends:
            w.wait()
        self._sends = []

    # -- skip tensors: direct rank-to-rank --------------------------------------------
    def _skip_send(self, i: int, tracker: "_RankSkipTracker", cell: _Cell, first_time: bool) -> None:
        """Ship this cell's remote stashes to the ranks that pop them (forward only; a recompute
        refreshes the graph-attached tensors but sends nothing)."""
        for key in self._skip_out_order:
            t = tracker.outbox.get(key)
            if t is None:
                raise RuntimeError(f"skip '{key[1]}' crosses ranks and must be a tensor (got None)")
            cell.skip_out[key] = t
            if not first_time:
                continue
            dst = self.ranks[self._skip_out[key]]
            spec = _Spec.of(t, cell.checkpointed and torch.is_grad_enabled())
            known = self._cur.skip_sent.get((i, key))
            if known is None:
                self._send_meta([spec], dst, self.group)
                self._cur.skip_sent[(i, key)] = spec
            elif known != spec:
                raise RuntimeError(f"shape of skip '{key[1]}' changed for an unchanged input; call reset() on every rank")
            self._sends += self._p2p([dist.P2POp(dist.isend, t.detach().contiguous(), dst, self.group)])

    def _skip_recv(self, i: int, cell: _Cell) -> None:
        for key in self._skip_in_order:
            src = self.ranks[self._skip_in[key]]
            spec = self._cur.skip_recv.get((i, key))
            if spec is None:
                spec = self._cur.skip_recv[(i, key)] = self._recv_meta(src, self.group)[0]
            t = torch.empty(spec.shape, dtype=spec.dtype, device=self.device)
            for w in self._p2p([dist.P2POp(dist.irecv, t, src, self.group)]):
                w.wait()
            cell.skip_in[key] = t

    def _skip_tracker(self, i: int, cell: _Cell) -> "_RankSkipTracker":
        tr = _RankSkipTracker(self._skip_out, self._skip_in)
        grad_on = torch.is_grad_enabled()
        for key, t in cell.skip_in.items():
            spec = self._cur.skip_recv[(i, key)]
            leaf = t.detach().requires_grad_(spec.requires_grad and grad_on)
            cell.skip_leaves[key] = leaf
            tr.inbox[key] = leaf
        return tr

    def _skip_grads_recv(self, i: int, cell: _Cell) -> List[Tuple[Tensor, Tensor]]:
        """Gradients of this cell's remote stashes, from the ranks that popped them."""
        pairs = []
        for key in reversed(self._skip_out_order):
            spec = self._cur.skip_sent[(i, key)]
            if not spec.requires_grad:
                continue
            src = self.ranks[self._skip_out[key]]
            g = torch.empty(spec.shape, dtype=spec.dtype, device=self.device)
            for w in self._p2p([dist.P2POp(dist.irecv, g, src, self.group)]):
                w.wait()
            t = cell.skip_out.get(key)
            if t is not None and t.requires_grad:
                pairs.append((t, g))
        return pairs

    def _skip_grads_send(self, i: int, cell: _Cell) -> None:
        for key in reversed(self._skip_in_order):
            spec = self._cur.skip_recv[(i, key)]
            if not spec.requires_grad:
                continue
            leaf = cell.skip_leaves[key]
            g = leaf.grad if leaf.grad is not None else torch.zeros_like(leaf)
            dst = self.ranks[self._skip_in[key]]
            self._sends += self._p2p([dist.P2POp(dist.isend, g.contiguous(), dst, self.group)])

    # -- timeline ------------------------------------------------------------------
    def _mark(self):
        if not self.timeline:
            return None
        if self.device.type == "cuda":
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            return ev
        return time.perf_counter()

    def _elapsed_ms(self, a, b) -> float:
        if a is None or b is None:
            return 0.0
        if self.device.type == "cuda":
            return float(a.elapsed_time(b))
        return (b - a) * 1e3

    def _span(self, phase: str, a, b) -> None:
        if self.timeline and a is not None and b is not None:
            self._t_spans.append((phase, a, b))

    def timeline_summary(self) -> Optional[List[Dict[str, float]]]:
        """Collective.  Per stage, for the last step: milliseconds this rank's stream spent inside
        forward cells, backward cells and recomputes (CUDA events on the rank's own stream; the marks sit
        after the receive wait and before the send, so waiting for a neighbour is *not* busy time), the
        span from its first forward cell to its last backward cell, and the measured bubble
        ``1 - busy / span of stage 0`` — stage 0 starts first and ends last, so its span is the length of
        the pipelined part of the step.  Compare with the ideal ``(n - 1) / (m + n - 1)`` of equal stages."""
        if not self.timeline or not self._t_spans:
            return None
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        mine: Dict[str, float] = {"fwd_busy_ms": 0.0, "bwd_busy_ms": 0.0, "recompute_ms": 0.0}
        for phase, a, b in self._t_spans:
            mine[{"fwd": "fwd_busy_ms", "bwd": "bwd_busy_ms", "rec": "recompute_ms"}[phase]] += self._elapsed_ms(a, b)
        mine["pipe_span_ms"] = self._elapsed_ms(self._t_spans[0][1], self._t_spans[-1][2])
        gathered: List[Any] = [None] * self.n_stages
        dist.all_gather_object(gathered, mine, group=self.group)
        length = gathered[0]["pipe_span_ms"]
        for g in gathered:
            busy = g["fwd_busy_ms"] + g["bwd_busy_ms"] + g["recompute_ms"]
            g["bubble"] = max(0.0, 1.0 - busy / length) if length > 0 else 0.0
        return gathered

    # -- forward -------------------------------------------------------------------
    def _checkpoint_stop(self, m: int) -> int:
        if not self.stage.training or not torch.is_grad_enabled():
            return 0
        basis = m if self.strict_except_last else self.chunks
        return {"always": basis, "except_last": basis - 1, "never": 0}[self.checkpoint]

    def _call(self, inputs: Tuple[Any, ...], tracker: Optional[SkipTracker] = None):
        if tracker is None:
            return self.stage(*inputs)
        with use_skip_tracker(tracker):
            return self.stage(*inputs)

    def forward(self, *inputs: Any, num_microbatches: Optional[int] = None):
        """Run the forward half of a step.  The first stage takes the mini-batch (split along
        dim 0); later stages take no inputs — they learn the micro-batch count from the step header.
        ``num_microbatches`` is accepted for compatibility and ignored."""
        if self.is_first:
            microbatch.check(self.device, *inputs)
            batches = microbatch.scatter(*inputs, chunks=self.chunks)
            m = len(batches)
        else:
            batches = []
            m = -1
        m = self._begin_step(inputs, m)
        if self._mailboxes_wanted():
            self._build_mailboxes()
        self._fwd_seq += 1

        stop = self._checkpoint_stop(m)
        grad_on = torch.is_grad_enabled()
        self._cells = []
        self._out_leaves = []
        self._t_spans = []
        self.last_bubble = bubble_fraction(m, self.n_stages)
        skips = self._has_skips

        for i in range(m):
            cell = _Cell()
            if self.is_first:
                b = batches[i]
                values = (b.tensor,) if b.atomic else tuple(b.values)
            else:
                values = self._recv_forward(i)
                if i + 1 < m:
                    self._post_recv_forward(i + 1)      # next micro-batch's transfer overlaps this cell
            cell.inputs = values
            tracker = None
            if skips:
                self._skip_recv(i, cell)
                tracker = self._skip_tracker(i, cell)

            t0 = self._mark()
            if i < stop:
                cell.checkpointed = True
                cell.cpu_rng = torch.get_rng_state()
                if self.device.type == "cuda":
                    cell.dev_rng = torch.cuda.get_rng_state(self.device)
                with torch.no_grad(), enable_checkpointing(), seed_scope(cell.seed_base):
                    out = self._call(values, tracker)
            else:
                with seed_scope(cell.seed_base):      # same seed derivation in every checkpoint mode
                    out = self._call(values, tracker)
            self._span("fwd", t0, self._mark())
            outs = out if isinstance(out, tuple) else (out,)

            if self.is_last:
                leaves = tuple(o.detach().requires_grad_(o.is_floating_point() and grad_on)
                               if torch.is_tensor(o) else o for o in outs)
                self._out_leaves.append(leaves)
                cell.outputs = outs if not cell.checkpointed else ()
                cell.leaves = leaves
            else:
                # a checkpointed cell ran without a graph; its recompute will produce grad-requiring outputs
                self._send_forward(i, outs, force_grad=cell.checkpointed and grad_on)
                cell.outputs = outs if not cell.checkpointed else ()
            if skips:
                # after the boundary send: the popping rank first waits for its neighbour's activation, which
                # (transitively) waits for the send above — the first step's blocking shape handshake of the
                # skip must not sit in front of it
                self._skip_send(i, tracker, cell, first_time=True)
                if cell.checkpointed:
                    cell.skip_out = {}
            self._cells.append(cell)

        self._drain_sends()
        self._cur.steps += 1
        if not self.is_last:
            return LocalRRef(None)
        gathered = microbatch.gather([Batch(l[0] if len(l) == 1 else l) for l in self._out_leaves])
        return LocalRRef(gathered)

    # -- backward ------------------------------------------------------------------
    def backward(self, loss: Optional[Tensor] = None) -> None:
        """Run the backward half.  On the last stage pass the loss computed from
        ``forward``'s output; elsewhere pass nothing."""
        from ..ops import wgrad          # pure-Python bookkeeping; no GPU needed to import
        with wgrad.deferred():           # one wgrad GEMM per Linear for the whole mini-batch
            self._backward(loss)

    def _backward(self, loss: Optional[Tensor]) -> None:
        cells = self._cells
        self._bwd_seq += 1
        skips = self._has_skips
        if self.is_last:
            if loss is None:
                raise ValueError("the last stage must be given the loss")
            loss.backward()

        for i in reversed(range(len(cells))):
            cell = cells[i]
            if self.is_last:
                grads = [l.grad if torch.is_tensor(l) else None for l in cell.leaves]
                works = []
            elif self._grads_by_mailbox(i):
                grads, works = None, []             # fetched below, right where they are needed
            else:
                grads, works = self._early_grads.pop(i, None) or self._post_recv_grads(i)
                if i > 0 and not self._grads_by_mailbox(i - 1):
                    self._early_grads[i - 1] = self._post_recv_grads(i - 1)   # overlaps this cell's backward

            if cell.checkpointed:
                t0 = self._mark()
                outs = self._recompute(i, cell)     # overlaps the gradient transfer
                self._span("rec", t0, self._mark())
            else:
                outs = cell.outputs
            for w in works:
                w.wait()
            if grads is None:
                grads = self._recv_grads_mailbox(i)  # one kernel that waits for the peer's flag on this stream

            pairs = [(o, g) for o, g in zip(outs, grads)
                     if torch.is_tensor(o) and o.requires_grad and g is not None]
            if skips:
                pairs += self._skip_grads_recv(i, cell)
            t0 = self._mark()
            if pairs:
                torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
            self._span("bwd", t0, self._mark())
            if not self.is_first:
                self._send_grads(i, cell.inputs)
            if skips:
                self._skip_grads_send(i, cell)
            cells[i] = _Cell()                      # drop the graph as early as possible
            for hook in self.cell_backward_hooks:
                hook(i)
        self._drain_sends()
        self._cells = []
        self._out_leaves = []

    def _recompute(self, i: int, cell: _Cell) -> Tuple[Any, ...]:
        devices = [self.device] if self.device.type == "cuda" else []
        tracker = self._skip_tracker(i, cell) if self._has_skips else None
        with torch.random.fork_rng(devices):
            torch.set_rng_state(cell.cpu_rng)
            if cell.dev_rng is not None:
                torch.cuda.set_rng_state(cell.dev_rng, self.device)
            with torch.enable_grad(), enable_recomputing(), seed_scope(cell.seed_base):
                out = self._call(cell.inputs, tracker)
        if tracker is not None:
            self._skip_send(i, tracker, cell, first_time=False)
        return out if isinstance(out, tuple) else (out,)
