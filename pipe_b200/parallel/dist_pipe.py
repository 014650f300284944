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

Each direction of each stage boundary has its own communicator, so a rank's
pending receive never queues ahead of its own sends on one NCCL stream (that
would serialise neighbours and stretch the bubble).  The gradient receive for
micro-batch ``i`` is posted before that micro-batch is recomputed, so recompute
overlaps the transfer — the property the reference gets from its two-node
Checkpoint/Recompute split.

With ``transport="mailbox"`` (or ``PB2_DIST_TRANSPORT=mailbox``) the process group only carries the
first step (it doubles as the shape handshake); from the second step on every boundary talks through an
*NVLink mailbox* (``parallel/mailbox.py``, ``csrc/mailbox.cu``): the sender's kernel stores the activation
straight into the receiver's memory and raises a flag, the receiver's kernel waits for the flag on its
own stream — no NCCL call, no host wait, no early posting.  The default is ``transport="nccl"``.

Works with ``gloo`` on CPU (tests) and ``nccl`` on GPUs.
"""
from __future__ import annotations

import os

from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor, nn

from .. import microbatch
from ..autograd.checkpoint import enable_checkpointing, enable_recomputing
from ..autograd.seeds import SeedBase, seed_scope
from ..microbatch import Batch
from ..rref import LocalRRef
from ..runtime.schedule import bubble_fraction

__all__ = ["DistPipe"]

_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32,
           torch.int16, torch.int8, torch.uint8, torch.bool]
_META_LEN = 16  # [n_tensors] then per tensor: dtype code, requires_grad, ndim, dims...


class _Spec:
    """Shape / dtype / grad flag of one tensor crossing a boundary."""

    __slots__ = ("shape", "dtype", "requires_grad")

    def __init__(self, shape: Sequence[int], dtype: torch.dtype, requires_grad: bool):
        self.shape = tuple(shape)
        self.dtype = dtype
        self.requires_grad = requires_grad

    @classmethod
    def of(cls, t: Tensor) -> "_Spec":
        return cls(t.shape, t.dtype, t.requires_grad and t.is_floating_point())

    def __eq__(self, other: object) -> bool:
        return (isinstance(other, _Spec) and self.shape == other.shape and self.dtype == other.dtype
                and self.requires_grad == other.requires_grad)


class _Cell:
    """What one micro-batch leaves behind on this stage between forward and backward."""

    __slots__ = ("inputs", "outputs", "checkpointed", "cpu_rng", "dev_rng", "leaves", "seed_base")

    def __init__(self) -> None:
        self.inputs: Tuple[Any, ...] = ()
        self.outputs: Tuple[Any, ...] = ()
        self.checkpointed = False
        self.cpu_rng: Optional[Tensor] = None
        self.dev_rng: Optional[Tensor] = None
        self.leaves: Tuple[Any, ...] = ()
        self.seed_base = SeedBase()


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
    """

    def __init__(self, stage: nn.Module, chunks: int = 1, checkpoint: str = "except_last",
                 group: Optional[dist.ProcessGroup] = None, device: Optional[torch.device] = None,
                 strict_except_last: bool = False, transport: str = "nccl",
                 pipelines: Optional[Sequence[Sequence[int]]] = None) -> None:
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
        self.stage = stage
        self.chunks = chunks
        self.checkpoint = checkpoint
        self.strict_except_last = strict_except_last
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

        self._recv_specs: Dict[int, List[_Spec]] = {}     # micro-batch -> specs of tensors from prev stage
        self._sent_specs: Dict[int, List[_Spec]] = {}
        self._cells: List[_Cell] = []
        self._out_leaves: List[Tuple[Any, ...]] = []
        self._sends: List[Any] = []
        self._early: Dict[int, Any] = {}                  # receives posted ahead of their micro-batch
        self._early_grads: Dict[int, Any] = {}
        self.last_bubble = 0.0

        # Boundary transport: "nccl" (default: process-group send / recv, receives pre-posted on their own
        # communicators), "mailbox" / "auto" (NVLink mailboxes, parallel/mailbox.py: the first step still goes
        # through the process group because it carries the shape handshake).  PB2_DIST_TRANSPORT overrides.
        # Measured (config 3, fused): N = 2 20.5 -> 20.1 ms and N = 4 13.06 -> 13.02 ms with mailboxes on
        # the same box, but N = 8 10.2 ms against 9.7 ms for NCCL on another box — not conclusive, so the
        # transport that produced the published scaling table stays the default.
        self.transport = os.environ.get("PB2_DIST_TRANSPORT", transport)
        self._mb: Optional[Dict[str, Any]] = None         # {"fwd_in", "fwd_out", "bwd_in", "bwd_out"} once built
        self._mb_failed = False
        self._steps = 0
        self._fwd_seq = 0
        self._bwd_seq = 0

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
        """Forget cached boundary shapes (call on every rank when the input
        shape changes)."""
        self._recv_specs.clear()
        self._sent_specs.clear()
        self._mb = None            # back to the process group until the new shapes are known
        self._mb_failed = False
        self._steps = 0

    # -- NVLink mailboxes ----------------------------------------------------------
    def _mailboxes_wanted(self) -> bool:
        return (self.transport in ("auto", "mailbox") and self.device.type == "cuda" and self._mb is None
                and not self._mb_failed and self._steps >= 1 and dist.get_backend() == "nccl")

    def _build_mailboxes(self) -> None:
        """Collective over the pipeline: every rank builds the mailboxes of its one or two boundaries,
        boundary by boundary in stage order (so the pairwise rendezvous cannot cycle), and all ranks
        then agree on whether everybody succeeded."""
        from .mailbox import PeerMailbox, mailbox_supported
        mb: Dict[str, Any] = {"fwd_in": None, "fwd_out": None, "bwd_in": None, "bwd_out": None}
        ok = True
        try:
            for b in range(self.n_stages - 1):
                if self.index not in (b, b + 1):
                    continue
                sending = self.index == b
                specs = self._sent_specs if sending else self._recv_specs
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
            self._fwd_seq = 0
            self._bwd_seq = 0
        else:
            self._mb_failed = True
            if self.transport == "mailbox":
                raise RuntimeError("transport='mailbox' requested but not available: "
                                   + getattr(self, "_mb_error", "a peer could not build its mailbox"))

    def _mb_slot(self, box, i: int, specs: Optional[List[_Spec]]) -> bool:
        return box is not None and specs is not None and (i, 0) in box.nbytes

    # -- boundary transfers ------------------------------------------------------
    def _send_meta(self, specs: List[_Spec], dst: int, group) -> None:
        meta = [len(specs)]
        for s in specs:
            meta += [_DTYPES.index(s.dtype), int(s.requires_grad), len(s.shape), *s.shape]
        if len(meta) > _META_LEN * 8:
            raise ValueError("too many / too high-rank tensors cross the stage boundary")
        meta += [0] * (_META_LEN * 8 - len(meta))
        dist.send(torch.tensor(meta, dtype=torch.int64, device=self.device), dst, group=group)

    def _recv_meta(self, src: int, group) -> List[_Spec]:
        buf = torch.empty(_META_LEN * 8, dtype=torch.int64, device=self.device)
        dist.recv(buf, src, group=group)
        meta = buf.tolist()
        specs, at = [], 1
        for _ in range(meta[0]):
            code, rg, nd = meta[at:at + 3]
            shape = meta[at + 3:at + 3 + nd]
            specs.append(_Spec(shape, _DTYPES[code], bool(rg)))
            at += 3 + nd
        return specs

    def _send_forward(self, i: int, values: Tuple[Any, ...]) -> None:
        tensors = [v for v in values if torch.is_tensor(v)]
        if len(tensors) != len(values):
            raise TypeError("only tensors can cross a DistPipe stage boundary")
        specs = [_Spec.of(t) for t in tensors]
        group = self._fwd_groups[self.index]
        known = self._sent_specs.get(i)
        if known is None:
            self._send_meta(specs, self._next(), group)
            self._sent_specs[i] = specs
        elif known != specs:
            raise RuntimeError("stage output shapes changed between steps; call DistPipe.reset() on every rank")
        box = self._mb["fwd_out"] if self._mb else None
        if known is not None and self._mb_slot(box, i, known):
            for k, t in enumerate(tensors):
                box.send(i, k, t.detach().contiguous(), self._fwd_seq)
            return
        for t in tensors:
            self._sends.append(dist.isend(t.detach().contiguous(), self._next(), group=group))

    def _post_recv_forward(self, i: int) -> None:
        """Queue the receives of micro-batch ``i`` now (its shapes must be known), so the transfer
        is already in flight — or done — when the stage gets to it."""
        if i in self._early or i not in self._recv_specs:
            return
        if self._mb and self._mb_slot(self._mb["fwd_in"], i, self._recv_specs.get(i)):
            return              # the sender pushes into the mailbox on its own; nothing to post
        group = self._fwd_groups[self.index - 1]
        out, works = [], []
        for s in self._recv_specs[i]:
            t = torch.empty(s.shape, dtype=s.dtype, device=self.device)
            works.append(dist.irecv(t, self._prev(), group=group))
            out.append(t)
        self._early[i] = (out, works)

    def _recv_forward(self, i: int) -> Tuple[Tensor, ...]:
        group = self._fwd_groups[self.index - 1]
        specs = self._recv_specs.get(i)
        box = self._mb["fwd_in"] if self._mb else None
        if self._mb_slot(box, i, specs):
            out = [torch.empty(s.shape, dtype=s.dtype, device=self.device) for s in specs]
            for k, t in enumerate(out):
                box.recv(i, k, t, self._fwd_seq)
            return tuple(t.requires_grad_(s.requires_grad) for t, s in zip(out, specs))
        if specs is None:
            specs = self._recv_specs[i] = self._recv_meta(self._prev(), group)
        if i in self._early:
            out, works = self._early.pop(i)
        else:
            out, works = [], []
            for s in specs:
                t = torch.empty(s.shape, dtype=s.dtype, device=self.device)
                works.append(dist.irecv(t, self._prev(), group=group))
                out.append(t)
        for w in works:
            w.wait()
        return tuple(t.requires_grad_(s.requires_grad) for t, s in zip(out, specs))

    def _grads_by_mailbox(self, i: int) -> bool:
        return bool(self._mb) and self._mb_slot(self._mb["bwd_in"], i, self._sent_specs.get(i))

    def _recv_grads_mailbox(self, i: int):
        box = self._mb["bwd_in"]
        bufs, k = [], 0
        for s in self._sent_specs[i]:
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
        bufs, works = [], []
        for s in self._sent_specs[i]:
            if not s.requires_grad:
                bufs.append(None)
                continue
            g = torch.empty(s.shape, dtype=s.dtype, device=self.device)
            works.append(dist.irecv(g, self._next(), group=group))
            bufs.append(g)
        return bufs, works

    def _send_grads(self, i: int, inputs: Tuple[Any, ...]) -> None:
        group = self._bwd_groups[self.index - 1]
        box = self._mb["bwd_out"] if self._mb else None
        by_mailbox = self._mb_slot(box, i, self._recv_specs.get(i))
        k = 0
        for t, s in zip(inputs, self._recv_specs[i]):
            if not s.requires_grad:
                continue
            g = t.grad if t.grad is not None else torch.zeros_like(t)
            if by_mailbox:
                box.send(i, k, g.contiguous(), self._bwd_seq)
                k += 1
            else:
                self._sends.append(dist.isend(g.contiguous(), self._prev(), group=group))

    def _drain_sends(self) -> None:
        for w in self._sends:
            w.wait()
        self._sends = []

    # -- forward -------------------------------------------------------------------
    def _checkpoint_stop(self, m: int) -> int:
        if not self.stage.training or not torch.is_grad_enabled():
            return 0
        basis = m if self.strict_except_last else self.chunks
        return {"always": basis, "except_last": basis - 1, "never": 0}[self.checkpoint]

    def _call(self, inputs: Tuple[Any, ...]):
        return self.stage(*inputs)

    def forward(self, *inputs: Any, num_microbatches: Optional[int] = None):
        """Run the forward half of a step.  The first stage takes the mini-batch
        (split along dim 0); later stages take no inputs but need the micro-batch
        count (``num_microbatches``; default: as many as ``Tensor.chunk`` yields
        is only knowable on the first stage, so it is broadcast once and cached)."""
        if self.is_first:
            microbatch.check(self.device, *inputs)
            batches = microbatch.scatter(*inputs, chunks=self.chunks)
            m = len(batches)
        else:
            batches = []
            m = -1
        m = self._agree_microbatches(m, num_microbatches)
        if self._mailboxes_wanted():
            self._build_mailboxes()
        self._fwd_seq += 1

        stop = self._checkpoint_stop(m)
        self._cells = []
        self._out_leaves = []
        self.last_bubble = bubble_fraction(m, self.n_stages)

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

            if i < stop:
                cell.checkpointed = True
                cell.cpu_rng = torch.get_rng_state()
                if self.device.type == "cuda":
                    cell.dev_rng = torch.cuda.get_rng_state(self.device)
                with torch.no_grad(), enable_checkpointing(), seed_scope(cell.seed_base):
                    out = self._call(values)
            else:
                out = self._call(values)
            outs = out if isinstance(out, tuple) else (out,)

            if self.is_last:
                leaves = tuple(o.detach().requires_grad_(o.is_floating_point() and torch.is_grad_enabled())
                               if torch.is_tensor(o) else o for o in outs)
                self._out_leaves.append(leaves)
                cell.outputs = outs if not cell.checkpointed else ()
                cell.leaves = leaves
            else:
                if cell.checkpointed:
                    # the recompute will produce grad-requiring outputs
                    for o in outs:
                        if torch.is_tensor(o) and o.is_floating_point():
                            o.requires_grad_(True)
                self._send_forward(i, outs)
                cell.outputs = outs if not cell.checkpointed else ()
            self._cells.append(cell)

        self._drain_sends()
        self._steps += 1
        if not self.is_last:
            return LocalRRef(None)
        gathered = microbatch.gather([Batch(l[0] if len(l) == 1 else l) for l in self._out_leaves])
        return LocalRRef(gathered)

    _m_cache: Optional[int] = None

    def _agree_microbatches(self, m: int, hint: Optional[int]) -> int:
        if hint is not None:
            return hint
        if self._m_cache is not None and (m == -1 or m == self._m_cache):
            return self._m_cache
        t = torch.tensor([m], dtype=torch.int64, device=self.device)
        dist.broadcast(t, self.ranks[0], group=self.group)
        self._m_cache = int(t.item())
        return self._m_cache

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
                outs = self._recompute(cell)        # overlaps the gradient transfer
            else:
                outs = cell.outputs
            for w in works:
                w.wait()
            if grads is None:
                grads = self._recv_grads_mailbox(i)  # one kernel that waits for the peer's flag on this stream

            pairs = [(o, g) for o, g in zip(outs, grads)
                     if torch.is_tensor(o) and o.requires_grad and g is not None]
            if pairs:
                torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
            if not self.is_first:
                self._send_grads(i, cell.inputs)
            cells[i] = _Cell()                      # drop the graph as early as possible
        self._drain_sends()
        self._cells = []
        self._out_leaves = []

    def _recompute(self, cell: _Cell) -> Tuple[Any, ...]:
        devices = [self.device] if self.device.type == "cuda" else []
        with torch.random.fork_rng(devices):
            torch.set_rng_state(cell.cpu_rng)
            if cell.dev_rng is not None:
                torch.cuda.set_rng_state(cell.dev_rng, self.device)
            with torch.enable_grad(), enable_recomputing(), seed_scope(cell.seed_base):
                out = self._call(cell.inputs)
        return out if isinstance(out, tuple) else (out,)
