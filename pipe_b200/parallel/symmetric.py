"""Symmetric gradient buffers: memory that every rank of a group can address.

A :class:`SymmetricBuffer` is one fp32 allocation per rank, of identical size, whose peer copies
are mapped into this process (and, where the NVSwitch supports it, aliased by one multicast
address).  The fused all-reduce kernel (``csrc/allreduce.cu``) takes the peer pointers and does the
collective with loads and stores; nothing else is needed at run time.

Allocation and handle exchange go through ``torch.distributed._symmetric_memory`` (cuMem virtual
memory + fabric / fd handles, multicast object).  If that path is unavailable the buffer falls back
to classic CUDA IPC handles exchanged with ``all_gather_object`` (no multicast, so the two-shot P2P
kernel is used).  On CPU / gloo — the host-logic tests — the "buffer" is a plain tensor and
:meth:`all_reduce_` falls back to ``dist.all_reduce``.

Reference counterpart: none in code — the reference only remarks that DDP may wrap ``Pipe`` with
``checkpoint='never'`` (``pipe.py:290-293``); SURVEY.md K18, W11, §5.8.

Layout: ``[data: n fp32][signal pad: pb2_allreduce_pad_bytes()]``; ``n`` is rounded up to a
multiple of ``4 * world`` elements (the kernel's slice granularity).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist
from torch import Tensor

__all__ = ["SymmetricBuffer"]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class SymmetricBuffer:
    def __init__(self, numel: int, group: Optional[dist.ProcessGroup], device: torch.device) -> None:
        self.group = group
        self.device = torch.device(device)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.numel = _round_up(max(numel, 1), 4 * self.world)
        self.epoch = 0
        self.mode = "torch"            # "symm" | "ipc" | "torch"
        self.multicast_ptr = 0
        self._keep: List[object] = []
        self.peer_data: List[int] = []
        self.peer_pads: List[int] = []

        if self.device.type != "cuda" or self.world == 1:
            self.data = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            return

        from ..ops import _lib
        lib = _lib.load()
        lib.pb2_allreduce_pad_bytes.restype = ctypes.c_size_t
        self.pad_bytes = int(lib.pb2_allreduce_pad_bytes())
        total_bytes = self.numel * 4 + self.pad_bytes
        if not self._try_symm_mem(total_bytes):
            self._use_ipc(total_bytes)

    # -- allocation paths ------------------------------------------------------------
    def _try_symm_mem(self, total_bytes: int) -> bool:
        try:
            import torch.distributed._symmetric_memory as symm_mem
            group = self.group if self.group is not None else dist.group.WORLD
            raw = symm_mem.empty(total_bytes, dtype=torch.uint8, device=self.device)
            hdl = symm_mem.rendezvous(raw, group)
            raw.zero_()
            self._keep += [raw, hdl]
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            self.peer_data = ptrs
            self.peer_pads = [p + self.numel * 4 for p in ptrs]
            self.data = raw[: self.numel * 4].view(torch.float32)
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self.multicast_ptr = mc
            self.mode = "symm"
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
            return True
        except Exception as e:  # noqa: BLE001 - any failure means "use the IPC path"
            self._symm_error = f"{type(e).__name__}: {e}"
            return False

    def _use_ipc(self, total_bytes: int) -> None:
        """cudaIpc handles of a plain allocation, exchanged through the process group."""
        raw = torch.zeros(total_bytes, dtype=torch.uint8, device=self.device)
        handle = raw.untyped_storage()._share_cuda_()
        gathered: List[object] = [None] * self.world
        dist.all_gather_object(gathered, handle, group=self.group)
        self._keep.append(raw)
        ptrs: List[int] = []
        for q, h in enumerate(gathered):
            if q == self.rank:
                ptrs.append(raw.data_ptr())
                continue
            storage = torch.UntypedStorage._new_shared_cuda(*h)
            peer = torch.empty(0, dtype=torch.uint8, device=self.device).set_(storage)
            self._keep.append(peer)
            ptrs.append(peer.data_ptr())
        self.peer_data = ptrs
        self.peer_pads = [p + self.numel * 4 for p in ptrs]
        self.data = raw[: self.numel * 4].view(torch.float32)
        self.mode = "ipc"
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    # -- the collective ----------------------------------------------------------------
    def all_reduce_(self, scale: float = 1.0, sumsq: Optional[Tensor] = None, cast_out: Optional[Tensor] = None,
                    ctas: int = 64, use_multimem: Optional[bool] = None, stream: Optional[int] = None,
                    lo: int = 0, hi: Optional[int] = None) -> None:
        """In-place ``data[lo:hi] <- scale * sum over ranks(data[lo:hi])`` with the fused kernel; every rank
        must call it at the same point of its stream.  ``lo`` and ``hi - lo`` are multiples of ``4 * world``
        (a *bucket*: successive calls on one stream may cover different ranges of the same buffer)."""
        hi = self.numel if hi is None else hi
        n = hi - lo
        if lo % (4 * self.world) or n % (4 * self.world) or n < 0 or hi > self.numel:
            raise ValueError(f"bucket [{lo}, {hi}) is not aligned to 4 * world = {4 * self.world} elements")
        if n == 0:
            return
        if self.mode == "torch":
            part = self.data[lo:hi]
            if self.world > 1:
                dist.all_reduce(part, group=self.group)
            if scale != 1.0:
                part.mul_(scale)
            if sumsq is not None:
                # every rank holds the full result: count each slice once, like the kernel does
                a = n // self.world * self.rank
                sumsq.add_(part[a:a + n // self.world].float().pow(2).sum())
            if cast_out is not None:
                cast_out.copy_(part)
            return

        from ..ops import _lib
        lib = _lib.load()
        self.epoch += 1
        if use_multimem is None:
            # in-switch reduction pays off once several peers contribute; with two, P2P pull/push wins
            use_multimem = bool(self.multicast_ptr) and self.world >= 4
        arr = ctypes.c_void_p * self.world
        bufs = arr(*[p + lo * 4 for p in self.peer_data])
        pads = arr(*self.peer_pads)
        if stream is None:
            stream = _lib.raw_stream(self.device)
        with _lib.on_device(self.device):
            mc = (self.multicast_ptr + lo * 4) if self.multicast_ptr else None
            rc = lib.pb2_allreduce_f32(bufs, pads, ctypes.c_void_p(mc), ctypes.c_int(self.world),
                                       ctypes.c_int(self.rank), ctypes.c_size_t(n), ctypes.c_float(scale),
                                       ctypes.c_void_p(_lib.ptr(sumsq)), ctypes.c_void_p(_lib.ptr(cast_out)),
                                       ctypes.c_uint(self.epoch), ctypes.c_int(ctas), ctypes.c_int(int(use_multimem)),
                                       ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_allreduce_f32")

    def barrier(self, phase: int = 0, stream: Optional[int] = None) -> None:
        """A device-side barrier among the ranks of the group on ``stream`` (one tiny kernel, no host wait):
        what this rank's earlier kernels stored into peers' copies is visible to the kernels a peer launches
        after its own call.  Uses this buffer's signal pad: do not mix with :meth:`all_reduce_` on one buffer."""
        if self.mode == "torch":
            if self.world > 1 and self.device.type != "cuda":
                dist.barrier(group=self.group)
            return
        from ..ops import _lib
        lib = _lib.load()
        self.epoch += 1
        arr = ctypes.c_void_p * self.world
        pads = arr(*self.peer_pads)
        if stream is None:
            stream = _lib.raw_stream(self.device)
        with _lib.on_device(self.device):
            rc = lib.pb2_group_barrier(pads, ctypes.c_int(self.world), ctypes.c_int(self.rank), ctypes.c_uint(self.epoch),
                                       ctypes.c_int(phase), ctypes.c_void_p(stream))
        _lib.check(rc, "pb2_group_barrier")
