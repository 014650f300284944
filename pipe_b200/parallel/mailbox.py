"""NVLink mailboxes: the stage-to-stage transport of :class:`DistPipe` without NCCL.

One :class:`PeerMailbox` carries the tensors that cross one stage boundary in one direction.  Its
memory is a symmetric allocation of the two ranks on either side of the boundary
(:class:`~pipe_b200.parallel.symmetric.SymmetricBuffer`: ``torch.distributed._symmetric_memory`` or
CUDA IPC), laid out identically on both::

    [slot 0: tensor 0 | tensor 1 | ...][slot 1: ...] ... [flags: slots x tensors x 32 u32][acks: same]

The receiver's copy holds payloads and flags, the sender's copy holds the acknowledgements.  A slot is
a micro-batch index, the sequence number is the step count, so nothing is ever reset.  ``send`` and
``recv`` each launch one small kernel (``csrc/mailbox.cu``) on the caller's current stream; there is
no host-side wait anywhere.

Reference counterpart: ``Copy`` / ``Wait`` on copy streams inside one process (README.md:185-237,
332-369); SURVEY.md K12 / W12 names NCCL p2p as the multi-process baseline, which
``DistPipe(transport="nccl")`` still is.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from .symmetric import SymmetricBuffer

__all__ = ["PeerMailbox", "mailbox_supported"]

_MAX_CTAS = 32          # PB2_MAILBOX_MAX_CTAS


def _round16(n: int) -> int:
    return (n + 15) // 16 * 16


def mailbox_supported(device: torch.device, shapes: Sequence[Tuple[Sequence[int], torch.dtype]]) -> bool:
    """CUDA device, and every message a multiple of 16 bytes (the kernels move 16-byte vectors)."""
    if torch.device(device).type != "cuda" or dist.get_backend() != "nccl":
        return False
    for shape, dtype in shapes:
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        if nbytes == 0 or nbytes % 16:
            return False
    return True


def _ctas_for(nbytes: int) -> int:
    """Both sides derive the grid from the message size: ~64 KB per CTA, 1..16 CTAs."""
    return max(1, min(16, nbytes // (64 * 1024)))


class PeerMailbox:
    """One direction of one stage boundary.

    Args:
        group: the two-rank process group of the boundary (sender and receiver).
        device: this rank's device.
        sender_rank: global rank of the sending side.
        layout: ``{slot: [(shape, dtype), ...]}`` — identical on both ranks.
    """

    def __init__(self, group: dist.ProcessGroup, device: torch.device, sender_rank: int,
                 layout: Dict[int, List[Tuple[Tuple[int, ...], torch.dtype]]]) -> None:
        self.device = torch.device(device)
        self.is_sender = dist.get_rank() == sender_rank
        self.n_slots = (max(layout) + 1) if layout else 0
        self.n_tensors = max((len(v) for v in layout.values()), default=0)
        self.nbytes: Dict[Tuple[int, int], int] = {}
        # per-tensor capacity = the largest message of that position over all slots
        cap = [0] * self.n_tensors
        for slot, tensors in layout.items():
            for k, (shape, dtype) in enumerate(tensors):
                n = 1
                for s in shape:
                    n *= int(s)
                b = n * torch.empty(0, dtype=dtype).element_size()
                if b == 0 or b % 16:
                    raise ValueError(f"mailbox messages must be positive multiples of 16 bytes, got {b}")
                self.nbytes[(slot, k)] = b
                cap[k] = max(cap[k], _round16(b))
        self.tensor_off = [sum(cap[:k]) for k in range(self.n_tensors)]
        self.slot_bytes = _round16(sum(cap))
        data_bytes = self.n_slots * self.slot_bytes
        words = self.n_slots * max(self.n_tensors, 1) * _MAX_CTAS
        self.flags_off = _round16(data_bytes)
        self.acks_off = self.flags_off + words * 4
        total = self.acks_off + words * 4
        self.buf = SymmetricBuffer((total + 3) // 4, group, self.device)
        if self.buf.mode == "torch":
            raise RuntimeError("peer-mapped memory is not available for this boundary")
        me = self.buf.rank
        self.local = self.buf.peer_data[me]
        self.remote = self.buf.peer_data[1 - me]

    def _words(self, slot: int, k: int) -> int:
        return (slot * max(self.n_tensors, 1) + k) * _MAX_CTAS * 4

    def send(self, slot: int, k: int, src: Tensor, seq: int) -> None:
        from ..ops import _lib
        nbytes = self.nbytes[(slot, k)]
        assert self.is_sender and src.is_contiguous() and src.numel() * src.element_size() == nbytes
        off = slot * self.slot_bytes + self.tensor_off[k]
        with _lib.on_device(self.device):
            rc = _lib.load().pb2_mailbox_send(
                ctypes.c_void_p(src.data_ptr()), ctypes.c_size_t(nbytes), ctypes.c_void_p(self.remote + off),
                ctypes.c_void_p(self.remote + self.flags_off + self._words(slot, k)),
                ctypes.c_void_p(self.local + self.acks_off + self._words(slot, k)),
                ctypes.c_uint(seq), ctypes.c_int(slot), ctypes.c_int(_ctas_for(nbytes)),
                ctypes.c_void_p(_lib.raw_stream(self.device)))
        _lib.check(rc, "pb2_mailbox_send")

    def recv(self, slot: int, k: int, dst: Tensor, seq: int) -> None:
        from ..ops import _lib
        nbytes = self.nbytes[(slot, k)]
        assert (not self.is_sender) and dst.is_contiguous() and dst.numel() * dst.element_size() == nbytes
        off = slot * self.slot_bytes + self.tensor_off[k]
        with _lib.on_device(self.device):
            rc = _lib.load().pb2_mailbox_recv(
                ctypes.c_void_p(dst.data_ptr()), ctypes.c_size_t(nbytes), ctypes.c_void_p(self.local + off),
                ctypes.c_void_p(self.local + self.flags_off + self._words(slot, k)),
                ctypes.c_void_p(self.remote + self.acks_off + self._words(slot, k)),
                ctypes.c_uint(seq), ctypes.c_int(slot), ctypes.c_int(_ctas_for(nbytes)),
                ctypes.c_void_p(_lib.raw_stream(self.device)))
        _lib.check(rc, "pb2_mailbox_recv")
