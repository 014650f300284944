"""Micro-batch container and the scatter / gather pair.

A :class:`Batch` is what flows between stages: either exactly one tensor
("atomic") or a tuple of arbitrary values with at least one tensor in it.
``scatter`` splits a mini-batch along dim 0 into at most ``chunks`` batches
(``Tensor.chunk`` semantics: fewer when the batch does not divide, e.g.
20 rows / 8 chunks -> 7 batches of 3,3,3,3,3,3,2), ``gather`` concatenates
them back.

Parity: reference ``pipe.py:446-464,476-477,484,490``, ``pipeline.py:44-54,229``,
``README.md:316-322,371-382``; SURVEY.md A1 / Appendix A "microbatch".
"""
from __future__ import annotations

import threading

from typing import Any, Callable, Iterator, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

__all__ = ["NoChunk", "Batch", "check", "scatter", "gather", "GatherBuffer", "CellContext", "use_cell", "current_cell"]

Tensors = Sequence[Tensor]
TensorOrTensors = Union[Tensor, Tensors]
Function = Callable[..., Any]


class NoChunk:
    """Marks a tensor input of ``Pipe.forward`` that must be handed whole to
    every micro-batch instead of being split along dim 0."""

    def __init__(self, inp: Tensor):
        if not torch.is_tensor(inp):
            raise TypeError(f"NoChunk only supported for tensors, found: {inp}")
        self._tensor = inp

    @property
    def tensor(self) -> Tensor:
        return self._tensor


class Batch:
    """One micro-batch: a tensor, or a tuple of values holding >= 1 tensor."""

    __slots__ = ("_values", "atomic")

    def __init__(self, values: Union[Tensor, Sequence[Any]]) -> None:
        self._values = values
        self.atomic = torch.is_tensor(values)
        if not self.atomic and not any(torch.is_tensor(v) for v in values):
            raise TypeError(f"No tensors found in batch: {values}")

    # -- accessors ---------------------------------------------------------
    @property
    def tensor(self) -> Tensor:
        if not self.atomic:
            raise AttributeError("not atomic batch")
        return self._values  # type: ignore[return-value]

    @property
    def values(self):
        return self._values

    def find_tensor_idx(self) -> int:
        """Index of the first tensor (0 for an atomic batch)."""
        if self.atomic:
            return 0
        for k, v in enumerate(self._values):
            if torch.is_tensor(v):
                return k
        raise TypeError("No tensor found!")

    def get_device(self) -> torch.device:
        if self.atomic:
            return self._values.device  # type: ignore[union-attr]
        return self._values[self.find_tensor_idx()].device

    def call(self, function: Function) -> "Batch":
        """Apply ``function`` to the payload (splatted when non-atomic)."""
        if self.atomic:
            return Batch(function(self._values))
        return Batch(function(*self._values))

    # -- sequence protocol ---------------------------------------------------
    def __repr__(self) -> str:
        return f"Batch[atomic={self.atomic!r}]({self._values!r})"

    def __len__(self) -> int:
        return 1 if self.atomic else len(self._values)

    def __iter__(self) -> Iterator[Any]:
        if self.atomic:
            yield self._values
        else:
            yield from self._values

    def __getitem__(self, index: int):
        if not self.atomic:
            return self._values[index]
        if index != 0:
            raise IndexError("atomic batch allows index 0 only")
        return self._values

    def __setitem__(self, index: Union[int, slice], value) -> None:
        if isinstance(index, int):
            self._set_one(index, value)
        else:
            self._set_all(index, value)

    def _set_one(self, index: int, value) -> None:
        if self.atomic:
            if index != 0:
                raise IndexError("atomic batch allows index 0 only")
            self._values = value
        else:
            self._values = self._values[:index] + (value,) + self._values[index + 1:]

    def _set_all(self, index: slice, value: Sequence[Any]) -> None:
        if not (index.start is index.stop is index.step is None):
            raise NotImplementedError("only slice [:] supported")
        if self.atomic:
            if len(value) != 1:
                raise IndexError("atomic batch cannot be replaced with multiple tensors")
            self._values = value[0]
        else:
            self._values = tuple(value)


def check(first_device: torch.device, *inputs: Any) -> None:
    """Validate ``Pipe.forward`` inputs: at least one tensor, and every tensor
    already on the first stage's device."""
    if not any(torch.is_tensor(x) for x in inputs):
        raise TypeError(f"inputs do not have any tensors: {inputs}")
    for x in inputs:
        if torch.is_tensor(x) and x.device != first_device:
            raise ValueError("All inputs should be on the same device as the first partition")


def scatter(*inputs: Any, chunks: int) -> List[Batch]:
    """Split ``inputs`` into micro-batches along dim 0.

    Tensors are chunked; ``NoChunk`` payloads and non-tensors are replicated to
    every micro-batch.  All tensors must yield the same number of chunks.
    """
    if len(inputs) == 1 and torch.is_tensor(inputs[0]):
        return [Batch(piece) for piece in inputs[0].chunk(chunks)]

    columns: List[List[Any]] = []
    n_pieces = -1
    for x in inputs:
        if torch.is_tensor(x):
            pieces = list(x.chunk(chunks))
            if n_pieces != -1 and n_pieces != len(pieces):
                raise RuntimeError(
                    f"Found different number of chunks produced for inputs: {n_pieces} and {len(pieces)}")
            n_pieces = len(pieces)
            columns.append(pieces)
        else:
            payload = x.tensor if isinstance(x, NoChunk) else x
            columns.append([payload] * chunks)

    if n_pieces == -1:
        n_pieces = chunks
    return [Batch(tuple(col[i] for col in columns)) for i in range(n_pieces)]


def _adjacent_base(tensors: Sequence[Tensor]) -> Optional[Tensor]:
    """If ``tensors`` are consecutive dim-0 slices of one buffer with identical strides (a last stage that
    wrote its micro-batch outputs straight into a :class:`GatherBuffer`; rows may be padded, i.e. the
    slices need not be contiguous), return the zero-copy view that spans them; else ``None``."""
    first = tensors[0]
    if first.dim() == 0 or first.stride(-1) != 1 and first.numel() > 0:
        return None
    storage = first.untyped_storage()
    strides = first.stride()
    at = first.storage_offset()
    rows = 0
    for t in tensors:
        if (t.dtype != first.dtype or t.device != first.device or t.shape[1:] != first.shape[1:]
                or t.stride() != strides or t.untyped_storage().data_ptr() != storage.data_ptr()
                or t.storage_offset() != at):
            return None
        at += t.size(0) * strides[0]
        rows += t.size(0)
    if at * first.element_size() > storage.nbytes():
        return None
    return torch.empty(0, dtype=first.dtype, device=first.device).set_(
        storage, first.storage_offset(), (rows, *first.shape[1:]), strides)


class _AdjacentCat(torch.autograd.Function):
    """``torch.cat`` of tensors that already sit next to each other in memory: forward is a view,
    backward splits the gradient (the gather copy of the mini-batch output goes away)."""

    @staticmethod
    def forward(ctx, base: Tensor, *tensors: Tensor) -> Tensor:  # type: ignore[override]
        ctx.sizes = [t.size(0) for t in tensors]
        return base.detach()

    @staticmethod
    def backward(ctx, grad: Tensor):  # type: ignore[override]
        return (None,) + tuple(grad.split(ctx.sizes))


def _cat(tensors: Sequence[Tensor]) -> Tensor:
    if len(tensors) > 1:
        base = _adjacent_base(tensors)
        if base is not None:
            return _AdjacentCat.apply(base, *tensors)
    return torch.cat(tuple(tensors))


def gather(outputs: List[Batch]):
    """Concatenate micro-batch outputs back into a mini-batch.

    Atomic outputs give one tensor.  Otherwise a tuple is returned: tensor
    positions are concatenated, non-tensor positions become a Python list with
    one entry per micro-batch.
    """
    if outputs[0].atomic:
        return _cat([b.tensor for b in outputs])

    width = len(outputs[0])
    gathered: List[Any] = []
    for k in range(width):
        kind = type(outputs[0][k])
        column = []
        for b in outputs:
            if type(b[k]) is not kind:
                raise TypeError(f"Types for microbatch outputs do not match, found: {kind} and {type(b[k])}")
            column.append(b[k])
        gathered.append(_cat(column) if torch.is_tensor(outputs[0][k]) else column)
    return tuple(gathered)


class GatherBuffer:
    """Preallocated mini-batch output that micro-batches write into in place.

    ``gather`` costs one extra read+write of the whole output (450 MiB for the
    tutorial logits).  A last-stage module that knows its slice
    can instead write straight into ``buffer[offset:offset+rows]``; ``cat`` then
    has nothing left to do.  The fused vocabulary projection does that for the logits it
    returns (``models/fused.py`` ``_DecoderFn``; it learns its slice from
    :func:`current_cell`); plain ``gather`` remains the path for everything else.
    """

    def __init__(self, rows: int, tail_shape: Sequence[int], *, dtype: torch.dtype, device: torch.device):
        self.buffer = torch.empty((rows, *tail_shape), dtype=dtype, device=device)
        self.offsets: List[Tuple[int, int]] = []

    def plan(self, sizes: Sequence[int]) -> None:
        start = 0
        self.offsets = []
        for s in sizes:
            self.offsets.append((start, start + s))
            start += s

    def slot(self, i: int) -> Tensor:
        """Micro-batch ``i``'s rows as an *alias* of the buffer (same memory, but not an autograd
        view of it: in-place writes into neighbouring slots must not trip view/version tracking)."""
        lo, hi = self.offsets[i]
        b = self.buffer
        row = b[0].numel() if b.size(0) else 0
        return torch.empty(0, dtype=b.dtype, device=b.device).set_(
            b.untyped_storage(), b.storage_offset() + lo * row, (hi - lo, *b.shape[1:]))


# ---------------------------------------------------------------------------------------
# which micro-batch is running: lets a last-stage module write into its slice of the mini-batch output
# ---------------------------------------------------------------------------------------
class CellContext:
    """Set by the pipeline executor around every stage call: micro-batch ``index``, the dim-0 row count of
    every micro-batch of this forward pass (``sizes``) and a dict shared by all cells of the pass."""

    __slots__ = ("index", "sizes", "shared")

    def __init__(self, index: int, sizes: Sequence[int], shared: dict) -> None:
        self.index = index
        self.sizes = list(sizes)
        self.shared = shared


_cell_local = threading.local()


class use_cell:
    def __init__(self, cell: Optional[CellContext]) -> None:
        self.cell = cell
        self.prev: Optional[CellContext] = None

    def __enter__(self):
        self.prev = getattr(_cell_local, "cell", None)
        _cell_local.cell = self.cell
        return self.cell

    def __exit__(self, *exc):
        _cell_local.cell = self.prev
        return False


def current_cell() -> Optional[CellContext]:
    """The running cell's context, or ``None`` outside a pipeline."""
    return getattr(_cell_local, "cell", None)
