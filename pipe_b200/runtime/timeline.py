"""Device-side timeline of pipeline cells -> measured bubble fraction.

When a :class:`Timeline` is attached to a ``Pipe`` (``pipe.timeline = Timeline()``)
the executor brackets every forward cell with CUDA events on the stage's compute
stream, and tensor hooks bracket the cell's backward (the hook on the cell's
output fires before its backward nodes run, the hook on its input after).  After
a step, :meth:`Timeline.summary` turns the events into per-stage busy time, the
step span, and the measured bubble ``1 - sum(busy) / (n * span)`` for forward
and backward, to be reported next to tokens/s (BASELINE.json ``metric``).

On CPU stages host timestamps are used instead of events.

Parity: the reference's only hook is ``record_function("chunk%d-part%d")``
(``pipeline.py:206-209,226-229``, removed by the author); SURVEY.md §5.1 asks for
the cell timeline and bubble % as additions.
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

__all__ = ["Timeline", "Span"]


class _Mark:
    """A point in time on a device stream (CUDA event) or on the host."""

    __slots__ = ("event", "host")

    def __init__(self, device: torch.device) -> None:
        self.host = time.perf_counter()
        self.event: Optional[torch.cuda.Event] = None
        if device.type == "cuda":
            self.event = torch.cuda.Event(enable_timing=True)
            self.event.record(torch.cuda.current_stream(device))


@dataclass
class Span:
    phase: str            # "fwd" | "bwd"
    i: int                # micro-batch
    j: int                # stage
    start_ms: float
    end_ms: float


class Timeline:
    def __init__(self) -> None:
        self._lock = threading.Lock()
        self._marks: Dict[Tuple[str, int, int, str], _Mark] = {}
        self._origin: Dict[int, _Mark] = {}   # per stage: reference mark
        self.enabled = True

    # -- recording ---------------------------------------------------------
    def reset(self) -> None:
        with self._lock:
            self._marks.clear()
            self._origin.clear()

    def origin(self, j: int, device: torch.device) -> None:
        """Record the per-stage time origin (call once per step, after a
        cross-device barrier so the origins line up)."""
        with self._lock:
            self._origin[j] = _Mark(device)

    def mark(self, phase: str, i: int, j: int, edge: str, device: torch.device) -> None:
        if not self.enabled:
            return
        m = _Mark(device)
        with self._lock:
            self._marks[(phase, i, j, edge)] = m

    def hook_backward(self, i: int, j: int, tensor_in: Optional[torch.Tensor],
                      tensor_out: torch.Tensor, device: torch.device) -> None:
        """Bracket the backward of cell ``(i, j)`` with tensor hooks."""
        if not self.enabled:
            return
        if tensor_out.requires_grad:
            tensor_out.register_hook(lambda g: self.mark("bwd", i, j, "start", device))
        if tensor_in is not None and tensor_in.requires_grad:
            tensor_in.register_hook(lambda g: self.mark("bwd", i, j, "end", device))

    # -- analysis ------------------------------------------------------------
    def spans(self) -> List[Span]:
        """Resolve marks to milliseconds relative to each stage's origin.
        Synchronises every CUDA device that has marks."""
        out: List[Span] = []
        with self._lock:
            marks = dict(self._marks)
            origin = dict(self._origin)
        if torch.cuda.is_available():
            torch.cuda.synchronize()

        def at(j: int, m: _Mark) -> float:
            o = origin.get(j)
            if m.event is not None and o is not None and o.event is not None:
                return o.event.elapsed_time(m.event)
            base = o.host if o is not None else 0.0
            return (m.host - base) * 1e3

        keys = {(p, i, j) for (p, i, j, _e) in marks}
        for p, i, j in sorted(keys):
            s = marks.get((p, i, j, "start"))
            e = marks.get((p, i, j, "end"))
            if s is None or e is None:
                continue
            out.append(Span(p, i, j, at(j, s), at(j, e)))
        return out

    def summary(self, n_stages: int) -> Dict[str, float]:
        spans = self.spans()
        res: Dict[str, float] = {}
        for phase in ("fwd", "bwd"):
            sel = [s for s in spans if s.phase == phase]
            if not sel:
                continue
            t0 = min(s.start_ms for s in sel)
            t1 = max(s.end_ms for s in sel)
            span = max(t1 - t0, 1e-9)
            busy = sum(s.end_ms - s.start_ms for s in sel)
            res[f"{phase}_span_ms"] = span
            res[f"{phase}_busy_ms"] = busy
            res[f"{phase}_bubble"] = max(0.0, 1.0 - busy / (n_stages * span))
        return res
