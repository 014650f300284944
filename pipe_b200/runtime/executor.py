"""The pipeline executor: fences, task dispatch, result collection.

For every clock tick of :func:`~.schedule.clock_cycles` the executor does two
things on the calling (main) thread:

``fence``    wires the autograd ordering edge ``batch[i-1] <- batch[i]`` (so the
             backward of micro-batch ``i-1`` on a stage runs after ``i``'s),
             moves skip tensors whose route ends at the stage, and copies the
             micro-batch from the previous stage on the ``(stage, micro-batch)``
             copy-stream pair.
``compute``  makes the stage's compute stream wait for that copy, wraps the
             stage call in a :class:`~.workers.Task` (checkpointed or plain) and
             hands it to the stage device's worker thread; then collects results
             in order, makes the copy stream wait for the compute stream, and
             attaches the Recompute node for checkpointed cells.

The first exception raised by any stage in a tick is re-raised — with its
original traceback — after the tick has drained; later ones in the same tick are
dropped.

Backward needs no scheduler: the Copy / Wait / fork-join / Checkpoint nodes
recorded here replay the schedule in reverse inside the autograd engine, whose
per-device threads provide the cross-device parallelism.

Parity: reference ``pipeline.py:43-60`` (helpers), ``82-117`` (``run``),
``119-142`` (``fence``), ``144-266`` (``compute``); SURVEY.md C8-C10.
"""
from __future__ import annotations

import sys
from queue import Queue
from typing import Callable, List, Optional, Sequence, Tuple

import torch
from torch import nn
from torch.autograd.profiler import record_function

from ..autograd.checkpoint import Checkpointing
from ..autograd.copy import Copy, Wait
from ..autograd.dependency import fork, join
from ..autograd.seeds import SeedBase, seed_scope
from ..autograd.stream import AbstractStream, current_stream, use_device
from ..microbatch import Batch, CellContext, use_cell
from ..skip.layout import SkipLayout
from ..skip.tracker import SkipTrackerThroughPotals, use_skip_tracker
from .schedule import Cell, clock_cycles
from .timeline import Timeline
from .workers import ExcInfo, Task, WorkerPool

__all__ = ["Pipeline", "depend", "copy_batch", "wait_batch"]


def depend(fork_from: Batch, join_to: Batch) -> None:
    """Order ``fork_from``'s backward after ``join_to``'s (in place)."""
    a = fork_from.find_tensor_idx()
    b = join_to.find_tensor_idx()
    fork_from[a], phony = fork(fork_from[a])
    join_to[b] = join(join_to[b], phony)


def _detach_non_float(values: Sequence) -> Tuple:
    # Integer / bool tensors cannot carry gradient; cutting them loose keeps the
    # engine from tracking them through Copy / Wait.
    return tuple(v.detach() if torch.is_tensor(v) and not v.is_floating_point() else v
                 for v in values)


def copy_batch(batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream) -> None:
    batch[:] = _detach_non_float(Copy.apply(prev_stream, next_stream, *batch))


def wait_batch(batch: Batch, prev_stream: AbstractStream, next_stream: AbstractStream) -> None:
    batch[:] = _detach_non_float(Wait.apply(prev_stream, next_stream, *batch))


# the reference's private name for the schedule generator (pipeline.py:63-79), kept for code that imports it
_clock_cycles = clock_cycles


class _cell_range:
    """``record_function`` label for one (micro-batch, stage) cell, doubled as an NVTX range on CUDA
    devices so the cells show up in ``ncu`` / Nsight timelines as well as in ``torch.profiler`` traces
    (reference: ``record_function("chunk%d-part%d")``, pipeline.py:206-207, 226-227)."""

    __slots__ = ("label", "nvtx", "rf")

    def __init__(self, label: Optional[str], device: torch.device) -> None:
        self.label = label
        self.nvtx = label is not None and device.type == "cuda"
        self.rf = None

    def __enter__(self):
        if self.label is not None:
            self.rf = record_function(self.label)
            self.rf.__enter__()
            if self.nvtx:
                torch.cuda.nvtx.range_push(self.label)
        return self

    def __exit__(self, *exc):
        if self.label is not None:
            if self.nvtx:
                torch.cuda.nvtx.range_pop()
            self.rf.__exit__(*exc)
        return False


class Pipeline:
    """Runs micro-batches through partitions on their devices, in place."""

    def __init__(self, partitions: List[nn.Sequential], devices: List[torch.device],
                 copy_streams: List[List[AbstractStream]], skip_layout: SkipLayout,
                 checkpoint_stop: int) -> None:
        self.partitions = partitions
        self.devices = devices
        self.copy_streams = copy_streams
        self.skip_layout = skip_layout
        self.checkpoint_stop = checkpoint_stop
        self.pool = WorkerPool(devices)
        self.in_queues: List[Queue] = self.pool.in_queues
        self.out_queues: List[Queue] = self.pool.out_queues
        #: label each cell ``chunk{i}-part{j}`` for torch.profiler / NVTX
        self.trace_cells = False
        self.timeline: Optional[Timeline] = None
        #: called as ``hook(i, j, batch)`` right after cell (i, j)'s forward was
        #: collected on the main thread (the DDP hybrid hangs its reducers here)
        self.cell_hooks: List[Callable[[int, int, Batch], None]] = []
        #: run the last cell of every tick on the calling thread instead of its worker
        self.inline_last_cell = True

    def close(self) -> None:
        self.pool.close()

    # -- one forward pass ---------------------------------------------------
    def run(self, batches: List[Batch]) -> None:
        m, n = len(batches), len(self.partitions)
        trackers = [SkipTrackerThroughPotals(self.skip_layout) for _ in batches]
        # row counts of the micro-batches + a dict shared by the cells of this pass: a last-stage module may
        # write its output straight into its slice of a mini-batch-wide buffer (microbatch.GatherBuffer)
        sizes = [b[b.find_tensor_idx()].size(0) if b[b.find_tensor_idx()].dim() else 1 for b in batches]
        shared: dict = {}
        self._cells = [CellContext(i, sizes, shared) for i in range(m)]
        for tick in clock_cycles(m, n):
            self.fence(batches, tick, trackers)
            self.compute(batches, tick, trackers)

    def fence(self, batches: List[Batch], tick: List[Cell],
              trackers: List[SkipTrackerThroughPotals]) -> None:
        streams = self.copy_streams
        for i, j in tick:
            # Stage 0 needs no edge: its backward is already ordered by stage 1's.
            if i != 0 and j != 0:
                depend(batches[i - 1], batches[i])

            here = streams[j][i]
            for prev_j, ns, name in self.skip_layout.copy_policy(j):
                trackers[i].copy(batches[i], streams[prev_j][i], here, ns, name)

            if j != 0:
                copy_batch(batches[i], streams[j - 1][i], here)

    def compute(self, batches: List[Batch], tick: List[Cell],
                trackers: List[SkipTrackerThroughPotals]) -> None:
        partitions, devices, copy_streams = self.partitions, self.devices, self.copy_streams
        n = len(partitions)

        # No checkpointing outside training: nothing will be back-propagated.
        checkpoint_stop = self.checkpoint_stop if partitions[0].training else 0

        compute_streams = [current_stream(d) for d in devices]
        timeline = self.timeline
        first_exc: Optional[ExcInfo] = None

        # -- dispatch ------------------------------------------------------
        inline = None
        for i, j in tick:
            batch = batches[i]
            if j != 0:
                wait_batch(batch, copy_streams[j][i], compute_streams[j])

            task = self._make_task(i, j, batch, partitions[j], trackers[i],
                                   compute_streams[j], checkpointed=i < checkpoint_stop)
            if self.inline_last_cell and (i, j) == tick[-1]:
                inline = (j, task)          # run below, on this thread, while the workers run the rest
            else:
                self.in_queues[j].put(task)

        # The last cell of the tick runs on the calling thread: one queue hop and one thread wake-up
        # less on the critical path (for a single-stage pipe that is every cell).  Workers still run
        # the other cells of the tick concurrently, so cross-device overlap is unchanged.
        inline_result = None
        if inline is not None:
            j_in, task_in = inline
            try:
                with use_device(devices[j_in]):
                    inline_result = (True, (task_in, task_in.compute()))
            except Exception:
                inline_result = (False, sys.exc_info())

        # -- collect -------------------------------------------------------
        for i, j in tick:
            if inline is not None and (i, j) == tick[-1]:
                ok, payload = inline_result
            else:
                ok, payload = self.out_queues[j].get()

            if first_exc is not None:
                continue  # keep draining so the queues stay in lock-step
            if not ok:
                first_exc = payload  # type: ignore[assignment]
                continue

            task, batch = payload  # type: ignore[misc]
            cell_in = batches[i]

            # The last stage's output is consumed on its compute stream (loss,
            # gather), so it needs no trailing Wait.
            if j != n - 1:
                wait_batch(batch, compute_streams[j], copy_streams[j][i])

            # Checkpointed cells: hang the Recompute node behind that Wait so it
            # is queued before backward blocks on the incoming gradient copy.
            with use_device(devices[j]):
                task.finalize(batch)

            if timeline is not None:
                k_in = cell_in.find_tensor_idx()
                k_out = batch.find_tensor_idx()
                timeline.hook_backward(i, j, cell_in[k_in], batch[k_out], devices[j])

            batches[i] = batch
            for hook in self.cell_hooks:
                hook(i, j, batch)

        if first_exc is not None:
            raise first_exc[0].with_traceback(first_exc[1], first_exc[2])

    # -- helpers -----------------------------------------------------------
    def _make_task(self, i: int, j: int, batch: Batch, partition: nn.Module,
                   tracker: SkipTrackerThroughPotals, stream: AbstractStream,
                   *, checkpointed: bool) -> Task:
        label = f"chunk{i}-part{j}" if self.trace_cells else None
        timeline = self.timeline
        device = self.devices[j]

        cell = self._cells[i] if i < len(getattr(self, "_cells", ())) else None

        def stage(*inputs):
            with use_skip_tracker(tracker), _cell_range(label, device), use_cell(cell):
                return partition(*inputs)

        if checkpointed:
            chk = Checkpointing(stage, batch)

            def compute_ckpt() -> Batch:
                if timeline is not None:
                    timeline.mark("fwd", i, j, "start", device)
                out = chk.checkpoint()
                if timeline is not None:
                    timeline.mark("fwd", i, j, "end", device)
                return out

            return Task(stream, compute=compute_ckpt, finalize=chk.recompute)

        def compute_plain() -> Batch:
            if timeline is not None:
                timeline.mark("fwd", i, j, "start", device)
            # same seed derivation as a checkpointed cell (one draw from the CPU generator per cell, the
            # call sites' seeds derived from it): 'never' and 'always' see identical dropout masks
            with seed_scope(SeedBase()):
                out = batch.call(stage)
            if timeline is not None:
                timeline.mark("fwd", i, j, "end", device)
            return out

        return Task(stream, compute=compute_plain, finalize=None)
