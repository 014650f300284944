"""Per-device worker threads and the ``Task`` unit they execute.

One daemon thread per *distinct* device pulls ``Task`` objects off an input
queue, runs ``task.compute()`` on the task's stream, and posts
``(True, (task, batch))`` or ``(False, exc_info)`` on its output queue.  Stages
that share a device share the thread (and the queue pair).  A worker that hits
an exception reports it and keeps serving, so one bad micro-batch does not
wedge the pipeline.  Unlike the reference (whose threads leak until interpreter
exit) a pool can be shut down with :meth:`close`.

Parity: reference ``pipeline.py:23,98,213,233,237,240,260`` (call sites) and
``README.md:39-47,291-314`` (excerpts of the absent ``worker`` module);
SURVEY.md A2.
"""
from __future__ import annotations

import atexit
import sys
import threading
import weakref
from queue import Queue
from types import TracebackType
from typing import Callable, Dict, List, Optional, Tuple, Type

import torch

from ..autograd.stream import AbstractStream, use_device, use_stream
from ..microbatch import Batch

__all__ = ["Task", "WorkerPool", "create_workers", "ExcInfo"]

ExcInfo = Tuple[Type[BaseException], BaseException, TracebackType]
InQueue = "Queue[Optional[Task]]"
OutQueue = "Queue[Tuple[bool, Union[Tuple[Task, Batch], ExcInfo, None]]]"


class Task:
    """A unit of stage work: ``compute`` runs on the worker thread, the optional
    ``finalize`` later on the thread that collects the result.

    The grad mode at construction time is captured and replayed on the worker,
    because grad mode is thread-local and the worker would otherwise always run
    with autograd enabled.
    """

    def __init__(self, stream: AbstractStream, *, compute: Callable[[], Batch],
                 finalize: Optional[Callable[[Batch], None]]) -> None:
        self.stream = stream
        self._compute = compute
        self._finalize = finalize
        self._grad_enabled = torch.is_grad_enabled()

    def compute(self) -> Batch:
        with use_stream(self.stream), torch.set_grad_enabled(self._grad_enabled):
            return self._compute()

    def finalize(self, batch: Batch) -> None:
        if self._finalize is None:
            return
        with use_stream(self.stream), torch.set_grad_enabled(self._grad_enabled):
            self._finalize(batch)


def _serve(in_queue: Queue, out_queue: Queue, device: torch.device) -> None:
    with use_device(device):
        while True:
            task = in_queue.get()
            if task is None:
                break
            try:
                batch = task.compute()
            except Exception:
                out_queue.put((False, sys.exc_info()))
                continue
            out_queue.put((True, (task, batch)))
    out_queue.put((False, None))


def _normalize(device: torch.device) -> torch.device:
    if device.type == "cuda" and device.index is None:
        return torch.device("cuda", index=torch.cuda.current_device())
    if device.type == "cpu" and device.index is not None:
        return torch.device("cpu")
    return device


_live_pools: "weakref.WeakSet[WorkerPool]" = weakref.WeakSet()


@atexit.register
def _close_all_pools() -> None:
    # Stop worker threads before the interpreter (and the CUDA runtime) tear down.
    for pool in list(_live_pools):
        try:
            pool.close(timeout=2.0)
        except Exception:
            pass


class WorkerPool:
    """Owns the threads; exposes per-*stage* queue lists (aliased when stages
    share a device)."""

    def __init__(self, devices: List[torch.device]) -> None:
        self.in_queues: List[Queue] = []
        self.out_queues: List[Queue] = []
        self._threads: List[threading.Thread] = []
        self._unique_in: List[Queue] = []
        self._closed = False

        by_device: Dict[torch.device, Tuple[Queue, Queue]] = {}
        for device in devices:
            device = _normalize(torch.device(device))
            pair = by_device.get(device)
            if pair is None:
                pair = (Queue(), Queue())
                by_device[device] = pair
                t = threading.Thread(target=_serve, args=(pair[0], pair[1], device),
                                     name=f"pipe-b200-worker[{device}]", daemon=True)
                t.start()
                self._threads.append(t)
                self._unique_in.append(pair[0])
            self.in_queues.append(pair[0])
            self.out_queues.append(pair[1])
        _live_pools.add(self)

    def close(self, timeout: Optional[float] = 5.0) -> None:
        """Stop the worker threads.  Idempotent."""
        if self._closed:
            return
        self._closed = True
        for q in self._unique_in:
            q.put(None)
        me = threading.current_thread()
        for t in self._threads:
            if t is not me:
                t.join(timeout)

    @property
    def alive(self) -> int:
        return sum(t.is_alive() for t in self._threads)

    def __del__(self) -> None:
        # Join, do not just signal: a worker that is still unwinding its CUDA device guard while the
        # interpreter finalises gets killed inside C++ frames, which aborts the process.
        try:
            self.close(timeout=2.0)
        except Exception:
            pass


def create_workers(devices: List[torch.device]) -> Tuple[List[Queue], List[Queue]]:
    """Reference-shaped helper: start a pool and return ``(in_queues, out_queues)``.
    The pool object is attached to the returned list as ``.pool`` so it can be
    closed."""
    pool = WorkerPool(devices)

    class _Queues(list):
        pass

    ins, outs = _Queues(pool.in_queues), _Queues(pool.out_queues)
    ins.pool = pool  # type: ignore[attr-defined]
    outs.pool = pool  # type: ignore[attr-defined]
    return ins, outs
