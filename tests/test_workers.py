"""Worker threads and Task."""
import threading
import time

import pytest
import torch

from pipe_b200.autograd.stream import CPUStream
from pipe_b200.microbatch import Batch
from pipe_b200.runtime.workers import Task, WorkerPool, create_workers

cpu = torch.device("cpu")


def test_compute_runs_on_worker_thread_and_finalize_on_caller():
    seen = {}

    def compute():
        seen["compute"] = threading.current_thread().name
        return Batch(torch.zeros(1))

    def finalize(batch):
        seen["finalize"] = threading.current_thread().name

    pool = WorkerPool([cpu])
    pool.in_queues[0].put(Task(CPUStream, compute=compute, finalize=finalize))
    ok, (task, batch) = pool.out_queues[0].get()
    assert ok
    task.finalize(batch)
    assert seen["compute"].startswith("pipe-b200-worker")
    assert seen["finalize"] == threading.current_thread().name
    pool.close()
    assert pool.alive == 0


def test_stages_on_one_device_share_a_thread():
    pool = WorkerPool([cpu, torch.device("cpu", 0), cpu])
    assert pool.in_queues[0] is pool.in_queues[1] is pool.in_queues[2]
    assert pool.alive == 1
    pool.close()


def test_exception_is_reported_and_worker_survives():
    class Boom(Exception):
        pass

    def bad():
        raise Boom("x")

    pool = WorkerPool([cpu])
    pool.in_queues[0].put(Task(CPUStream, compute=bad, finalize=None))
    ok, exc_info = pool.out_queues[0].get()
    assert not ok and exc_info[0] is Boom
    pool.in_queues[0].put(Task(CPUStream, compute=lambda: Batch(torch.ones(1)), finalize=None))
    ok, _ = pool.out_queues[0].get()
    assert ok
    pool.close()


@pytest.mark.parametrize("grad_mode", [True, False])
def test_grad_mode_is_replayed_on_worker(grad_mode):
    seen = {}

    def compute():
        seen["grad"] = torch.is_grad_enabled()
        return Batch(torch.zeros(1))

    pool = WorkerPool([cpu])
    with torch.set_grad_enabled(grad_mode):
        task = Task(CPUStream, compute=compute, finalize=None)
    pool.in_queues[0].put(task)
    pool.out_queues[0].get()
    assert seen["grad"] is grad_mode
    pool.close()


def test_create_workers_shape():
    ins, outs = create_workers([cpu, cpu])
    assert len(ins) == len(outs) == 2
    ins.pool.close()
