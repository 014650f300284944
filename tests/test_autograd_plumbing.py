"""stream / phony / fork-join / Copy / Wait on the CPU stand-in."""
import weakref

import pytest
import torch

from pipe_b200.autograd.copy import Copy, Wait
from pipe_b200.autograd.dependency import Fork, Join, fork, join
from pipe_b200.autograd.phony import get_phony
from pipe_b200.autograd.stream import (CPUStream, current_stream, default_stream, get_device,
                                       is_cuda, new_stream, record_stream, use_device, use_stream,
                                       wait_stream)

cpu = torch.device("cpu")


def test_cpu_stream_is_singleton_sentinel():
    assert new_stream(cpu) is CPUStream
    assert current_stream(cpu) is CPUStream
    assert default_stream(cpu) is CPUStream
    assert not is_cuda(CPUStream)
    assert get_device(CPUStream) == cpu
    with use_device(cpu), use_stream(CPUStream):
        pass
    wait_stream(CPUStream, CPUStream)
    record_stream(torch.zeros(2), CPUStream)


def test_phony_cached_and_empty():
    p = get_phony(cpu, requires_grad=True)
    assert p.numel() == 0 and p.requires_grad
    assert get_phony(cpu, requires_grad=True) is p
    assert get_phony(cpu, requires_grad=False) is not p
    assert not get_phony(cpu, requires_grad=False).requires_grad


def test_fork_join_order_backward():
    """b's upstream must run before a's upstream in backward."""
    order = []

    class Log(torch.autograd.Function):
        @staticmethod
        def forward(ctx, name, x):
            ctx.name = name
            return x.detach()

        @staticmethod
        def backward(ctx, g):
            order.append(ctx.name)
            return None, g

    a = Log.apply("a", torch.rand(1, requires_grad=True))
    b = Log.apply("b", torch.rand(1, requires_grad=True))
    a, phony = fork(a)
    b = join(b, phony)
    (a + b).backward()
    assert order == ["b", "a"]


def test_fork_join_enable_grad_and_no_grad():
    x = torch.rand(1, requires_grad=True)
    x2, p = fork(x)
    assert p.requires_grad and x2.grad_fn is not None and isinstance(x2.grad_fn, Fork._backward_cls)
    with torch.no_grad():
        x3, p3 = fork(x)
        assert not p3.requires_grad and x3 is x
    y = torch.rand(1)
    assert join(y, get_phony(cpu, requires_grad=False)) is y
    z = join(y, p)
    assert z.requires_grad and isinstance(z.grad_fn, Join._backward_cls)


def test_fork_leak():
    leak = None

    class F(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return x

        @staticmethod
        def backward(ctx, g):
            nonlocal leak
            leak = weakref.ref(ctx)
            return g

    a = F.apply(torch.rand(1, requires_grad=True))
    b = torch.rand(1)
    a, phony = fork(a)
    b = join(a, phony)
    b.backward()
    del a, b, phony
    assert leak() is None


def test_copy_wait_cpu_roundtrip():
    x = torch.rand(3, requires_grad=True)
    (y,) = Copy.apply(CPUStream, CPUStream, x)
    (z,) = Wait.apply(CPUStream, CPUStream, y)
    z.sum().backward()
    assert torch.equal(x.grad, torch.ones(3))


def test_copy_passes_non_tensors_and_none_grads():
    x = torch.rand(2, requires_grad=True)
    y, k, w = Copy.apply(CPUStream, CPUStream, x, 5, torch.arange(2))
    assert k == 5 and w.dtype == torch.int64
    y.sum().backward()
    assert x.grad is not None


@pytest.mark.gpu
def test_copy_wait_cuda():
    dev = torch.device("cuda", 0)
    s1, s2 = new_stream(dev), new_stream(dev)
    x = torch.rand(1 << 20, device=dev, requires_grad=True)
    wait_stream(s1, current_stream(dev))
    (y,) = Copy.apply(s1, s2, x)
    (y,) = Wait.apply(s2, current_stream(dev), y)
    y.sum().backward()
    torch.cuda.synchronize()
    assert torch.equal(x.grad, torch.ones_like(x))


@pytest.mark.gpu
def test_record_stream_shifted_view():
    dev = torch.device("cuda", 0)
    s = new_stream(dev)
    base = torch.zeros(1024, device=dev)
    view = base[100:]
    record_stream(view, s)      # must not raise; records on the whole storage
    ptr = base.data_ptr()
    with use_stream(s):
        torch.cuda._sleep(20_000_000)
    del base, view
    fresh = torch.zeros(1024, device=dev)
    assert fresh.data_ptr() != ptr      # block still owned by stream s
    torch.cuda.synchronize()
