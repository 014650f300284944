"""Batch / scatter / gather / check / NoChunk."""
import pytest
import torch

from pipe_b200.microbatch import Batch, GatherBuffer, NoChunk, check, gather, scatter


def test_batch_atomic():
    x = torch.tensor(42)
    b = Batch(x)
    assert b.atomic and b.tensor is x and b.values is x
    assert list(b) == [x] and len(b) == 1 and b[0] is x
    with pytest.raises(IndexError):
        b[1]
    assert b.find_tensor_idx() == 0
    assert b.get_device() == torch.device("cpu")


def test_batch_non_atomic():
    x, y = torch.tensor(42), torch.tensor(21)
    b = Batch((x, y))
    assert not b.atomic
    with pytest.raises(AttributeError):
        b.tensor
    assert b.values == (x, y) and len(b) == 2 and b[0] is x and b[1] is y
    b2 = Batch((1, "a", y))
    assert b2.find_tensor_idx() == 2


def test_batch_needs_a_tensor():
    with pytest.raises(TypeError):
        Batch((1, 2))


def test_batch_call():
    a = Batch(torch.tensor(42))
    b = Batch((torch.tensor(42), torch.tensor(21)))
    assert a.call(lambda x: x * 2).tensor == 84
    assert b.call(lambda x, y: (x + y, y))[0] == 63


def test_batch_setitem():
    a = Batch(torch.tensor(1))
    a[0] = torch.tensor(2)
    assert a.tensor == 2
    a[:] = (torch.tensor(3),)
    assert a.tensor == 3
    with pytest.raises(IndexError):
        a[:] = (torch.tensor(1), torch.tensor(2))
    b = Batch((torch.tensor(1), torch.tensor(2)))
    b[1] = torch.tensor(5)
    assert b[1] == 5
    b[:] = (torch.tensor(7), 3, torch.tensor(9))
    assert len(b) == 3
    with pytest.raises(NotImplementedError):
        b[0:1] = (torch.tensor(1),)


def test_check():
    check(torch.device("cpu"), torch.tensor(1))
    check(torch.device("cpu"), 1, torch.tensor(1))
    with pytest.raises(TypeError):
        check(torch.device("cpu"), 42)
    with pytest.raises(TypeError):
        check(torch.device("cpu"))


def test_scatter_tensor_and_uneven():
    parts = scatter(torch.zeros(20, 3), chunks=8)
    assert [p.tensor.size(0) for p in parts] == [3, 3, 3, 3, 3, 3, 2]     # 7, not 8 (Appendix B #1)
    assert len(scatter(torch.zeros(20, 3), chunks=16)) == 10
    assert len(scatter(torch.zeros(20, 3), chunks=32)) == 20


def test_scatter_tuple_nochunk_nontensor():
    a, b = torch.zeros(4, 1), torch.ones(4, 2)
    whole = torch.arange(3)
    parts = scatter(a, 7, NoChunk(whole), b, chunks=2)
    assert len(parts) == 2
    assert parts[0][0].shape == (2, 1) and parts[0][1] == 7 and parts[0][2] is whole and parts[1][3].shape == (2, 2)


def test_scatter_mismatch():
    with pytest.raises(RuntimeError):
        scatter(torch.zeros(4, 1), torch.zeros(1, 1), chunks=4)


def test_nochunk_rejects_non_tensor():
    with pytest.raises(TypeError):
        NoChunk(3)


def test_gather_tensor_and_tuple():
    assert gather([Batch(torch.zeros(2, 1)), Batch(torch.zeros(3, 1))]).shape == (5, 1)
    out = gather([Batch((torch.zeros(2, 1), "x")), Batch((torch.zeros(1, 1), "y"))])
    assert out[0].shape == (3, 1) and out[1] == ["x", "y"]
    with pytest.raises(TypeError):
        gather([Batch((torch.zeros(2, 1), "x")), Batch((torch.zeros(1, 1), 3))])


def test_scatter_gather_roundtrip():
    x = torch.randn(20, 5)
    assert torch.equal(gather(scatter(x, chunks=8)), x)


def test_gather_buffer():
    gb = GatherBuffer(5, (2,), dtype=torch.float32, device=torch.device("cpu"))
    gb.plan([3, 2])
    gb.slot(0).fill_(1)
    gb.slot(1).fill_(2)
    assert gb.buffer[:, 0].tolist() == [1, 1, 1, 2, 2]


def test_gather_is_zero_copy_for_adjacent_slices_and_keeps_gradients():
    gb = GatherBuffer(5, (2,), dtype=torch.float32, device=torch.device("cpu"))
    gb.plan([3, 2])
    w = torch.ones(2, requires_grad=True)

    class WriteInto(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, slot):
            slot.copy_(x)          # what a last-stage kernel does with the raw pointer
            return torch.empty(0).set_(slot.untyped_storage(), slot.storage_offset(), slot.shape)

        @staticmethod
        def backward(ctx, g):
            return g, None

    a = WriteInto.apply(torch.arange(6.0).view(3, 2) * w, gb.slot(0))
    b = WriteInto.apply(torch.arange(4.0).view(2, 2) * w, gb.slot(1))
    out = gather([Batch(a), Batch(b)])
    assert out.data_ptr() == gb.buffer.data_ptr() and out.shape == (5, 2)      # no copy
    out.sum().backward()
    assert w.grad.tolist() == [0 + 2 + 4 + 0 + 2, 1 + 3 + 5 + 1 + 3]
    # non-adjacent tensors still go through torch.cat
    c = gather([Batch(torch.zeros(2, 2)), Batch(torch.ones(1, 2))])
    assert c.shape == (3, 2)
