"""Pipe API: validation, partitioning rules, container protocol, device-move denial,
exception tunnelling, RRef return, state-dict layout."""
import pytest
import torch
from torch import nn

from pipe_b200 import LocalRRef, NoChunk, Pipe, PipeSequential, WithDevice, is_checkpointing, is_recomputing
from pipe_b200.partition import BalanceError
from pipe_b200.state import from_sequential_state_dict, key_map, to_sequential_state_dict


def test_parameters_and_len_iter_getitem():
    a, b, c = nn.Linear(1, 1), nn.Linear(1, 1), nn.Linear(1, 1)
    pipe = Pipe(nn.Sequential(a, b, c))
    assert len(list(pipe.parameters())) == 6
    assert len(pipe) == 3
    assert list(pipe) == [a, b, c]
    assert pipe[0] is a and pipe[-1] is c and pipe[-3] is a
    with pytest.raises(IndexError):
        pipe[3]
    with pytest.raises(IndexError):
        pipe[-4]


def test_every_cpu_child_is_its_own_stage():
    pipe = Pipe(nn.Sequential(nn.Linear(1, 1), nn.ReLU(), nn.Linear(1, 1)))
    assert len(pipe.partitions) == 3
    assert pipe.devices == [torch.device("cpu")] * 3
    assert all(isinstance(p, PipeSequential) for p in pipe.partitions)


def test_nested_sequential_is_flattened_one_level():
    inner = nn.Sequential(nn.Linear(2, 2), nn.Sequential(nn.ReLU(), nn.Linear(2, 2)))
    pipe = Pipe(nn.Sequential(inner, nn.Linear(2, 2)))
    assert len(pipe.partitions) == 2
    assert len(pipe.partitions[0]) == 2
    assert isinstance(pipe.partitions[0][1], nn.Sequential)      # deeper level stays nested
    keys = list(pipe.state_dict().keys())
    assert keys[0] == "partitions.0.0.weight" and "partitions.0.1.1.weight" in keys
    assert "partitions.1.0.bias" in keys


def test_chunks_and_checkpoint_validation():
    m = nn.Sequential(nn.Linear(1, 1))
    with pytest.raises(ValueError):
        Pipe(m, chunks=0)
    with pytest.raises(ValueError):
        Pipe(m, chunks=-1)
    with pytest.raises(ValueError):
        Pipe(m, checkpoint="sometimes")
    assert Pipe(m, chunks="3").chunks == 3


def test_module_must_be_sequential_without_duplicates():
    with pytest.raises(TypeError):
        Pipe(nn.Linear(1, 1))
    lin = nn.Linear(1, 1)
    with pytest.raises(ValueError):
        Pipe(nn.Sequential(lin, lin))


def test_device_moves_denied_dtype_allowed():
    pipe = Pipe(nn.Sequential(nn.Linear(1, 1)))
    for call in (lambda: pipe.cuda(), lambda: pipe.cpu(), lambda: pipe.to("cpu"),
                 lambda: pipe.to(torch.device("cpu")), lambda: pipe.to(0),
                 lambda: pipe.to(device="cpu"), lambda: pipe.to(torch.zeros(1)),
                 lambda: pipe.to(tensor=torch.zeros(1))):
        with pytest.raises(TypeError, match="denied to move"):
            call()
    pipe.to(torch.float64)
    assert next(pipe.parameters()).dtype == torch.float64
    pipe.to(dtype=torch.float32)


def test_forward_returns_rref_like():
    pipe = Pipe(nn.Sequential(nn.Linear(2, 2)), chunks=2)
    out = pipe(torch.rand(4, 2))
    assert isinstance(out, LocalRRef)
    assert out.is_owner()
    assert out.local_value().shape == (4, 2)
    assert out.to_here() is out.local_value()
    assert out.view(-1).shape == (8,)        # transparent use, as in the reference's patched main.py


def test_empty_sequential_passthrough():
    pipe = Pipe(nn.Sequential())
    x = torch.rand(2)
    assert pipe(x).local_value() is x
    a, b = pipe(x, x).local_value()
    assert a is x and b is x


def test_input_checks():
    pipe = Pipe(nn.Sequential(nn.Linear(1, 1)))
    with pytest.raises(TypeError):
        pipe(42)
    with pytest.raises(TypeError):
        pipe()


def test_multi_input_nochunk_non_tensor():
    class Add(nn.Module):
        def forward(self, a, k, whole, b):
            assert whole.shape == (3,)
            return a + b * k, k

    class Second(nn.Module):
        def forward(self, x, k):
            return x * k, "tag"

    pipe = Pipe(nn.Sequential(Add(), Second()), chunks=2)
    a, b = torch.ones(4, 1), torch.ones(4, 1)
    out, tags = pipe(a, 2, NoChunk(torch.arange(3)), b).local_value()
    assert torch.equal(out, torch.full((4, 1), 6.0))
    assert tags == ["tag", "tag"]


def test_first_exception_wins_with_traceback():
    class ExpectedError(Exception):
        pass

    class Raise(nn.Module):
        def __init__(self, tag):
            super().__init__()
            self.tag = tag

        def forward(self, x):
            raise ExpectedError(self.tag)

    pipe = Pipe(nn.Sequential(Raise("first")), chunks=4)
    with pytest.raises(ExpectedError, match="first") as info:
        pipe(torch.rand(4, 1))
    import traceback
    assert "forward" in "".join(traceback.format_tb(info.value.__traceback__))
    # the pipe stays usable afterwards
    pipe2 = Pipe(nn.Sequential(nn.Identity()), chunks=2)
    assert pipe2(torch.rand(2, 1)).local_value().shape == (2, 1)


def test_exception_early_stop_keeps_workers_in_lockstep():
    class Counter(nn.Module):
        def __init__(self):
            super().__init__()
            self.n = 0

        def forward(self, x):
            self.n += 1
            if self.n == 2:
                raise RuntimeError("boom")
            return x

    c = Counter()
    pipe = Pipe(nn.Sequential(c, nn.Identity()), chunks=3)
    with pytest.raises(RuntimeError, match="boom"):
        pipe(torch.rand(3, 1))
    c.n = 10
    assert pipe(torch.rand(3, 1)).local_value().shape == (3, 1)


def test_shared_parameters_same_device_ok():
    a = nn.Linear(1, 1)
    b = nn.Linear(1, 1)
    b.weight = a.weight
    Pipe(nn.Sequential(a, b))       # both on cpu: allowed


def test_with_device_unwraps():
    lin = nn.Linear(1, 1)
    pipe = Pipe(nn.Sequential(lin, WithDevice(nn.ReLU(), "cpu")))
    assert isinstance(pipe[1], nn.ReLU)
    wd = WithDevice(nn.ReLU(), "cpu")
    assert wd.device == torch.device("cpu") and isinstance(wd.module, nn.ReLU)
    assert wd(torch.tensor([-1.0])).item() == 0.0


def test_balance_and_devices_partitioning():
    m = nn.Sequential(nn.Linear(1, 1), nn.ReLU(), nn.Linear(1, 1), nn.ReLU())
    pipe = Pipe(m, balance=[3, 1], devices=["cpu", "cpu"], chunks=2)
    assert [len(p) for p in pipe.partitions] == [3, 1]
    assert pipe(torch.rand(2, 1)).local_value().shape == (2, 1)
    with pytest.raises(BalanceError):
        Pipe(m, balance=[2, 1], devices=["cpu", "cpu"])
    with pytest.raises(BalanceError):
        Pipe(m, balance=[4, 0], devices=["cpu", "cpu"])
    with pytest.raises(IndexError):
        Pipe(m, balance=[2, 2], devices=["cpu"])
    with pytest.raises(ValueError):
        Pipe(m, devices=["cpu"])


def test_checkpoint_mode_probe_and_except_last_quirk():
    seen = []

    class Probe(nn.Module):
        def forward(self, x):
            seen.append((is_checkpointing(), is_recomputing()))
            return x * 1.0

    def run(mode, chunks, rows, **kw):
        seen.clear()
        pipe = Pipe(nn.Sequential(Probe()), chunks=chunks, checkpoint=mode, **kw)
        x = torch.rand(rows, 1, requires_grad=True)
        pipe(x).local_value().sum().backward()
        fwd_ckpt = sum(1 for c, r in seen if c)
        recomputed = sum(1 for c, r in seen if r)
        return fwd_ckpt, recomputed

    assert run("never", 4, 4) == (0, 0)
    assert run("always", 4, 4) == (4, 4)
    assert run("except_last", 4, 4) == (3, 3)
    # 20 rows / 8 chunks -> 7 micro-batches: the reference checkpoints ALL of them (README.md:398)
    assert run("except_last", 8, 20) == (7, 7)
    assert run("except_last", 8, 20, strict_except_last=True) == (6, 6)


def test_eval_disables_checkpointing():
    seen = []

    class Probe(nn.Module):
        def forward(self, x):
            seen.append(is_checkpointing())
            return x

    pipe = Pipe(nn.Sequential(Probe()), chunks=2, checkpoint="always")
    pipe.eval()
    pipe(torch.rand(2, 1))
    assert seen == [False, False]


def test_non_float_tensors_detached_between_stages():
    class TwoOut(nn.Module):
        def forward(self, x):
            return x * 2, torch.arange(x.size(0))

    class Use(nn.Module):
        def forward(self, x, idx):
            assert not idx.requires_grad
            return x + idx.unsqueeze(1).float()

    pipe = Pipe(nn.Sequential(TwoOut(), Use()), chunks=2, checkpoint="always")
    x = torch.rand(4, 1, requires_grad=True)
    pipe(x).local_value().sum().backward()
    assert torch.equal(x.grad, torch.full((4, 1), 2.0))


def test_state_dict_roundtrip_and_converters():
    torch.manual_seed(1)
    seq = nn.Sequential(nn.Linear(2, 3), nn.ReLU(), nn.Linear(3, 2))
    ref_sd = {k: v.clone() for k, v in seq.state_dict().items()}
    pipe = Pipe(seq, chunks=2)
    sd = pipe.state_dict()
    assert list(sd.keys()) == ["partitions.0.0.weight", "partitions.0.0.bias",
                               "partitions.2.0.weight", "partitions.2.0.bias"]
    assert key_map(pipe, seq)["2.weight"] == "partitions.2.0.weight"
    back = to_sequential_state_dict(pipe, seq)
    assert list(back.keys()) == list(ref_sd.keys())
    for k in ref_sd:
        assert torch.equal(back[k], ref_sd[k])
    fresh = {k: torch.zeros_like(v) for k, v in ref_sd.items()}
    pipe.load_state_dict(from_sequential_state_dict(pipe, seq, fresh))
    assert all(float(p.abs().sum()) == 0 for p in pipe.parameters())
    # one nested Sequential per device: keys match the wrapped module plus the prefix
    nested = nn.Sequential(nn.Sequential(nn.Linear(1, 1), nn.ReLU()))
    expect = ["partitions." + k for k in nested.state_dict().keys()]
    assert list(Pipe(nested).state_dict().keys()) == expect


def test_close_stops_workers():
    pipe = Pipe(nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1)))
    assert pipe.pipeline.pool.alive == 1
    pipe.close()
    assert pipe.pipeline.pool.alive == 0


def test_trace_cells_labels(monkeypatch):
    from contextlib import contextmanager

    import pipe_b200.runtime.executor as ex
    seen = []

    @contextmanager
    def fake_record_function(label):
        seen.append(label)
        yield

    monkeypatch.setattr(ex, "record_function", fake_record_function)
    pipe = Pipe(nn.Sequential(nn.Linear(1, 1), nn.Linear(1, 1)), chunks=2)
    pipe(torch.rand(2, 1))
    assert seen == []                      # off by default, like the reference as shipped
    pipe.trace_cells(True)
    pipe(torch.rand(2, 1))
    assert sorted(seen) == ["chunk0-part0", "chunk0-part1", "chunk1-part0", "chunk1-part1"]


@pytest.mark.parametrize("inline", [True, False])
def test_inline_last_cell_toggle_is_transparent(inline):
    """Running the last cell of a tick on the calling thread or on its worker gives the same result,
    the same exception behaviour and the same thread for the *other* cells."""
    import threading
    torch.manual_seed(0)
    seen = []

    class Probe(nn.Module):
        def forward(self, x):
            seen.append(threading.current_thread().name)
            return x * 1.0

    seq = nn.Sequential(nn.Linear(4, 4), Probe(), nn.Linear(4, 2))
    ref = nn.Sequential(nn.Linear(4, 4), nn.Identity(), nn.Linear(4, 2))
    ref[0].load_state_dict(seq[0].state_dict())
    ref[2].load_state_dict(seq[2].state_dict())
    pipe = Pipe(seq, chunks=3, checkpoint="except_last")
    pipe.pipeline.inline_last_cell = inline
    x = torch.rand(6, 4)
    out = pipe(x).local_value()
    assert torch.allclose(out, ref(x), atol=1e-6)
    out.sum().backward()
    ref(x).sum().backward()
    assert torch.allclose(pipe[0].weight.grad, ref[0].weight.grad, atol=1e-6)
    main = threading.current_thread().name
    if inline:
        assert main in seen                      # some cells ran on the calling thread
    else:
        assert main not in seen[:3]              # forward cells all went through the worker
    pipe.close()
