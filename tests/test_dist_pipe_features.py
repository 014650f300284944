"""DistPipe over gloo on CPU: a changed batch size, evaluation after training, ``@skippable``
stash / pop across ranks (direct route), deferred batch norm, per-rank timeline.

Reference behaviour these pin: a short last batch just works (``pipe.py:484``); skip tensors go
straight from the stashing stage to the popping stage (``pipeline.py:134-138``); BatchNorm
conversion (``pipe.py:341-342``)."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

from pipe_b200.skip import Namespace, pop, skippable, stash


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, fn, *args):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, q, args)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get() for _ in range(world)]
    for p in procs:
        p.join(120)
    errs = [r for r in res if r[1] is not None]
    assert not errs, "\n".join(e[1] for e in errs)
    return {r[0]: r[2] for r in res}


def _entry(rank, world, port, fn, q, args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, None, fn(rank, world, *args)))
    except Exception:
        q.put((rank, traceback.format_exc(), None))
    finally:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------
def _stages(world):
    torch.manual_seed(0)
    layers = [nn.Linear(4, 8), nn.Tanh(), nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 8), nn.Linear(8, 2)]
    per = len(layers) // world
    return [nn.Sequential(*layers[k * per:(k + 1) * per if k < world - 1 else None]) for k in range(world)]


def _slice_params(full, world, rank):
    layers = list(full)
    per = len(layers) // world
    sel = layers[rank * per:(rank + 1) * per if rank < world - 1 else None]
    for l in sel:
        yield from l.parameters()


def _changing_batches(rank, world):
    from pipe_b200.parallel import DistPipe
    stages = _stages(world)
    pipe = DistPipe(stages[rank], chunks=4, checkpoint="except_last", device=torch.device("cpu"))
    full = nn.Sequential(*[l for s in _stages(world) for l in s])
    torch.manual_seed(3)
    batches = [torch.rand(10, 4), torch.rand(10, 4), torch.rand(3, 4), torch.rand(10, 4)]   # short batch: 3 micro-batches
    for x in batches:
        out = pipe(x) if pipe.is_first else pipe()
        loss = out.local_value().pow(2).sum() if pipe.is_last else None
        pipe.backward(loss)
        if pipe.is_last:
            assert torch.allclose(out.local_value(), full(x), atol=1e-6)
    # evaluation pass under no_grad after training steps, then training again
    pipe.eval()
    with torch.no_grad():
        out = pipe(batches[0]) if pipe.is_first else pipe()
    if pipe.is_last:
        assert not out.local_value().requires_grad
        assert torch.allclose(out.local_value(), full(batches[0]), atol=1e-6)
    pipe.train()
    for p in stages[rank].parameters():
        p.grad = None
    out = pipe(batches[0]) if pipe.is_first else pipe()
    pipe.backward(out.local_value().pow(2).sum() if pipe.is_last else None)
    full(batches[0]).pow(2).sum().backward()
    for a, b in zip(stages[rank].parameters(), _slice_params(full, world, rank)):
        assert torch.allclose(a.grad, b.grad, atol=1e-5)
    assert len(pipe._caches) == 3          # (10 rows, train) (3 rows, train) (10 rows, eval/no_grad)
    return True


def test_short_last_batch_and_eval_after_train():
    _run(2, _changing_batches)


def _static_shapes(rank, world):
    from pipe_b200.parallel import DistPipe
    stages = _stages(world)
    pipe = DistPipe(stages[rank], chunks=2, checkpoint="never", device=torch.device("cpu"), static_shapes=True)
    x = torch.rand(6, 4)
    for _ in range(3):
        out = pipe(x) if pipe.is_first else pipe()
        pipe.backward(out.local_value().sum() if pipe.is_last else None)
    if pipe.is_first:
        try:
            pipe(torch.rand(5, 4))
        except RuntimeError as e:
            assert "static_shapes" in str(e)
        else:
            raise AssertionError("a changed input must be refused with static_shapes=True")
    return True


def test_static_shapes_skips_header_and_refuses_changes():
    _run(2, _static_shapes)


# ---------------------------------------------------------------------------------------
@skippable(stash=["res"])
class Stash(nn.Module):
    def __init__(self):
        super().__init__()
        self.lin = nn.Linear(4, 4)

    def forward(self, x):
        y = self.lin(x)
        yield stash("res", y)
        return torch.tanh(y)


@skippable(pop=["res"])
class Pop(nn.Module):
    def __init__(self):
        super().__init__()
        self.lin = nn.Linear(4, 4)

    def forward(self, x):
        res = yield pop("res")
        return self.lin(x) + 2.0 * res


def _skip_model(ns_key):
    torch.manual_seed(1)
    s, p = Stash(), Pop()
    if ns_key is not None:
        ns = Namespace(key=ns_key)
        s, p = s.isolate(ns), p.isolate(ns)
    mid = nn.Sequential(nn.Linear(4, 4), nn.Tanh())
    return [nn.Sequential(s), mid, nn.Sequential(p, nn.Linear(4, 2))]


def _skip_across_ranks(rank, world, mode, ns_key):
    from pipe_b200.parallel import DistPipe
    stages = _skip_model(ns_key)
    pipe = DistPipe(stages[rank], chunks=3, checkpoint=mode, device=torch.device("cpu"))
    assert pipe._skip_out == ({(ns_key, "res"): 2} if rank == 0 else {})
    assert pipe._skip_in == ({(ns_key, "res"): 0} if rank == 2 else {})
    torch.manual_seed(7)
    x = torch.rand(6, 4)
    full = nn.Sequential(*[l for st in _skip_model(ns_key) for l in st])
    for _ in range(2):
        for p_ in stages[rank].parameters():
            p_.grad = None
        out = pipe(x) if pipe.is_first else pipe()
        pipe.backward(out.local_value().pow(2).sum() if pipe.is_last else None)
    ref_out = full(x)
    ref_out.pow(2).sum().backward()
    if pipe.is_last:
        assert torch.allclose(out.local_value(), ref_out, atol=1e-6)
    ref_layers = list(full)
    mine = {0: ref_layers[0:1], 1: ref_layers[1:3], 2: ref_layers[3:5]}[rank]
    ref_params = [p_ for l in mine for p_ in l.parameters()]
    for a, b in zip(stages[rank].parameters(), ref_params):
        assert torch.allclose(a.grad, b.grad, atol=1e-5), (rank, a.grad, b.grad)
    return True


def test_skip_travels_directly_between_ranks():
    for mode, key in (("never", None), ("always", "blockA"), ("except_last", None)):
        _run(3, _skip_across_ranks, mode, key)


def _unmatched_skip(rank, world):
    from pipe_b200.parallel import DistPipe
    torch.manual_seed(1)
    s, p = Stash().isolate(Namespace()), Pop().isolate(Namespace())      # process-local namespaces never match
    stage = nn.Sequential(s) if rank == 0 else nn.Sequential(p)
    try:
        DistPipe(stage, chunks=2, device=torch.device("cpu"))
    except TypeError as e:
        assert "Namespace(key=" in str(e) or "no later stage pops" in str(e)
        return True
    raise AssertionError("unmatched cross-rank skip must be rejected at construction")


def test_unmatched_cross_rank_skip_is_rejected():
    _run(2, _unmatched_skip)


# ---------------------------------------------------------------------------------------
def _deferred_bn(rank, world):
    from pipe_b200.batchnorm import DeferredBatchNorm
    from pipe_b200.parallel import DistPipe
    torch.manual_seed(0)
    stages = [nn.Sequential(nn.Conv2d(3, 4, 1), nn.BatchNorm2d(4)), nn.Sequential(nn.ReLU(), nn.Conv2d(4, 2, 1))]
    ref_bn = nn.BatchNorm2d(4)
    pipe = DistPipe(stages[rank], chunks=4, checkpoint="always", device=torch.device("cpu"), deferred_batch_norm=True,
                    timeline=True)
    torch.manual_seed(11)
    x = torch.rand(8, 3, 5, 5)
    out = pipe(x) if pipe.is_first else pipe()
    pipe.backward(out.local_value().sum() if pipe.is_last else None)
    summary = pipe.timeline_summary()
    assert summary is not None and len(summary) == world and all("bubble" in s for s in summary)
    assert summary[rank]["recompute_ms"] > 0.0
    if rank == 0:
        bn = pipe.stage[1]
        assert isinstance(bn, DeferredBatchNorm)
        ref_bn.train()
        ref_bn(stages[0][0](x))                     # one mini-batch-wide update
        assert torch.allclose(bn.running_mean, ref_bn.running_mean, atol=1e-5)
        assert torch.allclose(bn.running_var, ref_bn.running_var, atol=1e-4)
    return True


def test_deferred_batch_norm_and_timeline():
    _run(2, _deferred_bn)
