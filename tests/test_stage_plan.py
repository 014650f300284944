"""CPU tests of the stage-level execution plan (``PipeSequential._plan``), the per-cell context the executor provides,
and the strided zero-copy gather.  Reference: a stage is an ``nn.Sequential`` run layer by layer by its worker thread
(``pipe.py:121-133``, ``pipeline.py:225-230``); grouping runs of layers must not change what is computed."""
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.microbatch import GatherBuffer, current_cell, gather, Batch
from pipe_b200.partition import PipeSequential


class Chainable(nn.Module):
    """Stands in for the fused encoder layer: offers ``pb2_make_chain``; the chain doubles every layer's effect marker."""
    made = 0

    def __init__(self, k):
        super().__init__()
        self.k = k

    def forward(self, x, *extra):
        y = x + self.k
        return (y, *extra) if extra else y

    @staticmethod
    def pb2_make_chain(layers):
        Chainable.made += 1
        return _Chain(layers)


class _Chain:
    def __init__(self, layers):
        self.layers = layers
        self.calls = 0

    def usable(self, x):
        return x.dtype == torch.float32

    def __call__(self, x):
        self.calls += 1
        for l in self.layers:
            x = x + l.k
        return x


class Through(nn.Module):
    def forward(self, *xs):
        return xs if len(xs) > 1 else xs[0]


def test_runs_of_chainable_layers_become_one_call_and_results_do_not_change():
    seq = PipeSequential(Through(), Chainable(1.0), Chainable(2.0), Chainable(3.0), Through(), Chainable(4.0))
    plan = seq._plan()
    kinds = [type(p).__name__ for p in plan]
    assert kinds == ["Through", "tuple", "Through", "Chainable"]          # a run of three, a lone layer stays a layer
    x = torch.zeros(2)
    assert torch.equal(seq(x), torch.full((2,), 10.0))
    assert plan[1][0].calls == 1
    y, extra = seq(x, "tag")                                                # extra inputs pass through the chain
    assert extra == "tag" and torch.equal(y, torch.full((2,), 10.0))
    assert torch.equal(seq(x.double()), torch.full((2,), 10.0, dtype=torch.float64))   # not usable: layer by layer
    assert plan[1][0].calls == 2
    assert seq._plan() is plan                                              # cached until the children change
    seq.append(Chainable(5.0))
    assert seq._plan() is not plan and torch.equal(seq(x), torch.full((2,), 15.0))


def test_cell_context_reaches_the_stage_and_gather_buffer_makes_gather_a_view():
    seen = []

    class LastStage(nn.Module):
        """Writes its output into its slice of one mini-batch-wide buffer, found through the cell context."""

        def forward(self, x):
            cell = current_cell()
            seen.append((cell.index, tuple(cell.sizes)))
            gb = cell.shared.get("gb")
            if gb is None:
                gb = cell.shared["gb"] = GatherBuffer(sum(cell.sizes), (x.size(1) + 3,), dtype=x.dtype, device=x.device)
                gb.plan(cell.sizes)
            slot = gb.slot(cell.index)
            slot[:, : x.size(1)].copy_(x * 2)
            return slot[:, : x.size(1)]                   # padded rows: strided, not contiguous

    pipe = Pipe(nn.Sequential(nn.Identity(), LastStage()), chunks=3, checkpoint="never")
    x = torch.arange(10.0).view(5, 2)
    out = pipe(x).local_value()
    assert sorted(seen) == [(0, (2, 2, 1)), (1, (2, 2, 1)), (2, (2, 2, 1))]
    assert torch.equal(out, x * 2)
    assert out.stride() == (5, 1)                         # a view over the padded buffer: nothing was concatenated
    assert current_cell() is None
    pipe.close()


def test_gather_still_copies_when_slices_are_not_adjacent():
    a, b = torch.ones(2, 3), torch.zeros(1, 3)
    assert torch.equal(gather([Batch(a), Batch(b)]), torch.cat([a, b]))
