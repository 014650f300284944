"""Fused-dropout seeds under checkpointing with several stage threads (finding 1).

A stage that asks ``draw_seed()`` for its dropout seed must get the same value when the cell is
recomputed in backward, although other stage threads (and here a deliberately noisy thread) advance
the process-global CPU generator in between.  Reference: RNG restore of the absent ``checkpoint``
module, ``README.md:472-482, 525-537``."""
import threading

import pytest
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.autograd.checkpoint import is_recomputing
from pipe_b200.autograd.seeds import SeedBase, draw_seed, seed_scope
from pipe_b200.ops.functional import draw_seed as ops_draw_seed


class SeededDropout(nn.Module):
    """Counter-RNG dropout in miniature: the mask is a pure function of the drawn seed."""

    def __init__(self, name):
        super().__init__()
        self.name = name
        self.lin = nn.Linear(8, 8)
        self.log = []

    def forward(self, x):
        seed = ops_draw_seed()
        self.log.append((is_recomputing(), float(x.detach().sum()), seed))
        g = torch.Generator().manual_seed(seed % (2 ** 31))
        mask = (torch.rand(x.shape, generator=g) > 0.3).to(x.dtype)
        return self.lin(x) * mask


def test_scope_replays_and_isolates():
    base = SeedBase()
    with seed_scope(base):
        a = [draw_seed() for _ in range(3)]
    torch.manual_seed(999)                      # global generator state is irrelevant inside a scope
    with seed_scope(base):
        b = [draw_seed() for _ in range(3)]
    assert a == b and len(set(a)) == 3
    other = SeedBase()
    with seed_scope(other):
        assert draw_seed() != a[0]
    torch.manual_seed(5)
    x = draw_seed()
    torch.manual_seed(5)
    assert draw_seed() == x                     # outside a scope torch.manual_seed decides


@pytest.mark.parametrize("mode", ["always", "except_last"])
def test_recompute_sees_forward_seeds_with_noisy_threads(mode):
    torch.manual_seed(0)
    a, b = SeededDropout("a"), SeededDropout("b")
    pipe = Pipe(nn.Sequential(a, b), chunks=4, checkpoint=mode)        # two CPU stages
    stop = threading.Event()

    def noise():
        while not stop.is_set():
            draw_seed()                          # another "stage thread" advancing the CPU generator

    t = threading.Thread(target=noise, daemon=True)
    t.start()
    try:
        x = torch.randn(8, 8, requires_grad=True)
        out = pipe(x).local_value()
        out.sum().backward()
    finally:
        stop.set()
        t.join()
    for layer in (a, b):
        fwd = {key: seed for rec, key, seed in layer.log if not rec}
        rec = [(key, seed) for r, key, seed in layer.log if r]
        assert rec, "nothing was recomputed"
        for key, seed in rec:
            assert fwd[key] == seed, f"stage {layer.name}: recompute drew a different seed"
    pipe.close()


def test_checkpointed_gradients_match_never_given_same_seeds():
    """'always' and 'never' give identical gradients once the per-cell seeds are the same: the
    recompute regenerates the forward's masks."""
    torch.manual_seed(0)
    ref = nn.Sequential(SeededDropout("a"), SeededDropout("b"))
    x = torch.randn(8, 8)
    pipe = Pipe(ref, chunks=2, checkpoint="always")
    out = pipe(x.clone().requires_grad_()).local_value()
    out.sum().backward()
    # replay by hand: same seeds (taken from the log), no pipeline
    seeds = {(l.name, key): seed for l in ref for rec, key, seed in l.log if not rec}
    grads = {n: p.grad.clone() for n, p in pipe.named_parameters()}
    for p in pipe.parameters():
        p.grad = None
    total = 0
    for chunk in x.chunk(2):
        h = chunk
        for l in ref:
            seed = seeds[(l.name, float(h.detach().sum()))]
            g = torch.Generator().manual_seed(seed % (2 ** 31))
            mask = (torch.rand(h.shape, generator=g) > 0.3).to(h.dtype)
            h = l.lin(h) * mask
        total = total + h.sum()
    total.backward()
    for n, p in pipe.named_parameters():
        torch.testing.assert_close(p.grad, grads[n])
    pipe.close()
