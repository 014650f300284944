"""@skippable / stash / pop / verify / layout / tracker / portal."""
import gc
import weakref

import pytest
import torch
from torch import nn

from pipe_b200 import Pipe, is_checkpointing, is_recomputing
from pipe_b200.autograd.stream import CPUStream
from pipe_b200.microbatch import Batch
from pipe_b200.skip import (Namespace, Portal, SkipTracker, SkipTrackerThroughPotals,
                            current_skip_tracker, inspect_skip_layout, pop, skippable, stash,
                            use_skip_tracker, verify_skippables)
from pipe_b200.skip.layout import SkipLayout


@skippable(stash=["skip"])
class Stash(nn.Module):
    def forward(self, x):
        yield stash("skip", x)
        return x * 2


@skippable(pop=["skip"])
class Pop(nn.Module):
    def forward(self, x):
        skip = yield pop("skip")
        return x + skip


class Pass(nn.Module):
    def forward(self, x):
        return x + 0.0


# ---- plain (un-piped) use -------------------------------------------------------
def test_works_outside_pipe():
    seq = nn.Sequential(Stash(), Pass(), Pop())
    x = torch.rand(2, 2)
    assert torch.allclose(seq(x), x * 3)


def test_namespace_ordering_and_identity():
    a, b = Namespace(), Namespace()
    assert a == a and a != b and (a < b or b < a)
    assert hash(a) != hash(b)
    assert isinstance(None, Namespace)
    assert sorted([(a, "x"), (a, "a")])[0][1] == "a"


def test_isolate():
    ns1, ns2 = Namespace(), Namespace()
    seq = nn.Sequential(Stash().isolate(ns1), Stash().isolate(ns2), Pop().isolate(ns2), Pop().isolate(ns1))
    verify_skippables(seq)
    x = torch.ones(1)
    # s1 = 1, h=2 ; s2 = 2, h=4 ; pop ns2: 4+2=6 ; pop ns1: 6+1=7
    assert seq(x).item() == 7


# ---- runtime declaration errors -------------------------------------------------
def test_undeclared_stash_pop_and_missing():
    @skippable()
    class BadStash(nn.Module):
        def forward(self, x):
            yield stash("skip", x)
            return x

    with pytest.raises(RuntimeError, match="has not been declared as stashable"):
        BadStash()(torch.rand(1))

    @skippable()
    class BadPop(nn.Module):
        def forward(self, x):
            yield pop("skip")
            return x

    with pytest.raises(RuntimeError, match="has not been declared as poppable"):
        BadPop()(torch.rand(1))

    @skippable(stash=["skip"])
    class Forgets(nn.Module):
        def forward(self, x):
            return x

    with pytest.raises(RuntimeError, match="must be stashed but have not"):
        Forgets()(torch.rand(1))

    @skippable(pop=["skip"])
    class ForgetsPop(nn.Module):
        def forward(self, x):
            return x

    with use_skip_tracker(SkipTracker()):
        with pytest.raises(RuntimeError, match="has not been stashed"):
            ForgetsPop()(torch.rand(1))
        current_skip_tracker().save(Batch(torch.rand(1)), None, "skip", torch.rand(1))
        with pytest.raises(RuntimeError, match="must be popped but have not"):
            ForgetsPop()(torch.rand(1))

    @skippable()
    class Junk(nn.Module):
        def forward(self, x):
            yield 42
            return x

    with pytest.raises(TypeError, match="not a command from @skippable"):
        Junk()(torch.rand(1))


def test_stash_none_and_no_yield():
    @skippable(stash=["skip"])
    class StashNone(nn.Module):
        def forward(self, x):
            yield stash("skip", None)
            return x

    @skippable(pop=["skip"])
    class PopNone(nn.Module):
        def forward(self, x):
            s = yield pop("skip")
            assert s is None
            return x

    seq = nn.Sequential(StashNone(), PopNone())
    seq(torch.rand(1))
    pipe = Pipe(nn.Sequential(StashNone(), Pass(), PopNone()), chunks=2)
    pipe(torch.rand(2, 1))


# ---- static verification --------------------------------------------------------
def test_verify_matrix():
    verify_skippables(nn.Sequential(Stash(), Pop()))
    with pytest.raises(TypeError, match="no module declared 'skip' as poppable but stashed"):
        verify_skippables(nn.Sequential(Stash()))
    with pytest.raises(TypeError, match="declared 'skip' as poppable but it was not stashed"):
        verify_skippables(nn.Sequential(Pop()))
    with pytest.raises(TypeError, match="redeclared 'skip' as stashable"):
        verify_skippables(nn.Sequential(Stash(), Stash(), Pop()))
    with pytest.raises(TypeError, match="redeclared 'skip' as poppable"):
        verify_skippables(nn.Sequential(Stash(), Pop(), Pop()))

    @skippable(stash=["skip"], pop=["skip"])
    class Both(nn.Module):
        pass

    with pytest.raises(TypeError, match="both as stashable and as poppable"):
        verify_skippables(nn.Sequential(Both()))
    with pytest.raises(TypeError, match="do not match"):
        Pipe(nn.Sequential(Pop(), Stash()))
    # several problems are reported together
    with pytest.raises(TypeError) as info:
        verify_skippables(nn.Sequential(Pop(), Stash(), Stash()))
    assert str(info.value).count("* ") >= 3


# ---- layout -----------------------------------------------------------------------
def test_layout_routes_and_policy():
    p = [nn.Sequential(Stash()), nn.Sequential(Pass()), nn.Sequential(Pop())]
    layout = inspect_skip_layout(p)
    assert layout.requires_copy(None, "skip")
    assert list(layout.copy_policy(2)) == [(0, None, "skip")]
    assert list(layout.copy_policy(1)) == []
    same = inspect_skip_layout([nn.Sequential(Stash(), Pop())])
    assert not same.requires_copy(None, "skip")
    assert list(same.copy_policy(0)) == []
    assert not SkipLayout(1, {}).requires_copy(None, "nothing")


# ---- tracker / portal ---------------------------------------------------------------
def test_plain_tracker():
    t = SkipTracker()
    b = Batch(torch.rand(1))
    x = torch.rand(1)
    t.save(b, None, "k", x)
    assert t.load(b, None, "k") is x
    with pytest.raises(KeyError):
        t.load(b, None, "k")
    with pytest.raises(TypeError):
        t.copy(b, CPUStream, CPUStream, None, "k")


def test_portal_tensor_life_and_grad_mailbox():
    x = torch.rand(1)
    portal = Portal(x, tensor_life=2)
    assert portal.use_tensor() is x and portal.tensor is x
    assert portal.use_tensor() is x and portal.tensor is None
    with pytest.raises(RuntimeError, match="has been removed"):
        portal.use_tensor()
    with pytest.raises(RuntimeError, match="removed or never set"):
        portal.use_grad()
    portal.put_grad(x)
    assert portal.use_grad() is x
    assert Portal(x, tensor_life=0).tensor is None


def test_portal_blue_orange_copy_gradients():
    layout = SkipLayout(2, {(None, "k"): (0, 1)})
    tracker = SkipTrackerThroughPotals(layout)
    skip = torch.rand(2, requires_grad=True)
    main = torch.rand(2, requires_grad=True)
    b = Batch(main * 1.0)
    tracker.save(b, None, "k", skip * 1.0)
    assert (None, "k") in tracker.portals
    tracker.copy(b, CPUStream, CPUStream, None, "k")
    got = tracker.load(b, None, "k")
    (b[0] * got).sum().backward()
    assert torch.allclose(skip.grad, main.detach()) and torch.allclose(main.grad, skip.detach())
    assert tracker.portals[(None, "k")].tensor is None        # dropped after its last use


def test_portal_life_under_checkpointing():
    layout = SkipLayout(2, {(None, "k"): (0, 1)})
    tracker = SkipTrackerThroughPotals(layout)
    from pipe_b200.autograd.checkpoint import enable_checkpointing
    with enable_checkpointing():
        tracker.save(Batch(torch.rand(1)), None, "k", torch.rand(1))
    assert tracker.portals[(None, "k")].tensor_life == 2       # 3 planned uses, blue took one


# ---- inside Pipe ----------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["never", "always", "except_last"])
def test_skip_across_stages_matches_plain(mode):
    torch.manual_seed(0)

    def build():
        return nn.Sequential(nn.Linear(3, 3), Stash(), nn.Linear(3, 3), nn.Tanh(), Pop(), nn.Linear(3, 1))

    seq = build()
    ref = build()
    ref.load_state_dict(seq.state_dict())
    x = torch.rand(6, 3)
    ref(x).sum().backward()

    pipe = Pipe(seq, chunks=3, checkpoint=mode)
    assert len(pipe.partitions) == 6                         # every CPU child is a stage: skip crosses 3
    out = pipe(x).local_value()
    assert torch.allclose(out, ref(x), atol=1e-6)
    out.sum().backward()
    for p, q in zip(pipe.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-6)


def test_skip_tensors_do_not_leak():
    refs = []

    @skippable(stash=["skip"])
    class S(nn.Module):
        def forward(self, x):
            y = x * 2
            refs.append(weakref.ref(y))
            yield stash("skip", y)
            return x

    pipe = Pipe(nn.Sequential(S(), Pass(), Pop()), chunks=2, checkpoint="always")
    x = torch.rand(2, 1, requires_grad=True)
    out = pipe(x).local_value()
    out.sum().backward()
    del out
    gc.collect()
    assert all(r() is None for r in refs)


def test_skippable_repr_and_multi_input():
    assert repr(Stash()).startswith("@skippable(")

    @skippable(stash=["a"])
    class Two(nn.Module):
        def forward(self, x, y):
            yield stash("a", y)
            return x + y, y

    @skippable(pop=["a"])
    class Back(nn.Module):
        def forward(self, x, y):
            a = yield pop("a")
            return x + a

    pipe = Pipe(nn.Sequential(Two(), Back()), chunks=2)
    x, y = torch.ones(2, 1), torch.ones(2, 1)
    assert torch.equal(pipe(x, y).local_value(), torch.full((2, 1), 3.0))
