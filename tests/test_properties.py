"""Property-based checks (hypothesis) of the pure-Python pieces everything else leans on: the clock
schedule, scatter / gather, and the block partition solver."""
import itertools

import torch
from hypothesis import given, settings, strategies as st

from pipe_b200.balance import blockpartition
from pipe_b200.microbatch import Batch, gather, scatter
from pipe_b200.runtime.schedule import bubble_fraction, clock_cycles


@given(m=st.integers(1, 12), n=st.integers(1, 9))
def test_clock_schedule_visits_every_cell_once_in_dependency_order(m, n):
    ticks = [list(t) for t in clock_cycles(m, n)]
    assert len(ticks) == m + n - 1
    seen = {}
    for k, tick in enumerate(ticks):
        assert len({j for _, j in tick}) == len(tick)            # one cell per stage per tick
        for i, j in tick:
            assert 0 <= i < m and 0 <= j < n and i + j == k      # anti-diagonal k
            assert (i, j) not in seen
            seen[(i, j)] = k
    assert len(seen) == m * n
    for (i, j), k in seen.items():
        if j:
            assert seen[(i, j - 1)] < k                          # previous stage of the same micro-batch first
        if i:
            assert seen[(i - 1, j)] < k                          # previous micro-batch on the same stage first
    assert abs(bubble_fraction(m, n) - (n - 1) / (m + n - 1)) < 1e-12


@settings(deadline=None, max_examples=60)
@given(rows=st.integers(1, 40), cols=st.integers(1, 5), chunks=st.integers(1, 12))
def test_scatter_gather_round_trip_matches_tensor_chunk(rows, cols, chunks):
    x = torch.arange(rows * cols, dtype=torch.float32).view(rows, cols)
    flag = torch.arange(rows)
    batches = scatter(x, flag, chunks=chunks)
    assert len(batches) == len(x.chunk(chunks)) <= chunks
    for b, piece in zip(batches, x.chunk(chunks)):
        assert torch.equal(b[0], piece) and b[1].size(0) == piece.size(0)
    got_x, got_flag = gather(batches)
    assert torch.equal(got_x, x) and torch.equal(got_flag, flag)
    assert torch.equal(gather([Batch(b[0]) for b in batches]), x)


@settings(deadline=None, max_examples=80)
@given(costs=st.lists(st.integers(1, 30), min_size=1, max_size=9), parts=st.integers(1, 5))
def test_block_partition_is_optimal_for_the_bottleneck(costs, parts):
    if parts > len(costs):
        return
    blocks = blockpartition.solve(costs, parts)
    assert len(blocks) == parts and all(blocks) and [c for b in blocks for c in b] == costs
    got = max(sum(b) for b in blocks)
    best = min(max(sum(costs[a:b]) for a, b in zip((0,) + cut, cut + (len(costs),)))
               for cut in itertools.combinations(range(1, len(costs)), parts - 1))
    assert got == best
