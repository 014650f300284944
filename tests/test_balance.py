"""Balancers and partition_model."""
import time

import pytest
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.balance import balance_by_size, balance_by_time, balance_cost, blockpartition
from pipe_b200.utils import partition_model


def test_blockpartition_basic():
    assert blockpartition.solve([1, 2, 3, 4, 5, 6], partitions=2) == [[1, 2, 3, 4], [5, 6]]
    assert blockpartition.solve([1, 1, 1, 1], partitions=4) == [[1], [1], [1], [1]]
    assert blockpartition.solve([5], partitions=1) == [[5]]
    assert [sum(b) for b in blockpartition.solve([4, 1, 1, 1, 1, 4], 3)] == [4, 4, 4]


def test_blockpartition_errors():
    with pytest.raises(ValueError):
        blockpartition.solve([1, 2], partitions=0)
    with pytest.raises(ValueError):
        blockpartition.solve([1], partitions=2)


def test_blockpartition_is_optimal_on_random_inputs():
    import itertools
    import random
    rnd = random.Random(0)
    for _ in range(30):
        n = rnd.randint(3, 8)
        k = rnd.randint(1, n)
        seq = [rnd.randint(1, 9) for _ in range(n)]
        got = max(sum(b) for b in blockpartition.solve(seq, k))
        best = min(max(sum(seq[a:b]) for a, b in zip((0,) + cuts, cuts + (n,)))
                   for cuts in itertools.combinations(range(1, n), k - 1))
        assert got == best


def test_balance_cost():
    assert balance_cost([1, 1, 1, 1, 1, 1], 2) == [3, 3]
    assert balance_cost([10, 1, 1, 1], 2) == [1, 3]


def test_balance_by_time_cpu():
    class Delay(nn.Module):
        def __init__(self, s):
            super().__init__()
            self.s = s

        def forward(self, x):
            time.sleep(self.s)
            return x

    model = nn.Sequential(*[Delay(s) for s in (0.01, 0.01, 0.01, 0.03)])
    balance = balance_by_time(2, model, torch.rand(1, requires_grad=True), timeout=0, device="cpu")
    assert balance == [3, 1]


def test_balance_by_time_rejects_existing_grads():
    lin = nn.Linear(1, 1)
    lin(torch.rand(1, 1)).sum().backward()
    with pytest.raises(ValueError, match="already has gradient"):
        balance_by_time(1, nn.Sequential(lin), torch.rand(1, 1), device="cpu")


def test_balance_by_size_needs_cuda():
    with pytest.raises(ValueError, match="only CUDA"):
        balance_by_size(1, nn.Sequential(nn.Linear(1, 1)), torch.rand(1, 1), device="cpu")


@pytest.mark.gpu
def test_balance_by_size_cuda():
    model = nn.Sequential(nn.Linear(1024, 1024), nn.Linear(1024, 16), nn.Linear(16, 16))
    balance = balance_by_size(2, model, torch.rand(8, 1024), chunks=2, param_scale=4.0, device="cuda")
    assert balance == [1, 2]


def test_partition_model_then_pipe():
    flat = nn.Sequential(nn.Linear(2, 2), nn.ReLU(), nn.Linear(2, 2), nn.ReLU())
    nested = partition_model(flat, [2, 2], devices=["cpu", "cpu"])
    assert len(nested) == 2 and all(isinstance(g, nn.Sequential) for g in nested)
    with pytest.raises(ValueError):
        partition_model(flat, [1, 1], devices=["cpu", "cpu"])
    with pytest.raises(IndexError):
        partition_model(flat, [2, 2], devices=["cpu"])
    pipe = Pipe(nested, chunks=2)
    assert pipe(torch.rand(4, 2)).local_value().shape == (4, 2)
