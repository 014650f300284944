"""BASELINE.json config 1: the tutorial Transformer (tiny) over 2 CPU partitions, chunks=4 —
outputs and gradients must equal the un-piped nn.Sequential in all three modes."""
import copy

import pytest
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.models import CONFIGS, build_stages, count_parameters
from pipe_b200.models.transformer import stage_balance


def _tiny(dropout=0.0):
    cfg = copy.copy(CONFIGS["tiny"])
    cfg.dropout = dropout
    return cfg


@pytest.mark.parametrize("mode", ["never", "except_last", "always"])
def test_tiny_transformer_two_cpu_partitions(mode):
    cfg = _tiny()
    torch.manual_seed(0)
    model = build_stages(cfg, ["cpu", "cpu"])
    ref = copy.deepcopy(model)

    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    targets = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,))
    crit = nn.CrossEntropyLoss()

    want = ref(tokens)
    crit(want.reshape(-1, cfg.ntokens), targets).backward()

    pipe = Pipe(model, chunks=cfg.chunks, checkpoint=mode)
    # two nested Sequentials on cpu: each is ONE child on cpu => 2 stages
    assert len(pipe.partitions) == 2
    got = pipe(tokens).local_value()
    assert got.shape == (cfg.batch, cfg.bptt, cfg.ntokens)
    assert torch.allclose(got, want, atol=1e-5)
    crit(got.reshape(-1, cfg.ntokens), targets).backward()

    for (name, p), q in zip(pipe.named_parameters(), ref.parameters()):
        assert p.grad is not None, name
        assert torch.allclose(p.grad, q.grad, atol=1e-5), name

    # state dict is keyed like the wrapped Sequential, plus the prefix
    assert list(pipe.state_dict().keys()) == ["partitions." + k for k in ref.state_dict().keys()]


def test_dropout_training_step_runs_with_checkpointing():
    cfg = _tiny(dropout=0.2)
    model = build_stages(cfg, ["cpu", "cpu"])
    pipe = Pipe(model, chunks=cfg.chunks, checkpoint="always")
    opt = torch.optim.Adam(pipe.parameters(), lr=1e-3)
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    targets = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,))
    losses = []
    for _ in range(3):
        opt.zero_grad()
        out = pipe(tokens).local_value()
        loss = nn.functional.cross_entropy(out.reshape(-1, cfg.ntokens), targets)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(pipe.parameters(), 0.5)
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]


def test_parameter_count_of_shipped_config_matches_reference():
    # README.md:570 -> 520,900,718 parameters for E=H=2048, L=16, V=28782
    E, H, L, V = 2048, 2048, 16, 28782
    per_layer = 4 * E * E + 4 * E + 2 * E * H + H + E + 4 * E
    assert V * E + L * per_layer + V * E + V == 520_900_718


def test_stage_balance_mirrors_reference_layout():
    assert stage_balance(16, 2) == [9, 9]
    assert stage_balance(16, 8) == [3, 2, 2, 2, 2, 2, 2, 3]
    assert stage_balance(8, 1) == [10]
    assert count_parameters(build_stages(_tiny(), ["cpu"])) > 0
