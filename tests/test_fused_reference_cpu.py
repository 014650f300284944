"""The fused model's plain-PyTorch path (``reference_forward``): taken on non-CUDA devices and used as the
kernels' oracle.  Same parameter names as the eager model (SURVEY.md §7.1 "each with a torch-eager fallback")."""
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.models import TransformerConfig, build_layers


def _cfg(dropout=0.0):
    return TransformerConfig(ntokens=100, emsize=32, nhid=48, nlayers=2, nhead=4, dropout=dropout, bptt=6, batch=8, chunks=4)


def test_fused_modules_on_cpu_equal_the_eager_model():
    cfg = _cfg()
    fused = nn.Sequential(*build_layers(cfg, impl="fused", seed=3))
    eager = nn.Sequential(*build_layers(cfg, impl="eager", seed=3))
    assert [k for k, _ in fused.named_parameters()] == [k for k, _ in eager.named_parameters()]
    eager.load_state_dict(fused.state_dict())
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    want = eager(tok)
    got = fused(tok)
    assert got.shape == (cfg.batch, cfg.bptt, cfg.ntokens)
    assert torch.allclose(got, want, atol=1e-6)
    want.square().mean().backward()
    got.square().mean().backward()
    for (k, p), q in zip(fused.named_parameters(), eager.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-6), k


def test_fused_loss_head_on_cpu():
    cfg = _cfg()
    layers = build_layers(cfg, impl="fused", seed=1)
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    tgt = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    x = tok
    for m in layers[:-1]:
        x = m(x)
    logits = layers[-1](x)
    per_seq = layers[-1](x, tgt)
    want = torch.nn.functional.cross_entropy(logits.reshape(-1, cfg.ntokens), tgt.reshape(-1), reduction="none")
    assert torch.allclose(per_seq, want.view(cfg.batch, cfg.bptt).sum(1), atol=1e-5)


def test_pipe_over_fused_modules_on_cpu_with_targets_and_checkpointing():
    cfg = _cfg(dropout=0.1)
    layers = build_layers(cfg, impl="fused", seed=2)
    pipe = Pipe(nn.Sequential(*layers), chunks=cfg.chunks, checkpoint="always", balance=[2, 2], devices=["cpu", "cpu"])
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    tgt = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    loss = pipe(tok, tgt).local_value()          # targets ride through the stages to the fused loss head
    assert loss.shape == (cfg.batch,)
    (loss.sum() / tok.numel()).backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in pipe.parameters())
    pipe.close()
