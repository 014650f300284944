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


def test_fused_clip_adam_on_cpu_equals_clip_plus_torch_adam():
    from pipe_b200.ops.optim import FusedClipAdam
    cfg = _cfg()
    fused = nn.Sequential(*build_layers(cfg, impl="fused", seed=3))
    eager = nn.Sequential(*build_layers(cfg, impl="eager", seed=3))
    eager.load_state_dict(fused.state_dict())
    o1 = FusedClipAdam(fused, lr=1e-2, max_norm=0.5)
    o2 = torch.optim.Adam(eager.parameters(), lr=1e-2)
    tok = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt))
    for _ in range(3):
        o1.zero_grad()
        o2.zero_grad()
        fused(tok).square().mean().backward()
        eager(tok).square().mean().backward()
        norm = torch.nn.utils.clip_grad_norm_(eager.parameters(), 0.5)
        o1.step()
        o2.step()
        assert abs(o1.grad_norm() - float(norm)) < 1e-5
    for p, q in zip(fused.parameters(), eager.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
