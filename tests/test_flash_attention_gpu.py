"""Flash attention (tcgen05 / TMEM / TMA, ``csrc/attention_flash.cu``) against fp32 scaled-dot-product attention
on the packed QKV layout of the encoder layer.  Reference call site: ``nn.MultiheadAttention`` inside
``nn.TransformerEncoderLayer`` (``main.py:148``), sequence lengths up to ``max_len = 5000`` (``main.py:59``)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
dev = "cuda"


def rel(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / (want.norm() + 1e-12)).item()


def _reference(qkv, NB, S, H, causal):
    T, E3 = qkv.shape
    E = E3 // 3
    D = E // H
    x = qkv.float().view(S, NB, 3, H, D)
    q, k, v = (x[:, :, i].permute(1, 2, 0, 3) for i in range(3))          # [NB, H, S, D]
    out = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
    scores = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    if causal:
        scores = scores.masked_fill(torch.ones(S, S, device=qkv.device).triu(1).bool(), float("-inf"))
    lse = torch.logsumexp(scores, dim=-1)                                   # [NB, H, S]
    return out.permute(2, 0, 1, 3).reshape(T, E), lse.reshape(NB * H, S)


@pytest.mark.parametrize("S,NB,H,D", [(128, 2, 4, 64), (200, 3, 2, 64), (512, 2, 2, 128), (2048, 1, 2, 64),
                                      (384, 1, 2, 256), (77, 2, 2, 128), (1000, 1, 1, 256)])
@pytest.mark.parametrize("causal", [False, True])
def test_flash_forward_matches_sdpa(S, NB, H, D, causal):
    from pipe_b200.ops import functional as F2
    torch.manual_seed(S + D)
    qkv = (torch.randn(S * NB, 3 * H * D, device=dev) * 0.7).bfloat16()
    out, lse = F2.flash_attention_fwd(qkv, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0)
    torch.cuda.synchronize()
    want, want_lse = _reference(qkv, NB, S, H, causal)
    assert rel(out, want) < 1.5e-2, rel(out, want)
    assert torch.allclose(lse, want_lse, atol=2e-2, rtol=1e-3), (lse - want_lse).abs().max().item()


def test_flash_forward_dropout_keeps_expectation_and_is_reproducible():
    from pipe_b200.ops import functional as F2
    S, NB, H, D = 256, 2, 2, 64
    torch.manual_seed(0)
    qkv = (torch.randn(S * NB, 3 * H * D, device=dev) * 0.5).bfloat16()
    a, _ = F2.flash_attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.25, seed=11, rng_stream=3)
    b, _ = F2.flash_attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.25, seed=11, rng_stream=3)
    c, _ = F2.flash_attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.25, seed=12, rng_stream=3)
    clean, _ = F2.flash_attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.0, seed=0, rng_stream=0)
    torch.cuda.synchronize()
    assert torch.equal(a, b) and not torch.equal(a, c)
    # dropout on the probabilities is unbiased: averaged over many seeds the output approaches the clean one
    acc = torch.zeros_like(clean, dtype=torch.float32)
    n = 24
    for s in range(n):
        o, _ = F2.flash_attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.25, seed=100 + s, rng_stream=3)
        acc += o.float()
    assert rel(acc / n, clean) < 0.12


@pytest.mark.parametrize("S,NB,H,D", [(128, 2, 4, 64), (200, 3, 2, 64), (512, 2, 2, 128), (1024, 1, 2, 64), (77, 2, 2, 128)])
@pytest.mark.parametrize("causal", [False, True])
def test_flash_backward_matches_autograd(S, NB, H, D, causal):
    from pipe_b200.ops import functional as F2
    torch.manual_seed(S * 3 + D)
    E = H * D
    qkv = (torch.randn(S * NB, 3 * E, device=dev) * 0.7).bfloat16()
    dout = (torch.randn(S * NB, E, device=dev) * 0.5).bfloat16()
    out, lse = F2.flash_attention_fwd(qkv, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0)
    dqkv = F2.flash_attention_bwd(qkv, out, dout, lse, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0)
    torch.cuda.synchronize()
    ref_in = qkv.float().requires_grad_()
    x = ref_in.view(S, NB, 3, H, D)
    q, k, v = (x[:, :, i].permute(1, 2, 0, 3) for i in range(3))
    want = F.scaled_dot_product_attention(q, k, v, is_causal=causal).permute(2, 0, 1, 3).reshape(S * NB, E)
    want.backward(dout.float())
    g = ref_in.grad.view(S * NB, 3, E)
    got = dqkv.float().view(S * NB, 3, E)
    for i, name in enumerate("qkv"):
        assert rel(got[:, i], g[:, i]) < 3e-2, (name, rel(got[:, i], g[:, i]))


def test_flash_backward_dropout_is_consistent_with_forward():
    """With V = identity-like probes the forward output exposes the dropped probabilities; the backward must
    have used the same mask: finite-difference of the loss along a random direction matches <grad, direction>."""
    from pipe_b200.ops import functional as F2
    S, NB, H, D = 256, 1, 2, 64
    E = H * D
    torch.manual_seed(5)
    qkv = (torch.randn(S * NB, 3 * E, device=dev) * 0.5).bfloat16()
    dout = (torch.randn(S * NB, E, device=dev) * 0.5).bfloat16()
    kw = dict(causal=True, drop_p=0.2, seed=77, rng_stream=1)
    out, lse = F2.flash_attention_fwd(qkv, NB, S, H, **kw)
    dqkv = F2.flash_attention_bwd(qkv, out, dout, lse, NB, S, H, **kw).float()
    direction = torch.randn_like(dqkv)
    direction[:, :2 * E] = 0                                  # move V only: the output is linear in V
    eps = 0.25
    plus, _ = F2.flash_attention_fwd((qkv.float() + eps * direction).bfloat16(), NB, S, H, **kw)
    minus, _ = F2.flash_attention_fwd((qkv.float() - eps * direction).bfloat16(), NB, S, H, **kw)
    fd = ((plus.float() - minus.float()) * dout.float()).sum().item() / (2 * eps)
    an = (dqkv * direction).sum().item()
    assert abs(fd - an) / (abs(an) + 1e-6) < 5e-2, (fd, an)


@pytest.mark.parametrize("S,E,H", [(128, 256, 4), (200, 256, 2), (640, 256, 4)])
def test_fused_layers_use_flash_attention_beyond_one_tile(S, E, H):
    """The encoder layer (per-layer call and the stage-level stack call) routes S >= 64 with head_dim 64 / 128 to
    the flash kernels — including sequences the single-tile kernels refuse (S > 128) — and matches
    nn.TransformerEncoderLayer."""
    from torch import nn

    from pipe_b200.models.fused import FusedEncoderLayer, _uses_flash
    from pipe_b200.ops import wgrad
    from pipe_b200.partition import PipeSequential
    assert _uses_flash(S, E // H)
    torch.manual_seed(0)
    NB, FF = 2, 512
    refs = [nn.TransformerEncoderLayer(E, H, FF, 0.0).to(dev) for _ in range(2)]
    fused = [FusedEncoderLayer(E, H, FF, 0.0).to(dev) for _ in range(2)]
    for f, r in zip(fused, refs):
        f.load_state_dict(r.state_dict())
    x = torch.randn(S, NB, E, device=dev)
    g = torch.randn(S, NB, E, device=dev)
    xr = x.clone().requires_grad_()
    yr = refs[1](refs[0](xr))
    yr.backward(g)
    for runner in ("stack", "single"):
        xf = x.bfloat16().requires_grad_()
        if runner == "stack":
            yf = PipeSequential(*fused)(xf)
        else:
            yf = fused[1](fused[0](xf))
        assert rel(yf, yr) < 3e-2, (runner, rel(yf, yr))
        yf.backward(g.bfloat16())
        wgrad.flush_pending()
        assert rel(xf.grad, xr.grad) < 6e-2, (runner, rel(xf.grad, xr.grad))
        for f, r in zip(fused, refs):
            for (name, pf), pr in zip(f.named_parameters(), r.parameters()):
                assert rel(pf.grad, pr.grad) < 6e-2, (runner, name, rel(pf.grad, pr.grad))
                pf.grad = None
