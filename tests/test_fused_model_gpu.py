"""Fused modules vs their stock torch.nn counterparts (same weights), alone and inside Pipe."""
import copy

import pytest
import torch
import torch.nn.functional as F
from torch import nn

pytestmark = pytest.mark.gpu
dev = "cuda"


def rel(got, want):
    got, want = got.float(), want.float()
    return ((got - want).norm() / (want.norm() + 1e-12)).item()


def test_encoder_layer_matches_torch_layer():
    from pipe_b200.models.fused import FusedEncoderLayer
    from pipe_b200.ops import wgrad
    torch.manual_seed(0)
    E, H, FF, S, NB = 512, 8, 1024, 25, 3
    ref = nn.TransformerEncoderLayer(E, H, FF, 0.0).to(dev)
    fused = FusedEncoderLayer(E, H, FF, 0.0).to(dev)
    fused.load_state_dict(ref.state_dict())                    # same parameter names
    x = torch.randn(S, NB, E, device=dev)
    xr = x.clone().requires_grad_()
    xf = x.bfloat16().requires_grad_()
    yr = ref(xr)
    yf = fused(xf)
    assert yf.shape == yr.shape and yf.dtype == torch.bfloat16
    assert rel(yf, yr) < 2e-2
    g = torch.randn_like(yr)
    yr.backward(g)
    yf.backward(g.bfloat16())
    wgrad.flush_pending()
    assert rel(xf.grad, xr.grad) < 4e-2
    for (name, pf), pr in zip(fused.named_parameters(), ref.parameters()):
        assert pf.grad is not None, name
        assert rel(pf.grad, pr.grad) < 5e-2, (name, rel(pf.grad, pr.grad))


def test_causal_flag():
    from pipe_b200.models.fused import FusedEncoderLayer
    torch.manual_seed(0)
    E, H, S, NB = 256, 4, 16, 2
    ref = nn.TransformerEncoderLayer(E, H, 512, 0.0).to(dev)
    fused = FusedEncoderLayer(E, H, 512, 0.0, causal=True).to(dev)
    fused.load_state_dict(ref.state_dict())
    x = torch.randn(S, NB, E, device=dev)
    mask = nn.Transformer.generate_square_subsequent_mask(S, device=dev)
    assert rel(fused(x.bfloat16()), ref(x, src_mask=mask)) < 2e-2


def _tiny_cfg(dropout):
    from pipe_b200.models import TransformerConfig
    return TransformerConfig(ntokens=1000, emsize=256, nhid=512, nlayers=2, nhead=4, dropout=dropout,
                             bptt=25, batch=20, chunks=8)


def test_full_model_in_pipe_matches_eager_model():
    from pipe_b200 import Pipe
    from pipe_b200.models import build_stages
    cfg = _tiny_cfg(0.0)
    eager = build_stages(cfg, [0], impl="eager")
    fused = build_stages(cfg, [0], impl="fused")
    fused.load_state_dict(eager.state_dict())
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
    targets = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)

    want = eager(tokens)
    loss_ref = F.cross_entropy(want.reshape(-1, cfg.ntokens), targets.reshape(-1))
    loss_ref.backward()

    pipe = Pipe(fused, chunks=cfg.chunks, checkpoint="never")
    # (a) logits path: API parity with the reference's Decoder
    got = pipe(tokens).local_value()
    assert got.shape == want.shape
    assert rel(got, want) < 3e-2
    # (b) fused-loss path: targets ride through the pipe
    per_seq = pipe(tokens, targets).local_value()
    assert per_seq.shape == (cfg.batch,)
    loss = per_seq.sum() / (cfg.batch * cfg.bptt)
    assert abs(loss.item() - loss_ref.item()) < 2e-2 * abs(loss_ref.item())
    loss.backward()
    for (name, pf), pr in zip(pipe.named_parameters(), eager.parameters()):
        assert pf.grad is not None, name
        assert rel(pf.grad, pr.grad) < 8e-2, (name, rel(pf.grad, pr.grad))


def test_logits_path_backward():
    from pipe_b200 import Pipe
    from pipe_b200.models import build_stages
    cfg = _tiny_cfg(0.0)
    eager = build_stages(cfg, [0], impl="eager")
    fused = build_stages(cfg, [0], impl="fused")
    fused.load_state_dict(eager.state_dict())
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
    targets = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device=dev)
    F.cross_entropy(eager(tokens).reshape(-1, cfg.ntokens), targets).backward()
    pipe = Pipe(fused, chunks=4, checkpoint="except_last")
    out = pipe(tokens).local_value()
    F.cross_entropy(out.reshape(-1, cfg.ntokens).float(), targets).backward()
    for (name, pf), pr in zip(pipe.named_parameters(), eager.parameters()):
        assert rel(pf.grad, pr.grad) < 8e-2, (name, rel(pf.grad, pr.grad))


def test_dropout_recompute_is_bit_identical():
    """checkpoint='always' must reproduce the 'never' gradients with dropout on: the Philox seeds
    come from the CPU generator that checkpointing captures and replays."""
    from pipe_b200 import Pipe
    from pipe_b200.models import build_stages
    cfg = _tiny_cfg(0.2)
    grads = {}
    for mode in ("never", "always"):
        model = build_stages(cfg, [0], impl="fused", seed=3)
        pipe = Pipe(model, chunks=4, checkpoint=mode)
        torch.manual_seed(123)
        tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
        targets = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
        per_seq = pipe(tokens, targets).local_value()
        (per_seq.sum() / 500).backward()
        grads[mode] = {n: p.grad.clone() for n, p in pipe.named_parameters()}
    for n in grads["never"]:
        assert rel(grads["always"][n], grads["never"][n]) < 1e-3, n       # atomics reorder sums; masks identical


def test_training_reduces_loss_with_fused_optimizer():
    from pipe_b200.models.trainer import LMTrainer
    cfg = _tiny_cfg(0.1)
    tr = LMTrainer(cfg, n_stages=1, impl="fused", checkpoint="except_last", lr=2e-3)
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
    targets = torch.randint(0, cfg.ntokens, (cfg.batch * cfg.bptt,), device=dev)
    losses = [tr.step(tokens, targets).item() for _ in range(12)]
    assert all(l == l for l in losses)
    assert losses[-1] < losses[0] - 0.3, losses


def test_state_dict_moves_between_eager_and_fused():
    from pipe_b200.models import build_stages
    cfg = _tiny_cfg(0.0)
    eager = build_stages(cfg, [0], impl="eager")
    fused = build_stages(cfg, [0], impl="fused")
    assert list(eager.state_dict().keys()) == list(fused.state_dict().keys())
    fused.load_state_dict(eager.state_dict())
    eager.load_state_dict(fused.state_dict())


def test_skip_connection_across_fused_stages_and_inline_toggle():
    """@skippable stash/pop around fused layers in a 2-stage-on-one-GPU pipe, both dispatch modes."""
    from pipe_b200 import Pipe
    from pipe_b200.models import build_layer
    from pipe_b200.skip import pop, skippable, stash

    @skippable(stash=["emb"])
    class StashEmb(nn.Module):
        def forward(self, x):
            yield stash("emb", x)
            return x

    @skippable(pop=["emb"])
    class PopAdd(nn.Module):
        def forward(self, x):
            emb = yield pop("emb")
            return x + emb

    cfg = _tiny_cfg(0.0)
    outs = []
    for inline in (True, False):
        layers = [build_layer(cfg, k, "fused", 0, seed=5) for k in range(cfg.nlayers + 2)]
        seq = nn.Sequential(layers[0], StashEmb().cuda(), layers[1], layers[2], PopAdd().cuda(), layers[3])
        pipe = Pipe(seq, chunks=4, checkpoint="always", balance=[3, 3], devices=["cuda:0", "cuda:0"])
        assert len(pipe.partitions) == 2 and pipe._skip_layout.requires_copy(None, "emb")
        pipe.pipeline.inline_last_cell = inline
        tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev, generator=torch.Generator(dev).manual_seed(1))
        out = pipe(tokens).local_value()
        out.float().sum().backward()
        outs.append((out.detach().clone(), pipe[0].encoder.weight.grad.clone()))
        pipe.close()
    assert rel(outs[0][0], outs[1][0]) < 1e-3
    assert rel(outs[0][1], outs[1][1]) < 1e-2


@pytest.mark.parametrize("dropout", [0.0, 0.2])
def test_layer_stack_is_bit_identical_to_per_layer_calls(dropout, monkeypatch):
    """A run of layers through ``pb2_encoder_stack_*`` (one native call per pass) launches the same
    kernels in the same order as per-layer calls: outputs, input gradient and every parameter
    gradient must be bit-identical, over two steps (the second reuses cached descriptors and the
    grown wgrad stashes), also with the extra pass-through input used for fused-loss targets.
    (With the opt-in op-chain kernel off: it splits every reduction four ways, the one-GEMM kernel picks per shape —
    ``tests/test_chain_gpu.py`` compares the two.)"""
    monkeypatch.setenv("PB2_CHAIN", "0")
    from pipe_b200.models.fused import EncoderStack, FusedEncoderLayer
    from pipe_b200.ops import wgrad
    from pipe_b200.partition import PipeSequential
    E, H, FF, S, NB, L = 256, 4, 512, 25, 3, 3
    torch.manual_seed(1)
    layers_a = [FusedEncoderLayer(E, H, FF, dropout).to(dev) for _ in range(L)]
    layers_b = [FusedEncoderLayer(E, H, FF, dropout).to(dev) for _ in range(L)]
    for a, b in zip(layers_a, layers_b):
        b.load_state_dict(a.state_dict())
    stacked = PipeSequential(*layers_a)
    plan = stacked._plan()
    assert len(plan) == 1 and isinstance(plan[0][0], EncoderStack)
    extra = torch.arange(NB * S, device=dev).view(NB, S)
    for step in range(2):
        x = torch.randn(S, NB, E, device=dev).bfloat16()
        g = torch.randn(S, NB, E, device=dev).bfloat16()
        xa = x.clone().requires_grad_()
        xb = x.clone().requires_grad_()
        torch.manual_seed(100 + step)
        ya, ea = stacked(xa, extra)
        torch.manual_seed(100 + step)
        hb = xb
        for l in layers_b:
            hb = l(hb)
        assert ea is extra
        if dropout == 0.0:
            assert torch.equal(ya, hb)
        ya.backward(g)
        hb.backward(g)
        wgrad.flush_pending()
        if dropout == 0.0:                      # with dropout the two paths draw different seeds per layer
            assert torch.equal(xa.grad, xb.grad)
            for a, b in zip(layers_a, layers_b):
                for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
                    if pa.dim() >= 2:
                        assert torch.equal(pa.grad, pb.grad), (step, n)       # one GEMM, same operands
                    else:
                        # bias / LayerNorm gradients are summed with fp32 atomics: order varies run to run
                        assert rel(pa.grad, pb.grad) < 1e-5, (step, n, rel(pa.grad, pb.grad))
        else:
            assert torch.isfinite(ya.float()).all() and torch.isfinite(xa.grad.float()).all()
        for l in layers_a + layers_b:
            for p in l.parameters():
                p.grad = None


def test_layer_stack_replays_dropout_under_checkpoint():
    """Checkpoint 'always' must give the gradients of 'never' for the stacked path too: the n seeds of a
    cell come from the cell's base and are replayed by the recompute."""
    from pipe_b200 import Pipe
    from pipe_b200.models.fused import FusedEncoderLayer
    from pipe_b200.ops import wgrad
    E, H, FF, S, N = 256, 4, 512, 16, 8
    torch.manual_seed(3)
    layers = [FusedEncoderLayer(E, H, FF, 0.3).to(dev) for _ in range(3)]
    x = torch.randn(N, S, E, device=dev).bfloat16()

    class SeqFirst(nn.Module):
        def forward(self, t):
            return t.transpose(0, 1).contiguous()

    from pipe_b200 import WithDevice
    pipe = Pipe(nn.Sequential(WithDevice(SeqFirst(), "cuda:0"), *layers, WithDevice(SeqFirst(), "cuda:0")), chunks=2, checkpoint="always")
    out = pipe(x.clone().requires_grad_()).local_value()
    # forward under no_grad + recompute under grad must produce the same output bits
    out.float().sum().backward()
    wgrad.flush_pending()
    g_always = [p.grad.clone() for p in pipe.parameters()]
    assert all(torch.isfinite(g).all() for g in g_always)
    pipe.close()


# ---------------------------------------------------------------------------------------------------
# the autograd contract of the fused modules (INTEGRATION.md "Weight gradients and autograd")
# ---------------------------------------------------------------------------------------------------
def test_gradient_accumulation_over_two_backward_passes():
    """Two backward passes without zero_grad() accumulate (the second wgrad GEMM adds), and the optimizer clips
    with the norm of the accumulated gradient, not with a stale or double-counted one."""
    from pipe_b200.models.fused import FusedEncoderLayer
    from pipe_b200.ops import wgrad
    from pipe_b200.ops.optim import FusedClipAdam
    torch.manual_seed(0)
    E, H, FF, S, NB = 256, 4, 512, 16, 4
    layer = FusedEncoderLayer(E, H, FF, 0.0).to(dev)
    opt = FusedClipAdam(layer, lr=1e-3, max_norm=1e9)
    xa = torch.randn(S, NB, E, device=dev).bfloat16()
    xb = torch.randn(S, NB, E, device=dev).bfloat16()
    single = {}
    for name, x in (("a", xa), ("b", xb)):
        opt.zero_grad()
        layer(x).float().pow(2).sum().backward()
        wgrad.flush_pending()
        single[name] = {n: p.grad.clone() for n, p in layer.named_parameters()}
    opt.zero_grad()
    layer(xa).float().pow(2).sum().backward()
    layer(xb).float().pow(2).sum().backward()
    wgrad.flush_pending()
    total_sq = 0.0
    for n, p in layer.named_parameters():
        want = single["a"][n] + single["b"][n]
        assert rel(p.grad, want) < 1e-3, (n, rel(p.grad, want))
        total_sq += float(p.grad.double().pow(2).sum())
    opt.step()
    assert abs(opt.grad_norm() - total_sq ** 0.5) / total_sq ** 0.5 < 1e-3
    # ... and a step() repeated without zero_grad() measures the same gradients again instead of adding up
    opt.step()
    assert abs(opt.grad_norm() - total_sq ** 0.5) / total_sq ** 0.5 < 1e-3


def test_stock_optimizer_and_set_to_none_work_with_fused_modules():
    """torch.optim.Adam + zero_grad(set_to_none=True): gradient buffers are re-created every step, the bf16
    weight shadows follow the in-place update through the version counter."""
    from pipe_b200.models.fused import FusedEncoderLayer
    torch.manual_seed(0)
    E, H, FF, S, NB = 256, 4, 512, 16, 4
    layers = nn.Sequential(*[FusedEncoderLayer(E, H, FF, 0.0).to(dev) for _ in range(2)])
    from pipe_b200.partition import PipeSequential
    model = PipeSequential(*layers)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    x = torch.randn(S, NB, E, device=dev).bfloat16()
    target = torch.randn(S, NB, E, device=dev)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        loss = (model(x).float() - target).pow(2).mean()
        loss.backward()
        assert all(p.grad is not None for p in model.parameters())
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0], losses


def test_autograd_grad_does_not_see_weight_gradients():
    """Documented limit: weight gradients are written to ``.grad`` by the deferred wgrad GEMM, outside the
    autograd graph, so ``torch.autograd.grad`` with respect to a fused weight yields nothing."""
    from pipe_b200.models.fused import FusedEncoderLayer
    from pipe_b200.ops import wgrad
    layer = FusedEncoderLayer(256, 4, 512, 0.0).to(dev)
    x = torch.randn(8, 2, 256, device=dev).bfloat16().requires_grad_()
    y = layer(x)
    gx, gw = torch.autograd.grad(y.float().sum(), [x, layer.linear1.weight], allow_unused=True)
    wgrad.flush_pending()
    assert gx is not None and gw is None


def test_logits_path_writes_into_one_buffer_and_gather_is_a_view():
    """Without targets the fused decoder returns logits (reference main.py:55).  Inside Pipe every micro-batch
    writes its logits into its slice of one mini-batch-wide buffer, so the gathered output aliases the first
    micro-batch's memory (no cat), matches the eager model, and gradients still flow."""
    from pipe_b200 import Pipe
    from pipe_b200.models import build_stages
    cfg = _tiny_cfg(0.0)
    eager = build_stages(cfg, [0], impl="eager")
    fused = build_stages(cfg, [0], impl="fused")
    fused.load_state_dict(eager.state_dict())
    tokens = torch.randint(0, cfg.ntokens, (cfg.batch, cfg.bptt), device=dev)
    pipe = Pipe(fused, chunks=4, checkpoint="never")
    out = pipe(tokens).local_value()
    want = eager(tokens)
    assert out.shape == want.shape == (cfg.batch, cfg.bptt, cfg.ntokens)
    assert rel(out, want) < 3e-2
    ld = (cfg.ntokens + 63) // 64 * 64
    assert out.stride() == (cfg.bptt * ld, ld, 1)                 # rows of the padded shared buffer: a view, not a copy
    out.float().sum().backward()
    want.sum().backward()
    for (name, pf), pr in zip(pipe.named_parameters(), eager.parameters()):
        assert rel(pf.grad, pr.grad) < 8e-2, (name, rel(pf.grad, pr.grad))
    pipe.close()
