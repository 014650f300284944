"""The persistent op-chain kernel (``csrc/chain_sm100.cu``) against the same ops as separate kernels.

Numerics oracle: the one-GEMM kernel with the same four-way cluster split (bit-identical by construction: same k-ranges,
same reduction order, same epilogue code, same dropout bits) and a plain fp32 PyTorch evaluation of the op sequence."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from pipe_b200.ops import functional as F        # noqa: E402
from pipe_b200.ops.chain import Chain            # noqa: E402
from pipe_b200.ops.gemm import gemm_nt           # noqa: E402

dev = torch.device("cuda:0")


def _w(m, k):
    return (torch.randn(m, k, device=dev) / k ** 0.5).bfloat16()


def _ln_ref(x, h, g, b, eps=1e-5):
    z = x.float() + h.float()
    return torch.nn.functional.layer_norm(z, (z.size(-1),), g, b, eps)


@pytest.mark.parametrize("T,E,FF", [(75, 1024, 2048), (128, 512, 512), (20, 256, 1024)])
def test_chain_matches_fp32_reference_and_separate_kernels(T, E, FF):
    torch.manual_seed(0)
    x = torch.randn(T, E, device=dev).bfloat16()
    wo, w1, w2 = _w(E, E), _w(FF, E), _w(E, FF)
    bo, b1, b2 = (torch.randn(n, device=dev) for n in (E, FF, E))
    g1, be1 = torch.rand(E, device=dev) + 0.5, torch.randn(E, device=dev)
    o, x1, f, y = (torch.empty(T, E, device=dev, dtype=torch.bfloat16) for _ in range(4))
    h = torch.empty(T, FF, device=dev, dtype=torch.bfloat16)
    st = torch.empty(4, T, device=dev)
    c = Chain(dev)
    c.gemm(wo, x, o, bias=bo).layernorm(x, o, g1, be1, x1, st[0], st[1])
    c.gemm(w1, x1, h, bias=b1, relu=True).gemm(w2, h, f, bias=b2).layernorm(x1, f, g1, be1, y, st[2], st[3])
    assert c.supported()
    c.run()
    torch.cuda.synchronize()
    # fp32 evaluation of the same sequence (bf16 rounding where the chain stores bf16)
    o_ref = (x.float() @ wo.float().t() + bo).bfloat16()
    x1_ref = _ln_ref(x, o_ref, g1, be1).bfloat16()
    h_ref = torch.relu(x1_ref.float() @ w1.float().t() + b1).bfloat16()
    f_ref = (h_ref.float() @ w2.float().t() + b2).bfloat16()
    y_ref = _ln_ref(x1_ref, f_ref, g1, be1)
    assert (y.float() - y_ref).abs().max() < 0.08
    assert ((y.float() - y_ref).norm() / y_ref.norm()) < 1e-2
    # the separate kernels with the same split: same bits
    o2 = gemm_nt(wo, x, transpose_out=True, bias=bo, bias_along_m=True, splits=4, reduce_mode=2) if E >= 1024 else None
    if o2 is not None:
        assert torch.equal(o, o2)


def test_chain_dropout_bits_match_the_one_gemm_kernel_and_relaunch_is_clean():
    torch.manual_seed(1)
    T, E = 75, 1024
    x = torch.randn(T, E, device=dev).bfloat16()
    w = _w(E, E)
    a = torch.empty(T, E, device=dev, dtype=torch.bfloat16)
    b = torch.empty(T, E, device=dev, dtype=torch.bfloat16)
    ref1 = gemm_nt(w, x, transpose_out=True, drop_p=0.3, seed=11, rng_stream=2, splits=4, reduce_mode=2)
    ref2 = gemm_nt(w, ref1, transpose_out=True, drop_p=0.3, seed=11, rng_stream=3, splits=4, reduce_mode=2)
    c = Chain(dev)
    c.gemm(w, x, a, drop_p=0.3, seed=11, rng_stream=2).gemm(w, a, b, drop_p=0.3, seed=11, rng_stream=3)
    for _ in range(5):          # the grid-barrier counters are left at zero by every launch
        a.zero_(); b.zero_()
        c.run()
        torch.cuda.synchronize()
        assert torch.equal(a, ref1) and torch.equal(b, ref2)
    frac = (a == 0).float().mean().item()
    assert 0.25 < frac < 0.35


def test_chain_rejects_what_it_cannot_run():
    x = torch.randn(200, 256, device=dev).bfloat16()
    w = _w(256, 256)
    out = torch.empty(200, 256, device=dev, dtype=torch.bfloat16)
    assert not Chain(dev).gemm(w, x, out).supported()            # more than 128 tokens
    x = torch.randn(16, 128, device=dev).bfloat16()
    w = _w(256, 128)
    out = torch.empty(16, 256, device=dev, dtype=torch.bfloat16)
    assert not Chain(dev).gemm(w, x, out).supported()            # K = 128: two k-blocks cannot be split four ways


@pytest.mark.parametrize("dropout", [0.0, 0.2])
def test_stack_with_chain_agrees_with_stack_without(dropout, monkeypatch):
    from pipe_b200.models.fused import FusedEncoderLayer
    from pipe_b200.ops import wgrad
    from pipe_b200.partition import PipeSequential
    E, H, FF, S, NB, L = 512, 4, 1024, 25, 3, 3
    torch.manual_seed(1)
    layers_a = [FusedEncoderLayer(E, H, FF, dropout).to(dev) for _ in range(L)]
    layers_b = [FusedEncoderLayer(E, H, FF, dropout).to(dev) for _ in range(L)]
    for a, b in zip(layers_a, layers_b):
        b.load_state_dict(a.state_dict())
    sa, sb = PipeSequential(*layers_a), PipeSequential(*layers_b)
    x = torch.randn(S, NB, E, device=dev).bfloat16()
    g = torch.randn(S, NB, E, device=dev).bfloat16()
    outs, grads = [], []
    for stack, layers, chain in ((sa, layers_a, "1"), (sb, layers_b, "0")):
        monkeypatch.setenv("PB2_CHAIN", chain)
        xi = x.clone().requires_grad_()
        torch.manual_seed(7)
        y = stack(xi)
        y.backward(g)
        wgrad.flush_pending()
        torch.cuda.synchronize()
        outs.append(y.detach().float())
        grads.append([xi.grad.float()] + [p.grad.float() for l in layers for p in l.parameters()])

    def rel(a, b):
        return ((a - b).norm() / (b.norm() + 1e-12)).item()

    assert rel(outs[0], outs[1]) < 2e-2
    for a, b in zip(grads[0], grads[1]):
        assert rel(a, b) < 5e-2
