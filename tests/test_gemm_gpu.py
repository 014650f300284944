"""tcgen05 GEMM vs a plain fp32 PyTorch reference (SURVEY.md §4.4 'kernel numerics')."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b):
    return a.float() @ b.float().t()


def _close(got, want, rtol=2e-2, atol=None):
    want = want.float()
    got = got.float()
    atol = atol if atol is not None else 2e-2 * want.abs().max().item()
    assert torch.allclose(got, want, rtol=rtol, atol=atol), f"max err {(got - want).abs().max().item()} (scale {want.abs().max().item()})"


@pytest.mark.parametrize("M,N,K", [
    (128, 16, 64),          # one tile, one k-block
    (128, 80, 256),         # N not a power of two
    (4096, 75, 4096),       # B2 out-proj/FFN shape, swap-AB, split-K
    (12288, 75, 4096),      # B2 QKV
    (300, 50, 192),         # ragged M, ragged micro-batch, K = 3 blocks
    (4096, 4096, 500),      # wgrad shape: K tail (500 = 7*64 + 52), many N tiles
    (2048, 1024, 2048),     # reference-as-shipped micro-batch (1024 tokens)
    (28782, 75, 4096),      # vocab projection, M not a multiple of 128
])
def test_gemm_matches_fp32(M, N, K):
    from pipe_b200.ops.gemm import gemm_nt
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda").bfloat16()
    ldb = (K + 7) // 8 * 8
    b = torch.randn(N, K, device="cuda").bfloat16()
    if K % 8:
        pytest.skip("K must be a multiple of 8")
    out = gemm_nt(a, b, out_dtype=torch.float32)
    _close(out, _ref(a, b), rtol=1e-2, atol=1e-2 * (K ** 0.5))
    out_t = gemm_nt(a, b, transpose_out=True)
    _close(out_t, _ref(a, b).t(), rtol=2e-2, atol=2e-2 * (K ** 0.5))


def test_forced_splits_and_block_n_agree():
    from pipe_b200.ops.gemm import gemm_nt
    a = torch.randn(512, 1024, device="cuda").bfloat16()
    b = torch.randn(96, 1024, device="cuda").bfloat16()
    want = _ref(a, b)
    for splits in (1, 2, 4, 16):
        for bn in (0, 32, 96, 128):
            got = gemm_nt(a, b, out_dtype=torch.float32, splits=splits, block_n=bn)
            _close(got, want, rtol=1e-2, atol=0.5)
    # back-to-back split-K launches reuse the self-cleaning workspace
    for _ in range(3):
        got = gemm_nt(a, b, out_dtype=torch.float32, splits=4)
    _close(got, want, rtol=1e-2, atol=0.5)


@pytest.mark.parametrize("M,N,K,splits", [(4096, 75, 4096, 4), (4096, 75, 4096, 8), (1024, 250, 2048, 2),
                                          (384, 33, 1024, 4), (4096, 75, 12288, 4)])
def test_cluster_and_scratch_reductions_agree(M, N, K, splits):
    """Split-K through DSMEM inside a cluster (mode 2) vs through the global workspace (mode 1)."""
    from pipe_b200.ops.gemm import gemm_nt
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(M, device="cuda")
    want = _ref(a, b) + bias[:, None]
    for mode in (1, 2):
        got = gemm_nt(a, b, out_dtype=torch.float32, splits=splits, reduce_mode=mode, bias=bias, bias_along_m=True)
        _close(got, want, rtol=1e-2, atol=1e-2 * (K ** 0.5))
        got_t = gemm_nt(a, b, transpose_out=True, splits=splits, reduce_mode=mode, bias=bias, bias_along_m=True)
        _close(got_t, want.t(), rtol=2e-2, atol=2e-2 * (K ** 0.5))


def test_epilogue_bias_relu_transpose():
    from pipe_b200.ops.gemm import linear_forward
    x = torch.randn(75, 512, device="cuda").bfloat16()
    w = torch.randn(1024, 512, device="cuda").bfloat16()
    bias = torch.randn(1024, device="cuda")
    got = linear_forward(x, w, bias, relu=True)
    want = torch.relu(x.float() @ w.float().t() + bias)
    assert got.shape == (75, 1024)
    _close(got, want, atol=0.5)


def test_epilogue_dropout_is_deterministic_and_scaled():
    from pipe_b200.ops.gemm import linear_forward
    x = torch.randn(64, 256, device="cuda").bfloat16()
    w = torch.randn(512, 256, device="cuda").bfloat16()
    plain = linear_forward(x, w).float()
    a = linear_forward(x, w, drop_p=0.25, seed=11, rng_stream=3).float()
    b = linear_forward(x, w, drop_p=0.25, seed=11, rng_stream=3).float()
    c = linear_forward(x, w, drop_p=0.25, seed=12, rng_stream=3).float()
    assert torch.equal(a, b) and not torch.equal(a, c)
    kept = a != 0
    frac = kept.float().mean().item()
    assert 0.72 < frac < 0.78
    assert torch.allclose(a[kept], (plain / 0.75)[kept], rtol=2e-2, atol=0.3)


def test_epilogue_mask_addend_accumulate_sumsq_alpha():
    from pipe_b200.ops.gemm import gemm_nt
    a = torch.randn(256, 128, device="cuda").bfloat16()
    b = torch.randn(48, 128, device="cuda").bfloat16()
    base = _ref(a, b)
    mask = torch.randn(256, 48, device="cuda").bfloat16()
    add = torch.randn(256, 48, device="cuda").bfloat16()
    got = gemm_nt(a, b, mask_src=mask, mask_scale=1.25, addend=add, out_dtype=torch.float32)
    want = torch.where(mask.float() > 0, base * 1.25, torch.zeros_like(base)) + add.float()
    _close(got, want, atol=0.3)
    acc = torch.ones(256, 48, device="cuda")
    ss = torch.zeros((), device="cuda")
    gemm_nt(a, b, out=acc, accumulate=True, alpha=0.5, sumsq=ss)
    _close(acc, base * 0.5 + 1.0, atol=0.3)
    assert abs(ss.item() - (acc.double() ** 2).sum().item()) / (acc.double() ** 2).sum().item() < 1e-3


@pytest.mark.parametrize("M,N,K,bn", [(2560, 2048, 320, 0), (2500, 1000, 200, 128), (20000, 75, 512, 0)])
def test_persistent_schedule_is_bit_identical(M, N, K, bn):
    """More tiles than SMs: persistent CTAs with the double-buffered TMEM accumulator must give exactly
    what one CTA per tile gives, for every epilogue (same MMA order, same counter-RNG indices)."""
    from pipe_b200.ops.gemm import gemm_nt
    torch.manual_seed(1)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias_m = torch.randn(M, device="cuda")
    bias_n = torch.randn(N, device="cuda")
    mask = torch.randn(N, M, device="cuda").bfloat16()
    variants = [
        dict(out_dtype=torch.float32),
        dict(out_dtype=torch.float32, bias=bias_n, alpha=0.5),
        dict(transpose_out=True, bias=bias_m, bias_along_m=True, relu=True, drop_p=0.3, seed=5, rng_stream=2),
        dict(transpose_out=True, mask_src=mask, mask_scale=1.5, addend=mask),
    ]
    for kw in variants:
        one = gemm_nt(a, b, block_n=bn, persistent=2, splits=1, **kw)
        per = gemm_nt(a, b, block_n=bn, persistent=1, splits=1, **kw)
        assert torch.equal(one, per), kw.keys()
    want = _ref(a, b)
    _close(gemm_nt(a, b, block_n=bn, persistent=1, splits=1, out_dtype=torch.float32), want, rtol=1e-2, atol=1e-2 * K ** 0.5)
    # accumulate + sumsq across several tiles per CTA
    acc1 = torch.ones(M, N, device="cuda"); acc2 = torch.ones(M, N, device="cuda")
    s1 = torch.zeros((), device="cuda"); s2 = torch.zeros((), device="cuda")
    gemm_nt(a, b, out=acc1, accumulate=True, sumsq=s1, block_n=bn, persistent=2, splits=1)
    gemm_nt(a, b, out=acc2, accumulate=True, sumsq=s2, block_n=bn, persistent=1, splits=1)
    assert torch.equal(acc1, acc2)
    assert abs(s1.item() - s2.item()) <= 1e-4 * abs(s1.item())


@pytest.mark.parametrize("M,N,K", [(4096, 75, 4096), (512, 48, 1000), (4096, 75, 28782), (2048, 1024, 6144), (300, 20, 192)])
def test_mn_major_a_operand_matches_transposed_copy(M, N, K):
    """``a_mn=True`` reads A from its transpose ``[K, M]`` (M contiguous) as an MN-major UMMA operand: bit-identical to the
    same GEMM on a materialised ``[M, K]`` copy (same products, same accumulation order), on every scheduling path."""
    from pipe_b200.ops.gemm import gemm_nt
    torch.manual_seed(2)
    Mp = (M + 7) // 8 * 8
    at = torch.randn(K, Mp, device="cuda").bfloat16()[:, :M]          # [K, M], row stride multiple of 8
    Kp = (K + 7) // 8 * 8
    a = torch.zeros(M, Kp, device="cuda", dtype=torch.bfloat16)[:, :K]
    a.copy_(at.t())
    b = torch.randn(N, Kp, device="cuda").bfloat16()[:, :K]
    bias = torch.randn(M, device="cuda")
    for kw in (dict(transpose_out=True), dict(transpose_out=True, bias=bias, bias_along_m=True, relu=True),
               dict(out_dtype=torch.float32), dict(transpose_out=True, splits=1, persistent=1)):
        want = gemm_nt(a, b, **kw)
        got = gemm_nt(at, b, a_mn=True, **kw)
        assert torch.equal(got, want), kw
    ref = (a.float() @ b.float().t())
    got = gemm_nt(at, b, a_mn=True, out_dtype=torch.float32)
    assert ((got - ref).abs().max() / ref.abs().max()).item() < 1e-2
