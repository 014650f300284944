"""Every non-GEMM kernel of the fused path against the ATen sequence it replaces, at the shapes of
BASELINE config 3 (micro-batch = 3 sequences x 25 tokens, E = 4096, 16 heads, vocab 28782).
SURVEY.md §8 W6-W10: "baseline to beat on the same box".

Timing: each variant is captured into a CUDA graph (12 back-to-back calls) and the graph is replayed
between two events, so neither side pays Python or launch latency of the host.
    python scripts/bench_ops.py [--out gpurun_out/ops_vs_aten.txt]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from pipe_b200.ops import functional as F2

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/ops_vs_aten.txt")
args = ap.parse_args()
dev = torch.device("cuda")
torch.manual_seed(0)
lines = []


def graph_time_us(fn, n_iter=12):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(n_iter):
            fn()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n_iter * 1e3)
    return best


def report(name, ours, aten, note=""):
    lines.append(f"{name:44s} ours {ours:8.1f} us   ATen {aten:8.1f} us   x{aten / ours:5.2f}  {note}")
    print(lines[-1], flush=True)


NB, S, H, E, V = 3, 25, 16, 4096, 28782
T, D, p = NB * S, E // H, 0.2

# ---- dropout + residual + LayerNorm ------------------------------------------------------------
x = torch.randn(T, E, device=dev).bfloat16()
h = torch.randn(T, E, device=dev).bfloat16()
gamma = torch.rand(E, device=dev) + 0.5
beta = torch.randn(E, device=dev)
ln = torch.nn.LayerNorm(E, device=dev, dtype=torch.bfloat16)
report("add + LayerNorm fwd (75 x 4096)",
       graph_time_us(lambda: F2.add_layernorm_fwd(x, h, gamma, beta, 1e-5)),
       graph_time_us(lambda: ln(x + F.dropout(h, p, True))),
       "ATen side includes the dropout our GEMM epilogue already applied")
y, mean, rstd = F2.add_layernorm_fwd(x, h, gamma, beta, 1e-5)
dy = torch.randn(T, E, device=dev).bfloat16()
dg, db = torch.zeros(E, device=dev), torch.zeros(E, device=dev)
xa = x.clone().requires_grad_()
ha = h.clone().requires_grad_()


def aten_ln_bwd():
    out = ln(xa + F.dropout(ha, p, True))
    torch.autograd.grad(out, [xa, ha, ln.weight, ln.bias], dy)


report("add + LayerNorm fwd+bwd",
       graph_time_us(lambda: (F2.add_layernorm_fwd(x, h, gamma, beta, 1e-5),
                              F2.add_layernorm_bwd(dy, x, h, gamma, mean, rstd, dg, db, want_dh=True, drop_p=p, seed=1))),
       graph_time_us(aten_ln_bwd))

# ---- attention core ---------------------------------------------------------------------------------
qkv = (torch.randn(T, 3 * E, device=dev) * 0.5).bfloat16()
dout = torch.randn(T, E, device=dev).bfloat16()


def aten_attn_fwd():
    t = qkv.view(S, NB, 3, H, D)
    q, k, v = (t[:, :, i].permute(1, 2, 0, 3) for i in range(3))
    o = F.scaled_dot_product_attention(q, k, v, dropout_p=p)
    return o.permute(2, 0, 1, 3).reshape(T, E)


qkv_a = qkv.clone().requires_grad_()


def aten_attn_fwd_bwd():
    t = qkv_a.view(S, NB, 3, H, D)
    q, k, v = (t[:, :, i].permute(1, 2, 0, 3) for i in range(3))
    o = F.scaled_dot_product_attention(q, k, v, dropout_p=p).permute(2, 0, 1, 3).reshape(T, E)
    torch.autograd.grad(o, qkv_a, dout)


report("attention fwd (3 x 16 heads, S 25, D 256)",
       graph_time_us(lambda: F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)),
       graph_time_us(aten_attn_fwd), "ATen = SDPA on permuted views of the packed projection")
report("attention fwd+bwd",
       graph_time_us(lambda: (F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0),
                              F2.attention_bwd(qkv, dout, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0))),
       graph_time_us(aten_attn_fwd_bwd))

# ---- embedding + positions + dropout ---------------------------------------------------------------
tok = torch.randint(0, V, (NB, S), device=dev)
emb = torch.randn(V, E, device=dev)
pe = torch.randn(5000, 1, E, device=dev)
pe2 = pe.view(5000, E).contiguous()
report("embedding * sqrt(E) + PE + dropout fwd",
       graph_time_us(lambda: F2.embed_fwd(tok, emb, pe2, E ** 0.5, p, 1, 0)),
       graph_time_us(lambda: F.dropout(F.embedding(tok, emb).transpose(0, 1) * E ** 0.5 + pe[:S], p, True)),
       "ATen in fp32 (the tutorial's dtype)")

# ---- cross-entropy ----------------------------------------------------------------------------------
ld = (V + 63) // 64 * 64
logits = torch.zeros(T, ld, device=dev, dtype=torch.bfloat16)
logits[:, :V] = torch.randn(T, V, device=dev)
tgt = torch.randint(0, V, (T,), device=dev)
la = logits[:, :V].clone().requires_grad_()


def aten_ce():
    loss = F.cross_entropy(la.float(), tgt)
    torch.autograd.grad(loss, la)


work = logits.clone()
report("cross-entropy fwd+bwd (75 x 28782, bf16 logits)",
       graph_time_us(lambda: (work.copy_(logits), F2.cross_entropy(work, tgt, V, write_grad=True))),
       graph_time_us(aten_ce), "ours includes a 4.3 MB refresh copy of the logits it overwrites")

# ---- clip + Adam ------------------------------------------------------------------------------------
from pipe_b200.models.fused import FusedEncoderLayer  # noqa: E402
from pipe_b200.ops.optim import FusedClipAdam  # noqa: E402

layers = torch.nn.Sequential(*[FusedEncoderLayer(E, H, E, 0.0).to(dev) for _ in range(2)])
n_params = sum(q.numel() for q in layers.parameters())
opt = FusedClipAdam(layers, lr=1e-4, max_norm=0.5)
ref_params = [q.detach().clone().requires_grad_() for q in layers.parameters()]
ropt = torch.optim.Adam(ref_params, lr=1e-4, fused=True)
opt.zero_grad()
for q, r in zip(layers.parameters(), ref_params):
    q.grad.copy_(torch.randn_like(q) * 1e-3)
    r.grad = q.grad.clone()


def ours_adam():
    opt.step()          # squared norm of every gradient + clip factor + Adam + both bf16 shadows


def aten_adam():
    torch.nn.utils.clip_grad_norm_(ref_params, 0.5, foreach=True)
    ropt.step()


def ev_time_us(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3)
    return best


a_us, b_us = ev_time_us(ours_adam), ev_time_us(aten_adam)
report(f"clip_grad_norm + Adam ({n_params / 1e6:.0f} M params)", a_us, b_us,
       f"ours also writes both bf16 weight shadows; {n_params * 32 / a_us / 1e6:.2f} TB/s at 32 B/param")

os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "w").write("\n".join(lines) + "\n")
