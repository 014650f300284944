"""Attention core timings on one B200: the tcgen05 flash kernels (csrc/attention_flash.cu), the single-tile mma.sync
kernels (csrc/attention_tc.cu, S <= 128) and ATen's scaled_dot_product_attention on the same shapes, forward and
backward, CUDA-graph replay so launch overhead is out (same protocol as scripts/bench_gemm.py).

    python scripts/bench_attention.py > profiles/attention_bench.txt
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from pipe_b200.ops import functional as F2


def timed(fn, iters=20, graph=True):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if not graph:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(iters):
            fn()
        t1.record()
        torch.cuda.synchronize()
        return t0.elapsed_time(t1) * 1e3 / iters
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    g.replay()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) * 1e3 / iters


def main():
    dev = "cuda"
    shapes = [(128, 8, 32, 64, "reference as shipped: bptt 128, 8 rows per micro-batch, 32 heads"),
              (128, 8, 16, 128, ""), (512, 4, 16, 64, ""), (512, 4, 16, 128, ""), (2048, 1, 16, 64, ""),
              (2048, 1, 16, 128, ""), (4096, 1, 8, 128, "")]
    print("S NB H D causal | flash fwd us (TFLOP/s) | flash bwd us (TFLOP/s) | tile-kernel fwd / bwd us | ATen SDPA fwd / bwd us")
    for S, NB, H, D, note in shapes:
        for causal in (False, True):
            E = H * D
            qkv = (torch.randn(S * NB, 3 * E, device=dev) * 0.5).bfloat16()
            dout = (torch.randn(S * NB, E, device=dev) * 0.5).bfloat16()
            flops = 4.0 * NB * H * S * S * D * (0.5 if causal else 1.0)
            out, lse = F2.flash_attention_fwd(qkv, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0)
            t_f = timed(lambda: F2.flash_attention_fwd(qkv, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0))
            t_b = timed(lambda: F2.flash_attention_bwd(qkv, out, dout, lse, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0))
            tile = "- / -"
            if S <= 128 and not (S == 128 and D > 64):
                tf = timed(lambda: F2.attention_fwd(qkv, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0))
                tb = timed(lambda: F2.attention_bwd(qkv, dout, NB, S, H, causal=causal, drop_p=0.0, seed=0, rng_stream=0))
                tile = f"{tf:.1f} / {tb:.1f}"
            x = qkv.view(S, NB, 3, H, D)
            q, k, v = (x[:, :, i].permute(1, 2, 0, 3).contiguous().requires_grad_() for i in range(3))
            ts = timed(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=causal))
            o = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
            go = torch.randn_like(o)
            tsb = timed(lambda: torch.autograd.grad(o, (q, k, v), go, retain_graph=True), iters=50, graph=False)   # the engine cannot be captured
            print(f"{S} {NB} {H} {D} {int(causal)} | {t_f:.1f} ({flops / t_f / 1e6:.0f}) | {t_b:.1f} ({2.5 * flops / t_b / 1e6:.0f}) | "
                  f"{tile} | {ts:.1f} / {tsb:.1f}  {note}")


if __name__ == "__main__":
    main()
