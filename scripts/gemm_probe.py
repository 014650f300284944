"""Probe the GEMM kernel's per-CTA streaming rate and fixed costs (L2-cold operands)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops.gemm import gemm_nt
dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

def t(M, N, K, **kw):
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    out = torch.empty(N, M, dtype=torch.bfloat16, device=dev)
    from torch.profiler import ProfilerActivity, profile
    gemm_nt(a, b, out=out, transpose_out=True, **kw)
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(5):
            flush.zero_()
            gemm_nt(a, b, out=out, transpose_out=True, **kw)
        torch.cuda.synchronize()
    ts = sorted(e.device_time for e in prof.events() if "gemm_bf16" in e.name)
    us = ts[len(ts) // 2]
    print(f"M={M:6d} N={N:4d} K={K:5d} {str(kw):34s} {us:8.1f} us  {(M*K*2)/us/1e3:7.0f} GB/s  ctas={((M+127)//128)*max(1,kw.get('splits',1))}", flush=True)

t(128, 80, 512, splits=1)
t(128, 80, 4096, splits=1)
t(128, 80, 16384, splits=1)
t(128, 16, 16384, splits=1)
t(128, 80, 16384, splits=1, stages=2)
t(128, 80, 16384, splits=1, stages=4)
t(18944, 80, 4096, splits=1)
t(18944, 16, 4096, splits=1)
t(18944 * 2, 80, 4096, splits=1)
t(18944, 80, 4096, splits=1, stages=3)
t(4096, 80, 4096, splits=4)
t(4096, 80, 4096, splits=4, stages=3)
t(4096, 16, 4096, splits=4)
x = torch.empty(1 << 28, dtype=torch.uint8, device=dev); y = torch.empty_like(x)
for i in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y.copy_(x); e1.record(); torch.cuda.synchronize()
print("copy 256MB:", e0.elapsed_time(e1) * 1e3, "us ->", 2 * (1 << 28) / e0.elapsed_time(e1) / 1e6, "GB/s")
