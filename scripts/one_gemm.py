"""A tiny, deterministic GEMM workload for ncu: the out-proj / FFN shape of BASELINE config 3
(W[4096,4096] x 75 tokens, swap-AB, cluster split-K) launched a few times, plus the wgrad shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops.gemm import gemm_nt
dev = "cuda"
torch.manual_seed(0)
which = sys.argv[1] if len(sys.argv) > 1 else "skinny"
if which == "skinny":
    w = [torch.randn(4096, 4096, device=dev).bfloat16() for _ in range(6)]     # rotate: L2-cold weights
    x = torch.randn(75, 4096, device=dev).bfloat16()
    bias = torch.randn(4096, device=dev)
    for i in range(12):
        y = gemm_nt(w[i % 6], x, transpose_out=True, bias=bias, bias_along_m=True, relu=True, drop_p=0.2, seed=3, rng_stream=2)
else:
    dyt = torch.randn(4096, 512, device=dev).bfloat16()
    xt = torch.randn(4096, 512, device=dev).bfloat16()
    g = torch.empty(4096, 4096, device=dev)
    for i in range(6):
        gemm_nt(dyt[:, :500], xt[:, :500], out=g)
torch.cuda.synchronize()
print("ok")
