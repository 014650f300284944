"""A tiny attention workload for ncu: the config-3 micro-batch (3 sequences x 25 tokens, 16 heads of 256)
forward + backward, a few times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops import functional as F2
dev = "cuda"
torch.manual_seed(0)
NB, S, H, D = 3, 25, 16, 256
E = H * D
qkv = (torch.randn(S * NB, 3 * E, device=dev) * 0.5).bfloat16()
dout = torch.randn(S * NB, E, device=dev).bfloat16()
for _ in range(6):
    F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=0.2, seed=1, rng_stream=0)
    F2.attention_bwd(qkv, dout, NB, S, H, causal=False, drop_p=0.2, seed=1, rng_stream=0)
torch.cuda.synchronize()
print("ok")
