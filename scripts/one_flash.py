"""One flash-attention forward + backward (for ncu captures): python scripts/one_flash.py [S NB H D causal]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pipe_b200.ops import functional as F2

S, NB, H, D, causal = (int(x) for x in (sys.argv[1:6] if len(sys.argv) >= 6 else (2048, 1, 16, 64, 0)))
E = H * D
torch.manual_seed(0)
qkv = (torch.randn(S * NB, 3 * E, device="cuda") * 0.5).bfloat16()
dout = (torch.randn(S * NB, E, device="cuda") * 0.5).bfloat16()
for _ in range(3):
    out, lse = F2.flash_attention_fwd(qkv, NB, S, H, causal=bool(causal), drop_p=0.0, seed=0, rng_stream=0)
    dqkv = F2.flash_attention_bwd(qkv, out, dout, lse, NB, S, H, causal=bool(causal), drop_p=0.0, seed=0, rng_stream=0)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(dqkv.float().abs().mean()))
