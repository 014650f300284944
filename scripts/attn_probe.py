"""Attention core timings at the shapes that matter (CUPTI kernel durations; run with PB2_NO_PDL=1 for per-kernel
times that do not include waiting for the predecessor) plus the in-kernel phase timestamps of CTA 0."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from pipe_b200.ops import functional as F2
dev = "cuda"
def t(NB, S, H, D, p):
    E = H * D
    qkv = (torch.randn(S * NB, 3 * E, device=dev) * 0.5).bfloat16()
    dout = torch.randn(S * NB, E, device=dev).bfloat16()
    F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)
            F2.attention_bwd(qkv, dout, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)
        torch.cuda.synchronize()
    f = sorted(e.device_time for e in prof.events() if "attn" in e.name and "fwd" in e.name)
    b = sorted(e.device_time for e in prof.events() if "attn" in e.name and "bwd" in e.name)
    import ctypes
    from pipe_b200.ops import _lib
    F2.attention_fwd(qkv, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)
    torch.cuda.synchronize()
    ts = (ctypes.c_ulonglong * 16)()
    _lib.load().pb2_debug_attn_timestamps(ts)
    print("   fwd phases (us): gather %.2f  QK^T %.2f  softmax %.2f  dropout %.2f  PV+emit %.2f" % tuple((ts[i + 1] - ts[i]) / 1e3 for i in range(5)))
    F2.attention_bwd(qkv, dout, NB, S, H, causal=False, drop_p=p, seed=1, rng_stream=0)
    torch.cuda.synchronize()
    _lib.load().pb2_debug_attn_timestamps(ts)
    print("   bwd phases (us): gather %.2f  QK^T+dOV^T %.2f  softmax %.2f  dS %.2f  3 products+emit %.2f" % tuple((ts[i + 1] - ts[i]) / 1e3 for i in range(8, 13)))
    print(f"NB={NB} S={S} H={H} D={D} p={p}: fwd {f[len(f)//2]:.1f} us  bwd {b[len(b)//2]:.1f} us  ctas={NB*H}", flush=True)
t(3, 25, 16, 256, 0.2)
t(3, 25, 16, 256, 0.0)
t(1, 25, 1, 256, 0.0)
t(9, 25, 16, 256, 0.0)
t(3, 25, 16, 64, 0.0)
t(3, 25, 64, 64, 0.0)
t(8, 128, 32, 64, 0.2)
