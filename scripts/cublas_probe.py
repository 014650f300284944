"""Which kernels does cuBLAS pick for the skinny weight-streaming shapes (name, duration)?  The names encode tile,
stage count, cluster shape and 1-/2-CTA MMA — the reference point for csrc/gemm_sm100.cu."""
import torch
from torch.profiler import profile, ProfilerActivity
dev="cuda"
E,T=4096,75
ws=[torch.randn(E,E,device=dev).bfloat16() for _ in range(10)]
x=torch.randn(T,E,device=dev).bfloat16()
for i in range(5): torch.matmul(x, ws[i].t())
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(10): torch.matmul(x, ws[i].t())
    torch.cuda.synchronize()
for e in prof.events():
    if e.device_time>0: print(f"{e.device_time:7.1f} us  {e.name[:150]}")
w3=[torch.randn(3*E,E,device=dev).bfloat16() for _ in range(4)]
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(4): torch.matmul(x, w3[i].t())
    torch.cuda.synchronize()
for e in prof.events():
    if e.device_time>0: print(f"{e.device_time:7.1f} us  {e.name[:150]}")
