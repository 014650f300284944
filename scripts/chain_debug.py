import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops.chain import Chain
dev = torch.device("cuda:0")
def _w(m, k): return (torch.randn(m, k, device=dev) / k ** 0.5).bfloat16()
def _ln_ref(x, h, g, b, eps=1e-5):
    z = x.float() + h.float()
    return torch.nn.functional.layer_norm(z, (z.size(-1),), g, b, eps)
def err(a, b):
    a = a.float(); b = b.float()
    bad = ~torch.isfinite(a)
    return f"nan rows {sorted(set(bad.nonzero()[:, 0].tolist()))[:6]}" if bad.any() else f"{((a - b).norm() / b.norm()).item():.2e}"
for (T, E, FF) in [(128, 512, 512), (75, 512, 512), (128, 1024, 2048), (20, 256, 1024), (64, 512, 512), (112, 512, 512), (96, 512, 512)]:
    torch.manual_seed(0)
    x = torch.randn(T, E, device=dev).bfloat16()
    wo, w1, w2 = _w(E, E), _w(FF, E), _w(E, FF)
    bo, b1, b2 = (torch.randn(n, device=dev) for n in (E, FF, E))
    g1, be1 = torch.rand(E, device=dev) + 0.5, torch.randn(E, device=dev)
    o, x1, f, y = (torch.zeros(T, E, device=dev, dtype=torch.bfloat16) for _ in range(4))
    h = torch.zeros(T, FF, device=dev, dtype=torch.bfloat16)
    st = torch.zeros(4, T, device=dev)
    c = Chain(dev)
    c.gemm(wo, x, o, bias=bo).layernorm(x, o, g1, be1, x1, st[0], st[1])
    c.gemm(w1, x1, h, bias=b1, relu=True).gemm(w2, h, f, bias=b2).layernorm(x1, f, g1, be1, y, st[2], st[3])
    c.run(); torch.cuda.synchronize()
    o_ref = (x.float() @ wo.float().t() + bo).bfloat16()
    x1_ref = _ln_ref(x, o, g1, be1).bfloat16()
    h_ref = torch.relu(x1.float() @ w1.float().t() + b1).bfloat16()
    f_ref = (h.float() @ w2.float().t() + b2).bfloat16()
    y_ref = _ln_ref(x1, f, g1, be1)
    print((T, E, FF), "o", err(o, o_ref), "x1", err(x1, x1_ref), "h", err(h, h_ref), "f", err(f, f_ref), "y", err(y, y_ref), flush=True)
