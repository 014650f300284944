"""Phase timestamps inside the GEMM kernel (debug hook): where does a CTA's time go?"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops import _lib
from pipe_b200.ops.gemm import gemm_nt
lib = _lib.load()
dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
names = ["start", "setup done", "first stage landed", "last MMA issued", "accum ready", "epilogue done", "end", "cluster barrier passed", "partials sent", "epilogue operands ready", "partials received"]

def run(M, N, K, wgrad=False, **kw):
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    if wgrad:
        out = torch.empty(M, N, dtype=torch.float32, device=dev)
        bn = kw.get("block_n", 256)
        ctas = ((M + 127) // 128) * ((N + bn - 1) // bn)
        kw = dict(kw, transpose_out=False)
    else:
        out = torch.empty(N, M, dtype=torch.bfloat16, device=dev)
        ctas = max(((M + 127) // 128) * kw.get("splits", 1), 256)
        kw = dict(kw, transpose_out=True)
    ts = torch.zeros(ctas * 16, dtype=torch.int64, device=dev)
    gemm_nt(a, b, out=out, **kw)
    flush.zero_()
    lib.pb2_debug_gemm_timestamps(ctypes.c_void_p(ts.data_ptr()))
    gemm_nt(a, b, out=out, **kw)
    lib.pb2_debug_gemm_timestamps(ctypes.c_void_p(0))
    torch.cuda.synchronize()
    t = ts.view(ctas, 16).cpu()
    t = t[t[:, 0] > 0]                 # CTAs that actually ran (the buffer is sized generously)
    ctas = t.size(0)
    t0 = t[:, 0].min()
    kw = {k: (v if not torch.is_tensor(v) else "tensor") for k, v in kw.items()}
    print(f"--- M={M} N={N} K={K} {kw} ctas={ctas}: kernel span {(t[:, 6].max() - t0).item() / 1e3:.1f} us")
    for k, nm in enumerate(names):
        if (t[:, k] <= 0).all():
            continue
        col = (t[:, k] - t0).float() / 1e3
        print(f"   {nm:20s} min {col.min():7.2f}  median {col.median():7.2f}  max {col.max():7.2f} us")

if len(sys.argv) > 1 and sys.argv[1] == "big":
    run(4096, 4096, 512, wgrad=True)
    run(4096, 4096, 512, wgrad=True, stages=4)
    run(6144, 1024, 2048)
    run(6144, 1024, 2048, stages=4)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "auto":
    run(4096, 75, 4096)
    run(4096, 75, 4096, bias=torch.randn(4096, device=dev), bias_along_m=True, relu=True, drop_p=0.2, seed=5)
    run(4096, 75, 12288)
    run(12288, 75, 4096)
    sys.exit(0)
run(4096, 80, 4096, splits=4, reduce_mode=1)
run(4096, 80, 4096, splits=4, reduce_mode=2)
run(4096, 80, 4096, splits=2, reduce_mode=2)
run(4096, 80, 4096, splits=8, reduce_mode=2)
run(4096, 80, 4096, splits=4, reduce_mode=2, bias=torch.randn(4096, device=dev), bias_along_m=True, relu=True, drop_p=0.2, seed=5)
run(12288, 80, 4096, splits=1)
run(12288, 80, 4096, splits=2, reduce_mode=2)
