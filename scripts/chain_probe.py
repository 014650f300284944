"""Chain kernel against the same ops as separate kernels: bit comparison and timing (one GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.ops.chain import Chain
from pipe_b200.ops.gemm import gemm_nt
from pipe_b200.ops import functional as F

dev = torch.device("cuda:0")
torch.manual_seed(0)
T, E, FF = 75, 4096, 4096
mode = sys.argv[1] if len(sys.argv) > 1 else "all"

def w(m, k): return (torch.randn(m, k, device=dev) / k ** 0.5).bfloat16()

def time_us(fn, n=50, warm=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n

x = torch.randn(T, E, device=dev).bfloat16()
L = 6                                      # rotate weights through > L2
Ws = [dict(qkv=w(3 * E, E), bq=torch.randn(3 * E, device=dev), o=w(E, E), f1=w(FF, E), f2=w(E, FF), bo=torch.randn(E, device=dev), b1=torch.randn(FF, device=dev),
           b2=torch.randn(E, device=dev)) for _ in range(L)]
g1 = torch.rand(E, device=dev) + 0.5; be1 = torch.randn(E, device=dev)

def bufs():
    return dict(o=torch.empty(T, E, device=dev, dtype=torch.bfloat16), x1=torch.empty(T, E, device=dev, dtype=torch.bfloat16),
                h=torch.empty(T, FF, device=dev, dtype=torch.bfloat16), f=torch.empty(T, E, device=dev, dtype=torch.bfloat16),
                y=torch.empty(T, E, device=dev, dtype=torch.bfloat16), qkv=torch.empty(T, 3 * E, device=dev, dtype=torch.bfloat16),
                st=torch.empty(4, T, device=dev, dtype=torch.float32))

def separate(W, B, p=0.2, seed=7):
    gemm_nt(W["o"], x, out=B["o"], transpose_out=True, bias=W["bo"], bias_along_m=True, drop_p=p, seed=seed, rng_stream=1, a_static=True)
    ln_sep(x, B["o"], B["x1"], B["st"][0], B["st"][1])
    gemm_nt(W["f1"], B["x1"], out=B["h"], transpose_out=True, bias=W["b1"], bias_along_m=True, relu=True, drop_p=p, seed=seed, rng_stream=2, a_static=True)
    gemm_nt(W["f2"], B["h"], out=B["f"], transpose_out=True, bias=W["b2"], bias_along_m=True, drop_p=p, seed=seed, rng_stream=3, a_static=True)
    ln_sep(B["x1"], B["f"], B["y"], B["st"][2], B["st"][3])

import ctypes
from pipe_b200.ops import _lib
def ln_sep(xx, hh, yy, mean, rstd):
    rc = _lib.load().pb2_add_layernorm_fwd(ctypes.c_void_p(xx.data_ptr()), ctypes.c_void_p(hh.data_ptr()), ctypes.c_void_p(g1.data_ptr()),
                                           ctypes.c_void_p(be1.data_ptr()), ctypes.c_void_p(yy.data_ptr()), ctypes.c_void_p(mean.data_ptr()),
                                           ctypes.c_void_p(rstd.data_ptr()), ctypes.c_int(T), ctypes.c_int(E), ctypes.c_float(1e-5),
                                           ctypes.c_void_p(_lib.raw_stream(dev)))
    _lib.check(rc, "ln")

def chained(W, B, p=0.2, seed=7, upto=5, build_only=False):
    c = Chain(dev)
    c.gemm(W["o"], x, B["o"], bias=W["bo"], drop_p=p, seed=seed, rng_stream=1)
    if upto >= 2: c.layernorm(x, B["o"], g1, be1, B["x1"], B["st"][0], B["st"][1])
    if upto >= 3: c.gemm(W["f1"], B["x1"], B["h"], bias=W["b1"], relu=True, drop_p=p, seed=seed, rng_stream=2)
    if upto >= 4: c.gemm(W["f2"], B["h"], B["f"], bias=W["b2"], drop_p=p, seed=seed, rng_stream=3)
    if upto >= 5: c.layernorm(B["x1"], B["f"], g1, be1, B["y"], B["st"][2], B["st"][3])
    if upto >= 6: c.gemm(W["qkv"], B["y"], B["qkv"], bias=W["bq"])
    assert c.supported()
    if build_only: return c
    c.run()

A, Bc = bufs(), bufs()
separate(Ws[0], A)
torch.cuda.synchronize()
for upto in (1, 2, 3, 4, 5):
    for k in Bc: Bc[k].zero_()
    chained(Ws[0], Bc, upto=upto)
    torch.cuda.synchronize()
    names = ["o", "x1", "h", "f", "y"][:upto]
    print(f"chain of {upto} ops:", {n: bool(torch.equal(A[n], Bc[n])) for n in names},
          "stats", bool(torch.equal(A["st"][: 2 * (upto >= 2) + 2 * (upto >= 5)], Bc["st"][: 2 * (upto >= 2) + 2 * (upto >= 5)])), flush=True)
# repeated launches (barrier counters re-armed), rotating weights
for i in range(20):
    chained(Ws[i % L], Bc)
torch.cuda.synchronize()
separate(Ws[19 % L], A); torch.cuda.synchronize()
print("after 20 launches equal:", bool(torch.equal(A["y"], Bc["y"])), flush=True)
it = [0]
def run_sep():
    it[0] += 1; separate(Ws[it[0] % L], A)
UPTO = int(os.environ.get("UPTO", 6))
pre = [chained(Ws[i], Bc, build_only=True, upto=UPTO) for i in range(L)]
def run_chain():
    it[0] += 1; pre[it[0] % L].run()
def run_qkv():
    it[0] += 1; gemm_nt(Ws[it[0] % L]["qkv"], x, out=A["qkv"], transpose_out=True, bias=Ws[it[0] % L]["bq"], bias_along_m=True, a_static=True)
print(f"standalone QKV GEMM: {time_us(run_qkv):.1f} us")
print(f"separate kernels: {time_us(run_sep):.1f} us   chain of {UPTO}: {time_us(run_chain):.1f} us   (out-proj, LN, FFN1, FFN2, LN; 100 MB of weights = {100.7e6 / 6.447e12 * 1e6:.1f} us of HBM)")

# ---- in-kernel timeline of one launch (after a cold-L2 flush) ----
lib = _lib.load()
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
ts = torch.zeros(148 * 9 * 8, dtype=torch.int64, device=dev)
pre[1].run(); pre[2].run()
flush.zero_()
lib.pb2_debug_chain_timestamps(ctypes.c_void_p(ts.data_ptr()))
pre[3].run()
lib.pb2_debug_chain_timestamps(ctypes.c_void_p(0))
torch.cuda.synchronize()
t = ts.view(148, 9, 8).cpu()
live = t[:, 8, 0] > 0
t = t[live]
t0 = t[:, 8, 0].min()
print(f"CTAs {t.size(0)}; kernel span {(t[:, 8, 1].max() - t0).item() / 1e3:.1f} us; start skew {(t[:, 8, 0].max() - t0).item() / 1e3:.2f} us")
names = ["out-proj", "LN1", "FFN1", "FFN2", "LN2", "QKV"][:UPTO]
for op, nm in enumerate(names):
    for slot, what in ((0, "input ready (grid barrier passed)"), (3, "first MMA operands landed"), (1, "first accumulator ready"), (4, "partials sent"), (5, "operands prefetched"), (6, "partials received"), (7, "own chunks stored"), (2, "op done in this CTA")):
        col = t[:, op, slot]
        ok = col > 0
        if ok.any():
            c = (col[ok] - t0).float() / 1e3
            print(f"  {nm:9s} {what:36s} min {c.min():6.2f}  median {c.median():6.2f}  max {c.max():6.2f} us  ({int(ok.sum())} CTAs)")
