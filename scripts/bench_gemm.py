"""GEMM micro-benchmark: every shape class of BASELINE config 3 (E = H = 4096, 75-token micro-batches,
500-token mini-batch, V = 28782), CUDA-event timed, weights rotated through > 126 MB so nothing is
L2-resident between iterations.  Reports achieved GB/s against the measured HBM copy bandwidth and the
cuBLAS (torch.matmul) time for the same math.

    python scripts/bench_gemm.py [--sweep] [--out gpurun_out/gemm_bench.txt]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pipe_b200.ops.gemm import gemm_nt

ap = argparse.ArgumentParser()
ap.add_argument("--sweep", action="store_true")
ap.add_argument("--out", default="gpurun_out/gemm_bench.txt")
args = ap.parse_args()

peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = peaks.get("hbm_gbs", 6650.0)
dev = "cuda"
lines = []


def time_ms(fn, n_iter=12, n_warm=3):
    """Per-call device time with no host in the loop: the calls are captured once into a CUDA
    graph (on a side stream) and the graph is replayed between two events."""
    for i in range(n_warm):
        fn(i)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(n_warm):
            fn(i)                      # per-stream workspaces get allocated outside capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        for i in range(n_iter):
            fn(i)
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n_iter)
    return best


def case(name, M, N, K, out_f32=False, transpose=True, **kw):
    # rotate the big operand through enough copies to defeat the 126 MB L2
    a_bytes = M * K * 2
    copies = max(2, int(300e6 // a_bytes) + 1)
    Kp = (K + 7) // 8 * 8                      # row stride must be a multiple of 8 elements
    As = [torch.randn(M, Kp, device=dev).bfloat16()[:, :K] for _ in range(copies)]
    b = torch.randn(N, Kp, device=dev).bfloat16()[:, :K]
    out = torch.empty((N, M) if transpose else (M, N), dtype=torch.float32 if out_f32 else torch.bfloat16, device=dev)
    ours = time_ms(lambda i: gemm_nt(As[i % copies], b, out=out, transpose_out=transpose, **kw))
    ref = time_ms(lambda i: torch.matmul(b, As[i % copies].t()) if transpose else torch.matmul(As[i % copies], b.t()))
    nbytes = a_bytes + N * K * 2 + out.numel() * out.element_size()
    gbs = nbytes / ours / 1e6
    flops = 2.0 * M * N * K / ours / 1e9
    lines.append(f"{name:34s} M={M:6d} N={N:5d} K={K:5d} {str(kw):28s} ours {ours * 1e3:8.1f} us  {gbs:7.0f} GB/s "
                 f"({gbs / HBM * 100:5.1f}% of measured HBM)  {flops:7.1f} TFLOP/s   cuBLAS {ref * 1e3:8.1f} us")
    print(lines[-1], flush=True)


T, E, V = 75, 4096, 28782
if args.sweep and os.environ.get("SWEEP") == "epi":
    mask = torch.randn(T, E, device=dev).bfloat16()
    bias = torch.randn(E, device=dev)
    case("E x E plain", E, T, E)
    case("E x E bias", E, T, E, bias=bias, bias_along_m=True)
    case("E x E bias+relu+drop", E, T, E, bias=bias, bias_along_m=True, relu=True, drop_p=0.2, seed=3)
    case("E x E mask", E, T, E, mask_src=mask, mask_scale=1.25)
    case("E x E addend", E, T, E, addend=mask)
    case("E x E plain splits=2", E, T, E, splits=2)
    case("E x E plain splits=8", E, T, E, splits=8)
    case("3E x E plain (QKV)", 3 * E, T, E)
    case("3E x E splits=2", 3 * E, T, E, splits=2)
    case("E x 3E (QKV dgrad)", E, T, 3 * E)
    case("E x 3E splits=8", E, T, 3 * E, splits=8)
elif args.sweep and os.environ.get("SWEEP") == "persist":
    # persistent (double-buffered accumulator) against one-CTA-per-tile, with a numerics check
    for (nm, M, N, K, f32, tr) in (("wgrad ExE", E, E, 512, True, False), ("wgrad 3ExE", 3 * E, E, 512, True, False),
                                   ("wgrad vocab", V, E, 512, True, False), ("shipped 1024 tok", 3 * 2048, 1024, 2048, False, True),
                                   ("fwd vocab", V, T, E, False, True), ("dgrad vocab", E, T, V, False, True)):
        a = torch.randn(M, (K + 7) // 8 * 8, device=dev).bfloat16()[:, :K]
        b = torch.randn(N, (K + 7) // 8 * 8, device=dev).bfloat16()[:, :K]
        want = a.float() @ b.float().t()
        got = gemm_nt(a, b, transpose_out=tr, out_dtype=torch.float32 if f32 else torch.bfloat16, persistent=1).float()
        if tr:
            got = got.t()
        err = ((got - want).abs().max() / want.abs().max()).item()
        print(f"{nm}: persistent max rel err {err:.2e}", flush=True)
        assert err < 2e-2
        del a, b, want, got
        for pers in (2, 1):
            for stages in ((0, 3) if pers == 1 else (0, 4)):
                case(nm, M, N, K, out_f32=f32, transpose=tr, persistent=pers, stages=stages)
elif args.sweep and os.environ.get("SWEEP") == "shipped":
    # reference-as-shipped shapes (E = 2048, 1024-token micro-batches, 4096-token mini-batch): N-tile width vs wave quantisation
    E2, T2, TB = 2048, 1024, 4096
    for (nm, M, N, K, f32, tr) in (("qkv fwd", 3 * E2, T2, E2, False, True), ("out-proj / ffn fwd", E2, T2, E2, False, True),
                                   ("qkv dgrad", E2, T2, 3 * E2, False, True), ("wgrad 3ExE", 3 * E2, E2, TB, True, False),
                                   ("wgrad ExE", E2, E2, TB, True, False), ("vocab fwd", V, T2, E2, False, True)):
        for bn in (tuple(int(x) for x in os.environ['BNS'].split(',')) if os.environ.get('BNS') else (0, 256, 192, 128, 64)):
            case(nm, M, N, K, out_f32=f32, transpose=tr, block_n=bn)
elif args.sweep and os.environ.get("SWEEP") == "wgrad":
    for bn in (256, 128, 64):
        for stages in (0, 2, 4):
            case("sweep wgrad ExE", E, E, 512, out_f32=True, transpose=False, block_n=bn, stages=stages)
elif args.sweep:
    for splits in (1, 2, 4, 8, 16):
        for stages in (0, 3):
            case("sweep out-proj fwd", E, T, E, splits=splits, stages=stages)
    for splits in (1, 2, 4):
        case("sweep qkv fwd", 3 * E, T, E, splits=splits)
else:
    case("fwd QKV   (W[3E,E] x tokens)", 3 * E, T, E)
    case("fwd out-proj / FFN", E, T, E)
    case("dgrad QKV (Wt[E,3E] x tokens)", E, T, 3 * E)
    case("fwd vocab (W[V,E] x tokens)", V, T, E)
    case("dgrad vocab (Wt[E,V] x tokens)", E, T, V)
    case("wgrad E x E, K = 500 tokens", E, E, 512, out_f32=True, transpose=False)
    case("wgrad 3E x E", 3 * E, E, 512, out_f32=True, transpose=False)
    case("wgrad vocab", V, E, 512, out_f32=True, transpose=False)
    case("shipped cfg: 1024 tokens, E=2048", 3 * 2048, 1024, 2048)
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
open(args.out, "a").write("\n".join(lines) + "\n")
