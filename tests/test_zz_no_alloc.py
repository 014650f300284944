"""The workspace-free GEMM entry point, driven through the C ABI.

``pb2_gemm_bf16_noalloc`` exists for hosts of the native library that have no scratch memory to
offer — their HBM belongs to another component's pool, or they may not call cudaMalloc in the code
path at hand, or they are plain C callers without an allocator.  Its split-K scratch is an array
compiled into the library, shared by every call.  This file checks, through ctypes only, that it
computes what ``pb2_gemm_bf16`` (caller-supplied workspace) computes, on two calls in a row with
different tilings so the second call meets the first call's leftovers, and that an oversized request
is refused by name.  The file sorts last on purpose: a run that stops at the first failure has by
then executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _desc(_lib, a, b, out, splits, block_n):
    d = _lib.GemmDesc()
    d.a, d.b = a.data_ptr(), b.data_ptr()
    d.M, d.K = a.shape
    d.N = b.size(0)
    d.lda, d.ldb = a.stride(0), b.stride(0)
    d.out, d.ld_out = out.data_ptr(), out.stride(0)
    d.out_f32 = 1
    d.splits, d.block_n = splits, block_n
    d.reduce_mode = 1          # the caller-workspace split-K path, like the entry point under test
    return d


def _run_both(M, N, K, splits, block_n, seed):
    from pipe_b200.ops import _lib
    lib = _lib.load()
    lib.pb2_gemm_bf16_noalloc.restype = ctypes.c_int
    lib.pb2_gemm_noalloc_max_scratch_bytes.restype = ctypes.c_size_t
    torch.manual_seed(seed)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    want = torch.empty(M, N, device="cuda")
    got = torch.full((M, N), float("nan"), device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    ws = torch.zeros(int(lib.pb2_gemm_workspace_bytes()), dtype=torch.uint8, device="cuda")
    d1 = _desc(_lib, a, b, want, splits, block_n)
    rc = lib.pb2_gemm_bf16(ctypes.byref(d1), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(ws.numel()), stream)
    assert rc == 0, lib.pb2_last_error().decode()

    d2 = _desc(_lib, a, b, got, splits, block_n)
    rc = lib.pb2_gemm_bf16_noalloc(ctypes.byref(d2), stream)
    assert rc == 0, lib.pb2_last_error().decode()
    torch.cuda.synchronize()
    return got, want, (a.float() @ b.float().t())


def test_noalloc_entry_matches_workspace_entry_twice_in_a_row():
    # call 1: 32 tiles x 8 splits, N tile 80;  call 2: other shape, 4 splits, N tile 48 — it finds
    # call 1's partial tiles (and re-cleared counters) in the shared array
    for (M, N, K, splits, bn, seed) in [(4096, 75, 4096, 8, 80, 0), (1024, 48, 2048, 4, 48, 1)]:
        got, want, ref = _run_both(M, N, K, splits, bn, seed)
        assert torch.isfinite(got).all()
        # same kernel, same split order: agree to fp32 reassociation; and both match the fp32 reference
        assert torch.allclose(got, want, rtol=1e-2, atol=1e-2 * (K ** 0.5))
        assert torch.allclose(got, ref, rtol=1e-2, atol=1e-2 * (K ** 0.5))


def test_noalloc_entry_refuses_more_scratch_than_it_has():
    from pipe_b200.ops import _lib
    lib = _lib.load()
    lib.pb2_gemm_bf16_noalloc.restype = ctypes.c_int
    lib.pb2_gemm_noalloc_max_scratch_bytes.restype = ctypes.c_size_t
    limit = int(lib.pb2_gemm_noalloc_max_scratch_bytes())
    M, N, K, splits = 8192, 256, 1024, 16            # 64 tiles x 16 splits x 128 KB = 128 MiB of partials
    assert (M // 128) * splits * 256 * 128 * 4 > limit
    a = torch.zeros(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(N, K, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(M, N, device="cuda")
    d = _desc(_lib, a, b, out, splits, 256)
    rc = lib.pb2_gemm_bf16_noalloc(ctypes.byref(d), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.pb2_last_error().decode()
    assert "built-in scratch is limited" in msg and str(limit) in msg
    torch.cuda.synchronize()
