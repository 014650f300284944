"""The parts of the native library that run on the host only — routing decisions, buffer-size formulas, argument
validation — checked on the CPU tier (no kernel is launched; ``include/pipe_b200.h`` is the contract).
Reference counterpart: none (the reference has no native code)."""
import ctypes

import pytest

from pipe_b200.ops import _lib

try:
    LIB = _lib.load()
except Exception as exc:      # pragma: no cover
    pytest.skip(f"native library not available: {exc}", allow_module_level=True)

LIB.pb2_gemm_workspace_bytes.restype = ctypes.c_size_t
LIB.pb2_gemm_noalloc_max_scratch_bytes.restype = ctypes.c_size_t
LIB.pb2_flash_attention_bwd_scratch_bytes.restype = ctypes.c_size_t
LIB.pb2_encoder_stack_act_elems.restype = ctypes.c_size_t
LIB.pb2_encoder_stack_tmp_elems.restype = ctypes.c_size_t
LIB.pb2_encoder_stack_stats_floats.restype = ctypes.c_size_t
LIB.pb2_last_error.restype = ctypes.c_char_p


def test_attention_routing(monkeypatch):
    monkeypatch.delenv("PB2_ATTN_FLASH", raising=False)
    flash = lambda S, D: bool(LIB.pb2_attention_uses_flash(ctypes.c_int(S), ctypes.c_int(D)))
    assert not flash(25, 256)            # config 3: 25-token sequences, head_dim 256 -> tile kernels
    assert flash(128, 64)                # the reference as shipped -> flash
    assert flash(2048, 128)              # anything longer than one tile -> flash
    assert not flash(200, 256)           # no flash backward for head_dim 256
    assert not flash(32, 64)             # short sequences stay on the tile kernels


def test_workspace_and_scratch_sizes():
    ws = int(LIB.pb2_gemm_workspace_bytes())
    assert ws == 96 * 1024 * 1024 + 64 * 1024
    assert int(LIB.pb2_gemm_noalloc_max_scratch_bytes()) == 16 * 1024 * 1024
    small = int(LIB.pb2_flash_attention_bwd_scratch_bytes(ctypes.c_int(8), ctypes.c_int(128), ctypes.c_int(32), ctypes.c_int(64)))
    big = int(LIB.pb2_flash_attention_bwd_scratch_bytes(ctypes.c_int(8), ctypes.c_int(2048), ctypes.c_int(32), ctypes.c_int(64)))
    assert 0 < small < big and small % 16 == 0
    # fp32 dQ accumulator [tokens, H * D] is the bulk of it
    assert big >= 8 * 2048 * 32 * 64 * 4


def test_stack_slab_formulas():
    n, T, S, E, FF, H = 3, 75, 25, 4096, 4096, 16
    layers = (_lib.EncoderLayerDesc * n)()
    for l in layers:
        l.E, l.FF, l.heads = E, FF, H
    act = int(LIB.pb2_encoder_stack_act_elems(layers, ctypes.c_int(n), ctypes.c_int(T)))
    tmp = int(LIB.pb2_encoder_stack_tmp_elems(layers, ctypes.c_int(n), ctypes.c_int(T)))
    stats = int(LIB.pb2_encoder_stack_stats_floats(layers, ctypes.c_int(n), ctypes.c_int(S), ctypes.c_int(T)))
    assert act == n * T * (7 * E + FF)           # qkv (3E) + a + o + x1 + f (4E) + hmid (FF) per layer
    assert tmp == n * T * (10 * E + FF)          # dr2 df dx1 dr1 do da dx (7E) + dqkv (3E) + dz1 (FF)
    assert stats == n * 4 * T                    # two LayerNorms x (mean, rstd); no log-sum-exp on tile-attention shapes
    for l in layers:
        l.heads = 64                             # head_dim 64 at S = 128 routes to flash: + H floats per token
    stats_flash = int(LIB.pb2_encoder_stack_stats_floats(layers, ctypes.c_int(n), ctypes.c_int(128), ctypes.c_int(1024)))
    assert stats_flash == n * (4 * 1024 + 64 * 1024)


def test_gemm_rejects_bad_descriptors_before_touching_the_gpu():
    d = _lib.GemmDesc()
    d.a, d.b, d.out = 0x10000, 0x20000, 0x30000
    d.M, d.N, d.K = 0, 75, 4096
    d.lda = d.ldb = 4096
    d.ld_out = 128
    assert LIB.pb2_gemm_bf16(ctypes.byref(d), None, ctypes.c_size_t(0), None) == -1
    assert b"non-positive shape" in LIB.pb2_last_error()
    d.M = 128
    d.lda = 4100
    assert LIB.pb2_gemm_bf16(ctypes.byref(d), None, ctypes.c_size_t(0), None) == -1
    assert b"multiples of 8" in LIB.pb2_last_error()
    d.lda = 4096
    d.a = 0x10004
    assert LIB.pb2_gemm_bf16(ctypes.byref(d), None, ctypes.c_size_t(0), None) == -1
    assert b"16-byte aligned" in LIB.pb2_last_error()
    assert LIB.pb2_gemm_bf16(None, None, ctypes.c_size_t(0), None) == -1
    assert b"null descriptor" in LIB.pb2_last_error()
