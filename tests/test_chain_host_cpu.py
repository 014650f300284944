"""Host-side validation of the op-chain entry point (``pb2_chain_supported``): pure C++ bookkeeping, no GPU needed — the
CPU tier checks that a chain the kernel cannot run is refused before anything is launched (``csrc/chain_sm100.cu``,
``chain_reject``).  Reference counterpart: none (the reference launches every op as its own library call, main.py:148)."""
import ctypes

import pytest

from pipe_b200.ops import _lib
from pipe_b200.ops.chain import MAX_OPS, ChainOp

try:
    LIB = _lib.load()
except Exception as exc:      # pragma: no cover - the CPU tier of this repository always has the built library
    pytest.skip(f"native library not available: {exc}", allow_module_level=True)


def _gemm(op, M, N, K, *, a=0x10000, b=0x20000, lda=None, ldb=None):
    op.kind = 0
    d = op.gemm
    d.a, d.b, d.out = a, b, 0x30000
    d.M, d.N, d.K = M, N, K
    d.lda, d.ldb, d.ld_out = lda or K, ldb or K, M
    d.transpose_out, d.bias_along_m, d.accumulate, d.sumsq = 1, 1, 0, None
    return op


def _ln(op, rows, E):
    op.kind = 1
    op.ln.rows, op.ln.E = rows, E
    return op


def _ok(ops, n):
    return bool(LIB.pb2_chain_supported(ops, ctypes.c_int(n)))


def test_a_layer_tail_is_accepted():
    ops = (ChainOp * MAX_OPS)()
    _gemm(ops[0], 4096, 75, 4096)
    _ln(ops[1], 75, 4096)
    _gemm(ops[2], 4096, 75, 4096)
    _gemm(ops[3], 4096, 75, 4096)
    _ln(ops[4], 75, 4096)
    _gemm(ops[5], 12288, 75, 4096)
    assert _ok(ops, 6)


@pytest.mark.parametrize("why,mutate", [
    ("more than 128 tokens", lambda ops: _gemm(ops[0], 4096, 129, 4096)),
    ("reduction too short for four ranks", lambda ops: _gemm(ops[0], 4096, 75, 128)),
    ("row stride not a multiple of 8 elements", lambda ops: _gemm(ops[0], 4096, 75, 4096, lda=4100)),
    ("operand not 16-byte aligned", lambda ops: _gemm(ops[0], 4096, 75, 4096, a=0x10008)),
    ("GEMMs on different token counts", lambda ops: _gemm(ops[2], 4096, 64, 4096)),
    ("accumulating epilogue", lambda ops: setattr(ops[0].gemm, "accumulate", 1)),
    ("bias along tokens", lambda ops: (setattr(ops[0].gemm, "bias", 0x40000), setattr(ops[0].gemm, "bias_along_m", 0))),
    ("LayerNorm wider than the register tile", lambda ops: _ln(ops[1], 75, 16384)),
    ("unknown op kind", lambda ops: setattr(ops[1], "kind", 7)),
])
def test_chains_the_kernel_cannot_run_are_refused(why, mutate):
    ops = (ChainOp * MAX_OPS)()
    _gemm(ops[0], 4096, 75, 4096)
    _ln(ops[1], 75, 4096)
    _gemm(ops[2], 4096, 75, 4096)
    assert _ok(ops, 3)
    mutate(ops)
    assert not _ok(ops, 3), why


def test_op_count_limits():
    ops = (ChainOp * MAX_OPS)()
    for i in range(MAX_OPS):
        _ln(ops[i], 75, 4096)
    assert not _ok(ops, MAX_OPS)          # no GEMM at all
    assert not _ok(ops, 0)
    for i in range(7):
        _gemm(ops[i], 4096, 75, 4096)
    assert not _ok(ops, 7)                # more GEMMs than the kernel has tensor maps for
    assert _ok(ops, 6)
