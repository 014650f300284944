"""include/pipe_b200.h is the whole C interface of the native library: it has to be valid C (not only
C++), and the struct layouts the ctypes binding declares must match what a C compiler lays out."""
import ctypes
import os
import shutil
import subprocess
import tempfile
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_INC = "/usr/local/cuda/include"


def _have_toolchain():
    return shutil.which("gcc") and os.path.exists(os.path.join(CUDA_INC, "cuda_runtime_api.h"))


@pytest.mark.skipif(not _have_toolchain(), reason="needs gcc and the CUDA headers")
@pytest.mark.parametrize("lang,compiler", [("c", "gcc"), ("c++", "g++")])
def test_header_is_valid_c_and_cpp(lang, compiler):
    res = subprocess.run([compiler, "-fsyntax-only", "-Wall", "-Werror", "-I", CUDA_INC, "-x", lang,
                          os.path.join(ROOT, "include", "pipe_b200.h")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


@pytest.mark.skipif(not _have_toolchain(), reason="needs gcc and the CUDA headers")
def test_ctypes_struct_layouts_match_the_header():
    from pipe_b200.ops import _lib
    src = textwrap.dedent("""
        #include <stdio.h>
        #include <stddef.h>
        #include "pipe_b200.h"
        int main(void) {
            printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(pb2_gemm_desc), offsetof(pb2_gemm_desc, persistent),
                   sizeof(pb2_encoder_layer), offsetof(pb2_encoder_layer, side_stream),
                   offsetof(pb2_encoder_layer, peer_delta), sizeof(pb2_encoder_stack),
                   offsetof(pb2_encoder_stack, attn_scratch_bytes), sizeof(pb2_transpose_jobs),
                   sizeof(pb2_chain_op), offsetof(pb2_chain_op, gemm), offsetof(pb2_chain_op, ln), offsetof(pb2_chain_ln, eps));
            return 0;
        }
    """)
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "layout.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "layout")
        res = subprocess.run(["gcc", "-I", CUDA_INC, "-I", os.path.join(ROOT, "include"), c, "-o", exe],
                             capture_output=True, text=True)
        assert res.returncode == 0, res.stderr
        out = subprocess.run([exe], capture_output=True, text=True).stdout.split()
    (gemm_size, gemm_off, layer_size, layer_off, peer_off, stack_size, stack_off, jobs_size,
     chain_size, chain_gemm_off, chain_ln_off, ln_eps_off) = map(int, out)
    from pipe_b200.ops import chain
    assert ctypes.sizeof(chain.ChainOp) == chain_size
    assert chain.ChainOp.gemm.offset == chain_gemm_off and chain.ChainOp.ln.offset == chain_ln_off
    assert chain.ChainLn.eps.offset == ln_eps_off
    from pipe_b200.ops import functional as F2
    assert _lib.EncoderLayerDesc.peer_delta.offset == peer_off
    assert ctypes.sizeof(_lib.EncoderStackDesc) == stack_size
    assert _lib.EncoderStackDesc.attn_scratch_bytes.offset == stack_off
    assert ctypes.sizeof(F2._TransposeJobs) == jobs_size
    assert ctypes.sizeof(_lib.GemmDesc) == gemm_size
    assert _lib.GemmDesc.persistent.offset == gemm_off
    assert ctypes.sizeof(_lib.EncoderLayerDesc) == layer_size
    assert _lib.EncoderLayerDesc.side_stream.offset == layer_off
