/* pipe_b200 native kernels — C ABI.
 *
 * One shared library (pipe_b200/_native/libpipe_b200.so, sm_100a only).  Every
 * entry point takes raw device pointers and a cudaStream_t, launches on that
 * stream and returns 0 on success, -1 on failure (pb2_last_error() says why).
 * Nothing here allocates device memory unless stated: scratch is passed in by
 * the caller.  The Python package binds these with ctypes (pipe_b200/ops/_lib.py);
 * other runtimes can link the library directly.
 *
 * Dropout everywhere is counter-based: Philox4x32-10 keyed by `seed`, with the
 * element's flat index in the output tensor as counter and `rng_stream`
 * separating call sites, so a backward kernel (or a checkpoint recompute) given
 * the same (seed, rng_stream) regenerates the identical mask.
 */
#ifndef PIPE_B200_H_
#define PIPE_B200_H_

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------ */
const char* pb2_last_error(void);          /* message of the last failure on this thread */
long long pb2_launch_count(void);          /* kernels launched by this library so far */
int pb2_version(void);

/* ---- GEMM: D[M,N] = A[M,K] * B[N,K]^T, bf16 in, fp32 accumulate (tcgen05 + TMEM + TMA) ----
 * Reference counterpart: the cuBLAS calls behind nn.Linear / nn.MultiheadAttention
 * (main.py:45,148); SURVEY.md K2, K4, K6, K7. */
#define PB2_GEMM_COUNTER_BYTES (64u * 1024u)
#define PB2_GEMM_WORKSPACE_BYTES (96u * 1024u * 1024u + PB2_GEMM_COUNTER_BYTES)

typedef struct pb2_gemm_desc {
    const void* a;        /* bf16 [M, K], row stride lda (elements, multiple of 8): rides the UMMA M axis */
    const void* b;        /* bf16 [N, K], row stride ldb: rides the UMMA N axis */
    int M, N, K;
    long long lda, ldb;
    void* out;            /* bf16 or fp32 */
    long long ld_out;     /* row stride of the output tensor */
    int out_f32;          /* 1: fp32 output, 0: bf16 */
    int transpose_out;    /* 1: element (m, n) is stored at out[n * ld_out + m] (swap-AB) */
    const float* bias;    /* optional fp32 bias */
    int bias_along_m;     /* bias index is m (1) or n (0) */
    int relu;             /* apply max(0, x) after bias */
    float drop_p;         /* dropout probability applied after bias/ReLU (0 = off) */
    unsigned long long seed, rng_stream;
    const void* mask_src; /* optional bf16, output-shaped: out *= (mask_src > 0) * mask_scale  (ReLU+dropout backward) */
    float mask_scale;
    const void* addend;   /* optional bf16, output-shaped: out += addend (residual gradient) */
    int accumulate;       /* fp32 output only: out += result */
    float alpha;          /* scales the accumulator first (0 means 1) */
    float* sumsq;         /* optional device scalar: += sum of squares of the stored values (grad-norm fusion) */
    int block_n;          /* N tile, multiple of 16 in [16,256]; 0 = choose */
    int splits;           /* split-K factor; 0 = choose */
} pb2_gemm_desc;

size_t pb2_gemm_workspace_bytes(void);
/* `workspace` (device, zero-initialised once, reusable across launches on one stream) holds the
 * split-K partial tiles and per-tile arrival counters; may be NULL when splits == 1. */
int pb2_gemm_bf16(const pb2_gemm_desc* desc, void* workspace, size_t workspace_bytes, cudaStream_t stream);

#ifdef __cplusplus
}
#endif

#endif /* PIPE_B200_H_ */
