// Tensor-core attention core for short sequences (S <= 128 while the tiles fit shared memory): the five small matrix products of
// attention forward + backward on warp-level MMA (bf16 in, fp32 accumulate), everything else —
// softmax, dropout mask, the dS recurrence — on the scores held in shared memory.
//
// One CTA owns one (sequence, head) pair for the whole pass; Q, K, V (and dO) are gathered from the
// packed [T, 3E] projection output into zero-padded shared tiles, the S x S score matrix never
// leaves the SM, and results are scattered straight back into the packed layouts the neighbouring
// GEMMs read.  The SIMT kernels in attention.cu remain the fallback when the tiles do not fit
// (S = 128 with head_dim > 64).
//
// Why warp-level MMA and not tcgen05 here: at S = 25 the whole problem is a 32 x 32 x 256 product
// per head; a UMMA tile is 128 rows and needs TMA-swizzled operands, TMEM allocation and an mbarrier
// pipeline — fixed costs that exceed the work.  These products are latency-, not throughput-bound
// (attention is <= 0.2 % of the model's FLOPs).
//
// Reference counterpart: nn.MultiheadAttention core (main.py:148).  SURVEY.md K3.
#include <mma.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

using namespace nvcuda;

constexpr int kTcThreads = 256;

// debug: phase timestamps of CTA 0 of the last tensor-core attention launch (globaltimer ns)
__device__ unsigned long long g_attn_ts[16];
__device__ __forceinline__ void attn_ts(int slot) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_attn_ts[slot] = t;
    }
}
constexpr int kTcWarps = kTcThreads / 32;

struct AttnTc {
    int NB, S, H, D, E, SP;          // SP = S rounded up to 16
    float scale;
    int causal;
    float drop_p;
    unsigned long long seed, rng_stream;
};

typedef wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> FragA;
typedef wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> FragAT;
typedef wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> FragB;
typedef wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> FragBT;
typedef wmma::fragment<wmma::accumulator, 16, 16, 16, float> FragC;

// Gather up to four [S, D] head slices into zero-padded [SP, D + 8] shared tiles.  A warp covers
// whole rows (lane -> 16-byte vector of the row), four row-groups are requested before the first
// shared store, and the loop carries no integer division: the kernel is latency-bound, so both the
// L2 round trips and the index arithmetic are on its critical path.
__device__ __forceinline__ void tc_gather(const __nv_bfloat16* b0, const __nv_bfloat16* b1, const __nv_bfloat16* b2,
                                          const __nv_bfloat16* b3, long long ld012, long long ld3, __nv_bfloat16* d0,
                                          __nv_bfloat16* d1, __nv_bfloat16* d2, __nv_bfloat16* d3, int count, int NB, int S,
                                          int SP, int D, int n) {
    const int dpad = D + 8, vec_per_row = D / 8;              // 8 .. 32 vectors per row
    const int lanes_per_row = vec_per_row >= 32 ? 32 : vec_per_row;
    const int rows_per_warp = 32 / lanes_per_row;             // rows one warp instruction covers
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = lane / lanes_per_row, v = lane - sub * lanes_per_row;
    const int total_rows = count * SP;
    const int step = kTcWarps * rows_per_warp;
    constexpr int kBatch = 4;
    for (int r0 = warp * rows_per_warp + sub; r0 < total_rows; r0 += step * kBatch) {
        uint4 raw[kBatch];
#pragma unroll
        for (int u = 0; u < kBatch; ++u) {
            const int R = r0 + u * step;
            raw[u] = make_uint4(0u, 0u, 0u, 0u);
            if (R < total_rows && v < vec_per_row) {
                const int m = (R >= SP) + (R >= 2 * SP) + (R >= 3 * SP);
                const int s = R - m * SP;
                const __nv_bfloat16* base = m == 0 ? b0 : m == 1 ? b1 : m == 2 ? b2 : b3;
                const long long ld = m == 3 ? ld3 : ld012;
                if (s < S) raw[u] = *reinterpret_cast<const uint4*>(base + (static_cast<size_t>(s) * NB + n) * ld + v * 8);
            }
        }
#pragma unroll
        for (int u = 0; u < kBatch; ++u) {
            const int R = r0 + u * step;
            if (R < total_rows && v < vec_per_row) {
                const int m = (R >= SP) + (R >= 2 * SP) + (R >= 3 * SP);
                const int s = R - m * SP;
                __nv_bfloat16* dst = m == 0 ? d0 : m == 1 ? d1 : m == 2 ? d2 : d3;
                *reinterpret_cast<uint4*>(dst + s * dpad + v * 8) = raw[u];
            }
        }
    }
}

// C[SP x SP] (fp32, ld = SP + 8) = A[SP x D] * B[SP x D]^T, both row-major bf16 with ld = D + 8.
__device__ __forceinline__ void tc_abt(const __nv_bfloat16* A, const __nv_bfloat16* B, float* C, int SP, int D) {
    const int warp = threadIdx.x >> 5, tiles = SP / 16, dpad = D + 8, cpad = SP + 8;
    for (int t = warp; t < tiles * tiles; t += kTcWarps) {
        const int ti = t / tiles, tj = t % tiles;
        FragC acc;
        wmma::fill_fragment(acc, 0.f);
        for (int k = 0; k < D; k += 16) {
            FragA a;
            FragBT b;
            wmma::load_matrix_sync(a, A + ti * 16 * dpad + k, dpad);
            wmma::load_matrix_sync(b, B + tj * 16 * dpad + k, dpad);
            wmma::mma_sync(acc, a, b, acc);
        }
        wmma::store_matrix_sync(C + ti * 16 * cpad + tj * 16, acc, cpad, wmma::mem_row_major);
    }
}

// Softmax over the first S columns of the first S rows of `sc` (scaled by `scale`), in place.
// Eight lanes share a row, so 32 rows are in flight at once (a warp-per-row loop would walk the rows
// one latency chain after another).  SP <= 128: at most 16 columns per lane.
__device__ __forceinline__ void tc_softmax(float* sc, int S, int SP, float scale, int causal) {
    const int sub = threadIdx.x & 7, cpad = SP + 8;
    for (int i = threadIdx.x >> 3; i < SP; i += kTcThreads / 8) {
        float* r = sc + i * cpad;
        const int lim = i < S ? (causal ? i + 1 : S) : 0;
        float x[16];
        float mx = -INFINITY;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = sub + u * 8;
            x[u] = (j < lim) ? r[j] * scale : -INFINITY;
            mx = fmaxf(mx, x[u]);
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        float se = 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            x[u] = (sub + u * 8 < lim) ? __expf(x[u] - mx) : 0.f;
            se += x[u];
        }
        se += __shfl_xor_sync(0xffffffffu, se, 4);
        se += __shfl_xor_sync(0xffffffffu, se, 2);
        se += __shfl_xor_sync(0xffffffffu, se, 1);
        const float inv = lim > 0 ? 1.f / se : 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = sub + u * 8;
            if (j < SP) r[j] = x[u] * inv;
        }
    }
}

__device__ __forceinline__ float tc_keep(const CounterRng& rng, const AttnTc& a, uint32_t thr, float inv_keep, int n, int h,
                                         int i, int j) {
    if (a.drop_p <= 0.f) return 1.f;
    const uint64_t idx = ((static_cast<uint64_t>(n) * a.H + h) * a.S + i) * a.S + j;
    return dropout_keep(rng, a.rng_stream, idx, thr) ? inv_keep : 0.f;
}

// out rows (< S) of a [16 x 16] fp32 tile held in this warp's staging area -> bf16 global
__device__ __forceinline__ void tc_emit_tile(const float* stage, __nv_bfloat16* dst_row0, long long ld_dst, int NB,
                                             int row0, int S) {
    const int lane = threadIdx.x & 31;
    // 16 rows x 8 column pairs = 128 pairs, 4 per lane
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int p = lane + q * 32, r = p >> 3, c2 = p & 7;
        if (row0 + r < S) {
            const float2 v = *reinterpret_cast<const float2*>(stage + r * 24 + c2 * 2);
            reinterpret_cast<__nv_bfloat162*>(dst_row0 + static_cast<size_t>(r) * NB * ld_dst)[c2] =
                __floats2bfloat162_rn(v.x, v.y);
        }
    }
}

__global__ void __launch_bounds__(kTcThreads)
attn_tc_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, const AttnTc a) {
    pdl_launch_dependents();
    attn_ts(0);
    extern __shared__ __align__(128) uint8_t smem[];
    const int S = a.S, SP = a.SP, D = a.D, dpad = D + 8, cpad = SP + 8;
    __nv_bfloat16* q = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* k = q + SP * dpad;
    __nv_bfloat16* v = k + SP * dpad;
    float* sc = reinterpret_cast<float*>(v + SP * dpad);
    __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(sc + SP * cpad);
    float* stage = reinterpret_cast<float*>(pb + SP * cpad) + (threadIdx.x >> 5) * (16 * 24);
    const int n = blockIdx.x / a.H, h = blockIdx.x % a.H;
    const long long ld = 3LL * a.E;

    tc_gather(qkv + h * D, qkv + a.E + h * D, qkv + 2 * a.E + h * D, nullptr, ld, 0, q, k, v, nullptr, 3, a.NB, S, SP, D, n);
    __syncthreads();
    attn_ts(1);
    tc_abt(q, k, sc, SP, D);
    __syncthreads();
    attn_ts(2);
    tc_softmax(sc, S, SP, a.scale, a.causal);
    __syncthreads();
    attn_ts(3);
    {
        const CounterRng rng(a.seed);
        const uint32_t thr = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
        const float inv_keep = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
        for (int p = threadIdx.x; p < SP * SP; p += blockDim.x) {
            const int i = p / SP, j = p % SP;
            float w = sc[i * cpad + j];
            if (i < S && j < S) w *= tc_keep(rng, a, thr, inv_keep, n, h, i, j);
            pb[i * cpad + j] = __float2bfloat16(w);
        }
    }
    __syncthreads();
    attn_ts(4);

    // O = P V: each warp keeps up to four output tiles in flight (independent accumulators)
    const int warp = threadIdx.x >> 5, rt = SP / 16, ct = D / 16, total = rt * ct;
    for (int t0 = warp; t0 < total; t0 += kTcWarps * 4) {
        FragC acc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) wmma::fill_fragment(acc[u], 0.f);
        for (int kk = 0; kk < SP; kk += 16) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * kTcWarps;
                if (t < total) {
                    FragA pa;
                    FragB vb;
                    wmma::load_matrix_sync(pa, pb + (t / ct) * 16 * cpad + kk, cpad);
                    wmma::load_matrix_sync(vb, v + kk * dpad + (t % ct) * 16, dpad);
                    wmma::mma_sync(acc[u], pa, vb, acc[u]);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + u * kTcWarps;
            if (t < total) {
                const int ti = t / ct, tc = t % ct;
                wmma::store_matrix_sync(stage, acc[u], 24, wmma::mem_row_major);
                __syncwarp();
                tc_emit_tile(stage, out + (static_cast<size_t>(ti * 16) * a.NB + n) * a.E + h * D + tc * 16, a.E, a.NB,
                             ti * 16, S);
                __syncwarp();
            }
        }
    }
    attn_ts(5);
}

__global__ void __launch_bounds__(kTcThreads)
attn_tc_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dout,
                   __nv_bfloat16* __restrict__ dqkv, const AttnTc a) {
    pdl_launch_dependents();
    extern __shared__ __align__(128) uint8_t smem[];
    const int S = a.S, SP = a.SP, D = a.D, dpad = D + 8, cpad = SP + 8;
    __nv_bfloat16* q = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* k = q + SP * dpad;
    __nv_bfloat16* v = k + SP * dpad;
    __nv_bfloat16* dO = v + SP * dpad;
    float* P = reinterpret_cast<float*>(dO + SP * dpad);
    float* dP = P + SP * cpad;
    // The bf16 matrices the last three products read are written over the fp32 rows they are derived
    // from (same row pitch in bytes, so bpad = 2 * cpad bf16 elements): S = 128 with D = 64 then
    // fits the 227 KB of shared memory.
    __nv_bfloat16* Pd = reinterpret_cast<__nv_bfloat16*>(P);                  // dropped-out probabilities
    __nv_bfloat16* dS = reinterpret_cast<__nv_bfloat16*>(dP);                 // score gradients
    const int bpad = 2 * cpad;
    float* stage = dP + SP * cpad + (threadIdx.x >> 5) * (16 * 24);
    const int n = blockIdx.x / a.H, h = blockIdx.x % a.H;
    const long long ld = 3LL * a.E;

    tc_gather(qkv + h * D, qkv + a.E + h * D, qkv + 2 * a.E + h * D, dout + h * D, ld, a.E, q, k, v, dO, 4, a.NB, S, SP,
              D, n);
    __syncthreads();
    tc_abt(q, k, P, SP, D);          // raw scores
    tc_abt(dO, v, dP, SP, D);        // gradient w.r.t. the dropped-out probabilities
    __syncthreads();
    tc_softmax(P, S, SP, a.scale, a.causal);
    __syncthreads();

    // per row (8 lanes each): Pd = P * keep,  dPm = dP * keep,  delta = sum_j dPm * P,  dS = P * (dPm - delta) * scale
    {
        const CounterRng rng(a.seed);
        const uint32_t thr = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
        const float inv_keep = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
        const int sub = threadIdx.x & 7;
        for (int i = threadIdx.x >> 3; i < SP; i += kTcThreads / 8) {
            float p[16], dpm[16], keepv[16];
            float dlt = 0.f;
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const int j = sub + u * 8;
                p[u] = 0.f;
                dpm[u] = 0.f;
                keepv[u] = 0.f;
                if (j < SP) {
                    p[u] = P[i * cpad + j];
                    float ks = 0.f;
                    if (i < S && j < S) ks = tc_keep(rng, a, thr, inv_keep, n, h, i, j);
                    dpm[u] = dP[i * cpad + j] * ks;
                    dlt += dpm[u] * p[u];
                    keepv[u] = ks;
                }
            }
            dlt += __shfl_xor_sync(0xffffffffu, dlt, 4);
            dlt += __shfl_xor_sync(0xffffffffu, dlt, 2);
            dlt += __shfl_xor_sync(0xffffffffu, dlt, 1);
            __syncwarp();          // every lane of the row has read its fp32 values: rows may now be overwritten
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const int j = sub + u * 8;
                if (j < SP) {
                    Pd[i * bpad + j] = __float2bfloat16(p[u] * keepv[u]);
                    dS[i * bpad + j] = __float2bfloat16(p[u] * (dpm[u] - dlt) * a.scale);
                }
            }
        }
    }
    __syncthreads();

    // dV = Pd^T dO,  dQ = dS K,  dK = dS^T Q  — 3 x (SP/16) x (D/16) output tiles, four in flight per warp
    const int warp = threadIdx.x >> 5, rt = SP / 16, ct = D / 16, per = rt * ct, total = 3 * per;
    for (int t0 = warp; t0 < total; t0 += kTcWarps * 4) {
        FragC acc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) wmma::fill_fragment(acc[u], 0.f);
        for (int kk = 0; kk < SP; kk += 16) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * kTcWarps;
                if (t < total) {
                    const int which = t / per, r = t % per, ti = r / ct, tc = r % ct;
                    FragB bm;
                    if (which == 1) {            // dQ: A = dS (row-major), B = K
                        FragA am;
                        wmma::load_matrix_sync(am, dS + ti * 16 * bpad + kk, bpad);
                        wmma::load_matrix_sync(bm, k + kk * dpad + tc * 16, dpad);
                        wmma::mma_sync(acc[u], am, bm, acc[u]);
                    } else {                     // dV: A = Pd^T, B = dO;  dK: A = dS^T, B = Q
                        FragAT at;
                        wmma::load_matrix_sync(at, (which == 0 ? Pd : dS) + kk * bpad + ti * 16, bpad);
                        wmma::load_matrix_sync(bm, (which == 0 ? dO : q) + kk * dpad + tc * 16, dpad);
                        wmma::mma_sync(acc[u], at, bm, acc[u]);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + u * kTcWarps;
            if (t < total) {
                const int which = t / per, r = t % per, ti = r / ct, tc = r % ct;
                wmma::store_matrix_sync(stage, acc[u], 24, wmma::mem_row_major);
                __syncwarp();
                const int col0 = (which == 0 ? 2 * a.E : which == 1 ? 0 : a.E) + h * D + tc * 16;
                tc_emit_tile(stage, dqkv + (static_cast<size_t>(ti * 16) * a.NB + n) * ld + col0, ld, a.NB, ti * 16, S);
                __syncwarp();
            }
        }
    }
}

size_t attn_tc_smem(int SP, int D, bool bwd) {
    const size_t dpad = D + 8, cpad = SP + 8;
    size_t b = (bwd ? 4 : 3) * SP * dpad * 2;             // Q K V (dO)
    b += (bwd ? 2 : 1) * SP * cpad * 4;                    // fp32 score matrices (backward: bf16 copies alias them)
    if (!bwd) b += SP * cpad * 2;                          // forward: bf16 probabilities
    b += kTcWarps * 16 * 24 * 4;                           // per-warp output staging
    return b + 128;
}

bool attn_tc_supported(int S, int D) {
    if (S > 128 || D % 16 != 0 || D > 256) return false;
    const int SP = (S + 15) / 16 * 16;
    return attn_tc_smem(SP, D, true) <= 227 * 1024;        // e.g. S = 128 needs D <= 64; S <= 64 any D <= 256
}

int attn_tc_fwd(const void* qkv, void* out, int NB, int S, int H, int D, int causal, float drop_p,
                unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream) {
    const int SP = (S + 15) / 16 * 16;
    AttnTc a{NB, S, H, D, H * D, SP, 1.f / sqrtf(static_cast<float>(D)), causal, drop_p, seed, rng_stream};
    const size_t smem = attn_tc_smem(SP, D, false);
    if (ensure_max_smem(reinterpret_cast<const void*>(attn_tc_fwd_kernel), 227 * 1024)) return -1;
    attn_tc_fwd_kernel<<<NB * H, kTcThreads, smem, stream>>>(static_cast<const __nv_bfloat16*>(qkv),
                                                             static_cast<__nv_bfloat16*>(out), a);
    PB2_LAUNCHED("attn_tc_fwd_kernel");
    return 0;
}

int attn_tc_bwd(const void* qkv, const void* dout, void* dqkv, int NB, int S, int H, int D, int causal, float drop_p,
                unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream) {
    const int SP = (S + 15) / 16 * 16;
    AttnTc a{NB, S, H, D, H * D, SP, 1.f / sqrtf(static_cast<float>(D)), causal, drop_p, seed, rng_stream};
    const size_t smem = attn_tc_smem(SP, D, true);
    if (ensure_max_smem(reinterpret_cast<const void*>(attn_tc_bwd_kernel), 227 * 1024)) return -1;
    attn_tc_bwd_kernel<<<NB * H, kTcThreads, smem, stream>>>(static_cast<const __nv_bfloat16*>(qkv),
                                                             static_cast<const __nv_bfloat16*>(dout),
                                                             static_cast<__nv_bfloat16*>(dqkv), a);
    PB2_LAUNCHED("attn_tc_bwd_kernel");
    return 0;
}

}  // namespace pb2

extern "C" int pb2_debug_attn_timestamps(unsigned long long* out16) {
    return cudaMemcpyFromSymbol(out16, pb2::g_attn_ts, sizeof(unsigned long long) * 16) == cudaSuccess ? 0 : -1;
}
