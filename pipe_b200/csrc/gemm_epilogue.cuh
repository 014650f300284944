// Pieces of the tcgen05 GEMM shared by the one-GEMM kernels (gemm_sm100.cu) and the persistent op-chain kernel
// (chain_sm100.cu): tile constants, the kernel argument block, operand-stage helpers and the fused epilogue.
// Reference counterpart: none — bias, ReLU, dropout, residual add are separate ATen kernels behind nn.Linear /
// nn.TransformerEncoderLayer there (main.py:45,148).  SURVEY.md K2/K4/K6.
#pragma once

#include "common.cuh"

namespace pb2 {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;          // 64 bf16 = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kMaxStages = 8;
constexpr int kStagePitch = 20;         // floats per row of a staged 16-column chunk (80 B: 16-byte aligned, conflict-free)
constexpr int kEpilogueWarps = 8;       // two per TMEM lane quarter, alternating 16-column chunks
constexpr int kThreads = 64 + 32 * kEpilogueWarps;

struct GemmArgs {
    int M, N, K;
    int block_n;
    int k_blocks_total;
    int k_blocks_per_split;
    int splits;
    int stages;
    int ring_bytes;                  // bytes reserved for the TMA ring (>= the epilogue staging area)
    int tmem_cols;
    // epilogue
    void* out;
    long long ld_out;
    int out_f32;
    int transpose_out;
    const float* bias;
    int bias_along_m;
    int relu;
    float drop_p;
    unsigned long long seed, stream;
    const __nv_bfloat16* mask_src;   // out *= (mask_src[idx] > 0) * mask_scale
    float mask_scale;
    const __nv_bfloat16* addend;     // out += addend[idx]
    int accumulate;                  // fp32 out += (else overwrite)
    float alpha;                     // scales the accumulator before anything else
    float* sumsq;                    // optional: += sum of squares of what was written
    float* partials;
    unsigned int* counters;
    int a_static;                    // A (weights) is not produced by the preceding kernel: prefetch before the PDL wait
    int cluster_reduce;              // splits CTAs of a tile form a cluster and reduce through DSMEM
    int local_chunks;                // 16-column chunks each cluster rank owns (ceil(n_chunks / splits))
    unsigned long long* debug_ts;    // optional: per-CTA phase timestamps (globaltimer ns), 8 per CTA
    int tma_store;                   // persistent kernel: plain fp32 row-major result leaves through TMA stores
    int a_mn;                        // A is stored [K, M] (M contiguous): MN-major operand, two 64-column boxes per stage
};

// One stage of A.  K-major: one [128 rows, 64 k] box.  MN-major (A given as [K, M]): two [64 k-rows, 64 m-columns] boxes,
// 8 KB apart — the two 64-element MN atoms of the 128-row tile.
__device__ __forceinline__ void load_a_stage(uint8_t* sa, const CUtensorMap* tmap, uint64_t* bar, int k_elem, int m0,
                                             uint64_t hint, int a_mn) {
    if (a_mn) {
        tma_load_2d(sa, tmap, bar, m0, k_elem, hint);
        tma_load_2d(sa + 8192, tmap, bar, m0 + 64, k_elem, hint);
    } else {
        tma_load_2d(sa, tmap, bar, k_elem, m0, hint);
    }
}
__device__ __forceinline__ uint64_t a_desc(uint32_t sa, int a_mn) {
    return a_mn ? umma_desc_mn_sw128(sa, 8192) : umma_desc_k_sw128(sa);
}

__device__ __forceinline__ float bf16_bits_to_float(unsigned short b) {
    return __uint_as_float(static_cast<uint32_t>(b) << 16);
}

// Everything after the raw accumulator for one 16-column chunk of one accumulator row, written as
// straight-line predicated code: the epilogue runs with one or two warps per scheduler, so every
// dependent instruction or divergent branch is exposed latency; sixteen independent elements in
// flight hide it.  (m, n) are accumulator coordinates; the output element lives at
// out[n * ld + m] when the store is transposed, else out[m * ld + n].
struct EpiCtx {
    CounterRng rng;
    uint32_t thr;
    float inv_keep;
    float bias_m;        // bias for this thread's row when bias_along_m
    bool row_ok;
};

// Everything in a chunk's epilogue that does not depend on the accumulator — the dropout
// keep-bits and the mask / residual operands from global memory — gathered ahead of time, so that in
// the cluster path it is computed / in flight while the split-K partials cross the cluster instead of
// after them.
struct ChunkPre {
    uint32_t keep;       // bit j: element j survives dropout
    float mk[16];        // mask_src values
    float ad[16];        // addend values
    bool have;
};

// Dropout keep-bits (bit j: element idx0 + j * stride survives) of one thread's 16 elements.
__device__ __forceinline__ uint32_t dropout_bits_generic(const GemmArgs& g, const EpiCtx& e, long long idx0, long long stride) {
    uint32_t bits = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const uint64_t idx = static_cast<uint64_t>(idx0 + j * stride);
        const uint4 r = e.rng(idx >> 2, g.stream);
        const uint32_t sel = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
        bits |= (sel >= e.thr ? 1u : 0u) << j;
    }
    return bits;
}
// Transposed store: thread = accumulator row m, its 16 elements are ld_out apart, and one generator call covers the four
// elements idx & ~3 .. idx | 3 — which belong to four NEIGHBOURING LANES, not to this thread.  So each lane of an aligned
// group of four evaluates 4 of the 16 counters (for all four lanes) and the group exchanges the decisions with four
// shuffles: 4 generator calls per thread instead of 16, the same bits.  Must be called by all 32 lanes.
__device__ __forceinline__ uint32_t dropout_bits_transposed(const GemmArgs& g, const EpiCtx& e, long long idx0) {
    const int lane = threadIdx.x & 31, q = lane & 3;
    if ((g.ld_out & 3) || (static_cast<int>(idx0) & 3) != q) return dropout_bits_generic(g, e, idx0, g.ld_out);   // warp-uniform
    uint32_t mine = 0;                     // nibble i: decisions of counter j = q + 4 i for lanes base .. base + 3
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint64_t idx = static_cast<uint64_t>(idx0 + static_cast<long long>(q + 4 * i) * g.ld_out);
        const uint4 r = e.rng(idx >> 2, g.stream);
        mine |= ((r.x >= e.thr ? 1u : 0u) | (r.y >= e.thr ? 2u : 0u) | (r.z >= e.thr ? 4u : 0u) | (r.w >= e.thr ? 8u : 0u)) << (4 * i);
    }
    uint32_t bits = 0;
#pragma unroll
    for (int src = 0; src < 4; ++src) {
        const uint32_t theirs = __shfl_sync(0xffffffffu, mine, (lane & ~3) | src);
#pragma unroll
        for (int i = 0; i < 4; ++i) bits |= ((theirs >> (4 * i + q)) & 1u) << (src + 4 * i);
    }
    return bits;
}

__device__ __forceinline__ void epilogue_prefetch(const GemmArgs& g, const EpiCtx& e, int m_glob, int n_first,
                                                  ChunkPre& pre) {
    const bool tr = g.transpose_out != 0;
    const long long stride = tr ? g.ld_out : 1;
    const long long idx0 = tr ? static_cast<long long>(n_first) * g.ld_out + m_glob
                              : static_cast<long long>(m_glob) * g.ld_out + n_first;
    const int n_valid = g.N - n_first;
    pre.have = true;
    pre.keep = 0xFFFFu;
    if (e.row_ok) {
        if (g.mask_src) {
#pragma unroll
            for (int j = 0; j < 16; ++j) pre.mk[j] = (j < n_valid) ? __bfloat162float(g.mask_src[idx0 + j * stride]) : 0.f;
        }
        if (g.addend) {
#pragma unroll
            for (int j = 0; j < 16; ++j) pre.ad[j] = (j < n_valid) ? __bfloat162float(g.addend[idx0 + j * stride]) : 0.f;
        }
    }
    if (g.drop_p > 0.f) pre.keep = tr ? dropout_bits_transposed(g, e, idx0) : dropout_bits_generic(g, e, idx0, stride);
}

__device__ __forceinline__ void epilogue_chunk(const GemmArgs& g, const EpiCtx& e, float (&v)[16], int m_glob,
                                               int n_first, float& ss, const ChunkPre* pre = nullptr) {
    const bool tr = g.transpose_out != 0;
    const long long stride = tr ? g.ld_out : 1;
    const long long idx0 = tr ? static_cast<long long>(n_first) * g.ld_out + m_glob
                              : static_cast<long long>(m_glob) * g.ld_out + n_first;
    const int n_valid = g.N - n_first;                 // columns of this chunk inside the matrix (may exceed 16)

#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] *= g.alpha;
    if (g.bias) {
        if (g.bias_along_m) {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] += e.bias_m;
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] += (j < n_valid) ? g.bias[n_first + j] : 0.f;
        }
    }
    if (g.relu) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    if (g.drop_p > 0.f && pre) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = ((pre->keep >> j) & 1u) ? v[j] * e.inv_keep : 0.f;
    } else if (g.drop_p > 0.f) {
        if (tr) {
            // element index n * ld + m: lanes m..m+3 share a counter
            const uint32_t keep = dropout_bits_transposed(g, e, idx0);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * e.inv_keep : 0.f;
        } else {
            // 16 consecutive indices: 4 (aligned) or 5 counters
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint64_t idx = static_cast<uint64_t>(idx0 + j);
                const uint4 r = e.rng(idx >> 2, g.stream);
                const uint32_t sel = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
                v[j] = sel >= e.thr ? v[j] * e.inv_keep : 0.f;
            }
        }
    }
    if (!e.row_ok) return;
    if (g.mask_src) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float mk = pre ? pre->mk[j] : ((j < n_valid) ? __bfloat162float(g.mask_src[idx0 + j * stride]) : 0.f);
            v[j] = mk > 0.f ? v[j] * g.mask_scale : 0.f;
        }
    }
    if (g.addend) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (pre) v[j] += pre->ad[j];
            else if (j < n_valid) v[j] += __bfloat162float(g.addend[idx0 + j * stride]);
        }
    }

    if (g.out_f32) {
        float* dst = static_cast<float*>(g.out) + idx0;
        if (!tr && n_valid >= 16 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
            float4* d4 = reinterpret_cast<float4*>(dst);
            if (g.accumulate) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 o = d4[q];
                    v[4 * q] += o.x; v[4 * q + 1] += o.y; v[4 * q + 2] += o.z; v[4 * q + 3] += o.w;
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) d4[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                if (j < n_valid) {
                    if (g.accumulate) v[j] += dst[j * stride];
                    dst[j * stride] = v[j];
                }
            }
        }
    } else {
        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(g.out) + idx0;
        if (!tr && n_valid >= 16 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
            uint4 pk[2];
            __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
            for (int q = 0; q < 8; ++q) h2[q] = __floats2bfloat162_rn(v[2 * q], v[2 * q + 1]);
            reinterpret_cast<uint4*>(dst)[0] = pk[0];
            reinterpret_cast<uint4*>(dst)[1] = pk[1];
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
                if (j < n_valid) dst[j * stride] = __float2bfloat16(v[j]);
        }
    }
    if (g.sumsq) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (j < n_valid) ss += v[j] * v[j];
    }
}

}  // namespace pb2
