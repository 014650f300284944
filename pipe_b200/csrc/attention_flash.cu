// Flash attention for sm_100a on tcgen05 / TMEM / TMA: any sequence length, head_dim 64 / 128 / 256 (forward),
// causal flag, probability dropout, on the packed QKV layout of the encoder layer ([T, 3E], row = s * NB + n).
//
// Forward, one CTA per (128 query rows, head, sequence):
//   warp 0      TMA producer: Q once, then K / V blocks of BN keys through a 2-stage ring.  The tiles come
//               straight out of the packed tensor with a 3-D tensor map (column, n, s) and land 128-byte
//               swizzled; nothing is gathered or transposed by threads.
//   warp 1      one elected lane issues the MMAs: S = Q K^T (both operands K-major) into one of two S
//               accumulators in TMEM, and O += P V with P read from shared memory (K-major) and V used as it
//               lies in memory, [keys, D] with D contiguous, as an MN-major B operand — no V^T copy.
//   warps 2-5   online softmax, one thread per query row (the TMEM lane it can read): S from TMEM, running
//               max / sum in registers, dropout from the counter RNG, P to shared memory as bf16; when the
//               running max moved, the O accumulator is rescaled in place (tcgen05.ld / tcgen05.st).
// The two S accumulators let the tensor core compute S of block j + 1 while the softmax warps work on block j.
// Per row the kernel also writes LSE = log sum exp of the scaled scores, which is all the backward needs to
// recompute the probabilities.
//
// Backward: see the second half of this file.
//
// Reference counterpart: the attention core of nn.MultiheadAttention (main.py:148) for sequences up to the
// positional table's 5000 (main.py:59); SURVEY.md K3, W6.  The S <= 32 single-tile kernels of
// attention_tc.cu stay the path for BASELINE's 25-token sequences (a UMMA pipeline costs more than it saves
// there); this file is the path beyond.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kFaBM = 128;          // query rows per CTA = TMEM lanes
constexpr int kFaThreads = 192;     // producer warp, MMA warp, four softmax warps

// 3-D tile load: coordinates (column, n, s)
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int32_t c0,
                                            int32_t c1, int32_t c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

struct FaArgs {
    int NB, S, H, D;
    int causal;
    float scale_log2;        // 1/sqrt(D) * log2(e)
    float drop_p;
    unsigned long long seed, stream;
    __nv_bfloat16* out;      // [T, E]
    float* lse;              // [NB * H, S] natural log of the row sums of exp(scaled scores)
    long long ld_out;        // E
};

struct __align__(8) FaCtl {
    uint64_t q_full;
    uint64_t kv_full[2];
    uint64_t kv_empty[2];
    uint64_t s_full[2];
    uint64_t p_full;
    uint64_t pv_done;
    uint32_t tmem_base;
    uint32_t pad;
};

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Dropout decisions of eight consecutive keys of one attention row from ONE generator call: 16 random bits per
// element (p is quantised to 1/65536; the kernels scale by the exact keep probability of the quantised value).
// counter = row * ceil(S / 8) + key / 8, so forward, backward and recompute agree for any S.
__device__ __forceinline__ uint32_t drop_keep8(const CounterRng& rng, uint64_t stream, uint64_t counter, uint32_t thr16) {
    const uint4 r = rng(counter, stream);
    uint32_t bits = 0;
    bits |= ((r.x & 0xFFFFu) >= thr16) << 0;
    bits |= ((r.x >> 16) >= thr16) << 1;
    bits |= ((r.y & 0xFFFFu) >= thr16) << 2;
    bits |= ((r.y >> 16) >= thr16) << 3;
    bits |= ((r.z & 0xFFFFu) >= thr16) << 4;
    bits |= ((r.z >> 16) >= thr16) << 5;
    bits |= ((r.w & 0xFFFFu) >= thr16) << 6;
    bits |= ((r.w >> 16) >= thr16) << 7;
    return bits;
}
__host__ __device__ __forceinline__ uint32_t drop_thr16(float p) {
    const float t = p * 65536.f + 0.5f;
    return t < 0.f ? 0u : (t > 65535.f ? 65535u : static_cast<uint32_t>(t));
}

// element (row, col) of a [rows][64] bf16 tile with the 128-byte swizzle, as a byte offset
__device__ __forceinline__ uint32_t sw128_offset(uint32_t row, uint32_t col) {
    return row * 128u + ((((col >> 3) ^ (row & 7u)) << 4) | ((col & 7u) << 1));
}

template <int D, int BN>
__global__ void __launch_bounds__(kFaThreads, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_kv, const FaArgs a) {
    constexpr int KB = D / 64;                       // 64-wide blocks of the head dimension
    constexpr uint32_t kQBytes = kFaBM * D * 2;
    constexpr uint32_t kKBytes = BN * D * 2;
    constexpr uint32_t kPBytes = kFaBM * BN * 2;
    constexpr uint32_t kTmemCols = (2 * 128 + D) > 256 ? 512 : 256;
    constexpr uint32_t kOCol = 256;                  // O accumulator behind the two S accumulators
    static_assert(2 * 128 + D <= 512, "TMEM budget");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kQBytes;                      // [2][KB][BN][128 B]
    uint8_t* sV = sK + 2 * kKBytes;                  // [2][KB][BN][128 B]
    uint8_t* sP = sV + 2 * kKBytes;                  // [BN / 64][128][128 B]
    FaCtl* ctl = reinterpret_cast<FaCtl*>(sP + kPBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * kFaBM;
    const int h = blockIdx.y, n = blockIdx.z;
    const int E = a.H * D;
    int n_blocks = (a.S + BN - 1) / BN;
    if (a.causal) {
        const int last = (q0 + kFaBM - 1 < a.S - 1 ? q0 + kFaBM - 1 : a.S - 1);
        n_blocks = last / BN + 1;
    }

    pdl_launch_dependents();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv);
        tma_prefetch_desc(&tmap_kv);
        mbar_init(&ctl->q_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl->kv_full[s], 1);
            mbar_init(&ctl->kv_empty[s], 1);
            mbar_init(&ctl->s_full[s], 1);
        }
        mbar_init(&ctl->p_full, 128);
        mbar_init(&ctl->pv_done, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&ctl->tmem_base, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            pdl_wait();
            mbar_expect_tx(&ctl->q_full, kQBytes);
            for (int kb = 0; kb < KB; ++kb)
                tma_load_3d(sQ + kb * (kFaBM * 128), &tmap_qkv, &ctl->q_full, h * D + kb * 64, n, q0);
            for (int j = 0; j < n_blocks; ++j) {
                const int s = j & 1;
                mbar_wait(&ctl->kv_empty[s], ((j >> 1) & 1) ^ 1);
                mbar_expect_tx(&ctl->kv_full[s], 2 * kKBytes);
                for (int kb = 0; kb < KB; ++kb) {
                    tma_load_3d(sK + s * kKBytes + kb * (BN * 128), &tmap_kv, &ctl->kv_full[s], E + h * D + kb * 64, n, j * BN);
                    tma_load_3d(sV + s * kKBytes + kb * (BN * 128), &tmap_kv, &ctl->kv_full[s], 2 * E + h * D + kb * 64, n, j * BN);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t idesc_s = umma_idesc_bf16_major(kFaBM, BN, 0, 0);
        const uint32_t idesc_o = umma_idesc_bf16_major(kFaBM, D, 0, 1);
        auto issue_qk = [&](int j) {
            const int s = j & 1;
            mbar_wait(&ctl->kv_full[s], (j >> 1) & 1);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t d_tmem = tmem + (j & 1) * 128;
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t da = umma_desc_k_sw128(smem_u32(sQ + kb * (kFaBM * 128)));
                    const uint64_t db = umma_desc_k_sw128(smem_u32(sK + s * kKBytes + kb * (BN * 128)));
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc_s, (kb | k) ? 1u : 0u);
                }
                umma_commit(&ctl->s_full[j & 1]);
            }
            __syncwarp();
        };
        mbar_wait(&ctl->q_full, 0);
        issue_qk(0);
        for (int j = 0; j < n_blocks; ++j) {
            if (j + 1 < n_blocks) issue_qk(j + 1);          // S of the next block while softmax works on this one
            mbar_wait(&ctl->p_full, j & 1);
            tc_fence_after();
            if (elect_one()) {
                const int s = j & 1;
                const uint32_t v_base = smem_u32(sV + s * kKBytes);
#pragma unroll
                for (int t = 0; t < BN / 16; ++t) {
                    // A: P[128, BN] K-major, 16 keys = 32 bytes inside the 64-key block t / 4
                    const uint64_t da = umma_desc_k_sw128(smem_u32(sP + (t >> 2) * (kFaBM * 128))) + 2 * (t & 3);
                    // B: V[keys, D] MN-major, 16 keys = two 8-row groups = 2048 bytes; D atoms BN * 128 bytes apart
                    const uint64_t db = umma_desc_mn_sw128(v_base + t * 2048, BN * 128);
                    umma_bf16(tmem + kOCol, da, db, idesc_o, (j | t) ? 1u : 0u);
                }
                umma_commit(&ctl->kv_empty[s]);
                umma_commit(&ctl->pv_done);
            }
            __syncwarp();
        }
    } else {
        // ===================== softmax / correction / epilogue (warps 2..5) =====================
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;               // query row inside the tile = TMEM lane
        const int q = q0 + row;
        const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
        const CounterRng rng(a.seed);
        const uint32_t thr16 = drop_thr16(a.drop_p);
        const bool drop = a.drop_p > 0.f;
        const uint64_t row_base = (static_cast<uint64_t>(n * a.H + h) * a.S + (q < a.S ? q : 0)) * static_cast<uint64_t>((a.S + 7) >> 3);
        float m_run = -INFINITY, l_run = 0.f;
        for (int j = 0; j < n_blocks; ++j) {
            mbar_wait(&ctl->s_full[j & 1], (j >> 1) & 1);
            tc_fence_after();
            const uint32_t s_addr = tmem + lane_addr + (j & 1) * 128;
            // Blocks that lie entirely inside the sequence (and, with a causal mask, entirely below the diagonal)
            // need no per-element tests: most of a long sequence takes the plain path.
            const bool need_mask = (j * BN + BN > a.S) || (a.causal && j * BN + BN - 1 > q0);
            // pass 1: row maximum of this block (raw scores; the scale is positive, so it commutes with max)
            float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(s_addr + c, r);
                tmem_ld_wait();
                if (need_mask) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int key = j * BN + c + i;
                        const bool ok = key < a.S && (!a.causal || key <= q);
                        mx[i & 3] = fmaxf(mx[i & 3], ok ? __uint_as_float(r[i]) : -INFINITY);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(r[i]));
                }
            }
            const float m_blk = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * a.scale_log2;
            const float m_new = fmaxf(m_run, m_blk);
            const float m_use = m_new == -INFINITY ? 0.f : m_new;
            const float alpha = fast_exp2(m_run - m_use);          // 0 on the first block (m_run = -inf)
            // the previous block's P V must be complete before P is overwritten and O rescaled
            if (j > 0) {
                mbar_wait(&ctl->pv_done, (j - 1) & 1);
                tc_fence_after();
                if (__any_sync(0xffffffffu, m_new != m_run)) {
                    const uint32_t o_addr = tmem + lane_addr + kOCol;
#pragma unroll
                    for (int c = 0; c < D; c += 16) {
                        uint32_t r[16];
                        tmem_ld_32x16(o_addr + c, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tmem_st_32x16(o_addr + c, r);
                    }
                    tmem_st_wait();
                }
            }
            // pass 2: probabilities -> bf16 P tile (dropout applied), row sum before dropout
            float ls[4] = {0.f, 0.f, 0.f, 0.f};
            const float neg_m = -m_use;
#pragma unroll
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(s_addr + c, r);
                tmem_ld_wait();
                float p[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) p[i] = fast_exp2(fmaf(__uint_as_float(r[i]), a.scale_log2, neg_m));
                if (need_mask) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int key = j * BN + c + i;
                        if (!(key < a.S && (!a.causal || key <= q))) p[i] = 0.f;
                    }
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) ls[i & 3] += p[i];
                if (drop) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                        const uint32_t keep = drop_keep8(rng, a.stream, row_base + static_cast<uint64_t>((j * BN + c + i) >> 3), thr16);
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            if (!((keep >> e) & 1u)) p[i + e] = 0.f;
                    }
                }
                uint32_t packed[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(p[i], p[i + 1]);
                    packed[i >> 1] = *reinterpret_cast<uint32_t*>(&h2);
                }
                // 32 keys = four 16-byte chunks of this row in the 64-key block c / 64
                uint8_t* tile = sP + (c >> 6) * (kFaBM * 128);
                const uint32_t col = c & 63;
#pragma unroll
                for (int v4 = 0; v4 < 4; ++v4)
                    *reinterpret_cast<uint4*>(tile + sw128_offset(row, col + 8 * v4)) =
                        make_uint4(packed[4 * v4], packed[4 * v4 + 1], packed[4 * v4 + 2], packed[4 * v4 + 3]);
            }
            const float l_blk = (ls[0] + ls[1]) + (ls[2] + ls[3]);
            l_run = l_run * alpha + l_blk;
            m_run = m_new;
            fence_proxy_async();            // P was written by the generic proxy, the tensor core reads it through the async proxy
            tc_fence_before();
            mbar_arrive(&ctl->p_full);
        }
        // epilogue: O / l (and the dropout rescale), bf16, straight to [T, E]
        mbar_wait(&ctl->pv_done, (n_blocks - 1) & 1);
        tc_fence_after();
        const float inv = (l_run > 0.f ? 1.f / l_run : 0.f) * (drop ? 65536.f / (65536.f - static_cast<float>(thr16)) : 1.f);
        const uint32_t o_addr = tmem + lane_addr + kOCol;
        __nv_bfloat16* dst = a.out + (static_cast<size_t>(q) * a.NB + n) * a.ld_out + h * D;
#pragma unroll
        for (int c = 0; c < D; c += 16) {
            uint32_t r[16];
            tmem_ld_32x16(o_addr + c, r);
            tmem_ld_wait();
            if (q < a.S) {
                uint4 pk[2];
                __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    h2[i] = __floats2bfloat162_rn(__uint_as_float(r[2 * i]) * inv, __uint_as_float(r[2 * i + 1]) * inv);
                reinterpret_cast<uint4*>(dst + c)[0] = pk[0];
                reinterpret_cast<uint4*>(dst + c)[1] = pk[1];
            }
        }
        if (q < a.S && a.lse)
            a.lse[static_cast<size_t>(n * a.H + h) * a.S + q] = (m_run == -INFINITY ? 0.f : m_run) * 0.6931471805599453f + logf(l_run);
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, kTmemCols);
    }
}


// =====================================================================================================
// Backward.  One CTA per (128 keys, head, sequence) walks the query blocks that attend to them:
//   S  = Q K^T, dP = dO V^T                       (two MMAs into TMEM, both operands K-major)
//   P  = exp(S * scale - LSE), dropout, dS = P o (dP - delta) * scale       (softmax warps, one row each,
//                                                   P and dS to shared memory as bf16)
//   dV += P^T dO, dK += dS^T Q                    (A and B both MN-major: the tiles are used as they were
//                                                   loaded / written, nothing is transposed)
//   dQ  = dS K                                     (A K-major, B MN-major) -> fp32 atomics into a dQ accumulator
// dK and dV stay in TMEM for the whole walk and are written once.  delta = rowsum(dO o O) comes from a small
// kernel in front, the bf16 dQ from one behind (the accumulator is fp32 because several key blocks add to it).
// TMEM: S | dP | dK | dV = 256 + 2 D columns, dQ reuses the S columns once S has been read: D <= 128.
// =====================================================================================================
struct FaBwdArgs {
    int NB, S, H, D;
    int causal;
    float scale, scale_log2;
    float drop_p;
    unsigned long long seed, stream;
    const float* lse;        // [NB * H, S]
    const float* delta;      // [NB * H, S]
    float* dq_acc;           // [T, E] fp32, zeroed
    __nv_bfloat16* dqkv;     // [T, 3E]
    long long ld_dqkv;       // 3E
};

struct __align__(8) FaBwdCtl {
    uint64_t kv_full;
    uint64_t q_full[2];
    uint64_t q_empty[2];
    uint64_t sdp_full;
    uint64_t pds_full;
    uint64_t dq_full;
    uint64_t dq_free;
    uint32_t tmem_base;
    uint32_t pad;
};

template <int D, int STAGES>
__global__ void __launch_bounds__(kFaThreads, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do, const FaBwdArgs a) {
    constexpr int KB = D / 64;
    constexpr int BN = 128;                              // keys per CTA
    constexpr uint32_t kTile = 128 * D * 2;              // one [128, D] bf16 tile
    constexpr uint32_t kPTile = 128 * 128 * 2;
    constexpr uint32_t kDkCol = 256, kDvCol = 256 + D;
    static_assert(256 + 2 * D <= 512, "TMEM budget");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + kTile;
    uint8_t* sQ = sV + kTile;                            // [STAGES]
    uint8_t* sdO = sQ + STAGES * kTile;                  // [STAGES]
    uint8_t* sP = sdO + STAGES * kTile;                  // [2 key blocks][128 q rows][128 B]
    uint8_t* sdS = sP + kPTile;
    FaBwdCtl* ctl = reinterpret_cast<FaBwdCtl*>(sdS + kPTile);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k0 = blockIdx.x * BN;
    const int h = blockIdx.y, n = blockIdx.z;
    const int E = a.H * D;
    const int q_blocks = (a.S + kFaBM - 1) / kFaBM;
    const int first = a.causal ? k0 / kFaBM : 0;         // query blocks before the key block see none of its keys
    const int n_it = q_blocks - first;

    pdl_launch_dependents();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv);
        tma_prefetch_desc(&tmap_do);
        mbar_init(&ctl->kv_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl->q_full[s], 1);
            mbar_init(&ctl->q_empty[s], 1);
        }
        mbar_init(&ctl->sdp_full, 1);
        mbar_init(&ctl->pds_full, 128);
        mbar_init(&ctl->dq_full, 1);
        mbar_init(&ctl->dq_free, 128);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&ctl->tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;

    if (warp == 0) {
        if (elect_one()) {
            pdl_wait();
            mbar_expect_tx(&ctl->kv_full, 2 * kTile);
            for (int kb = 0; kb < KB; ++kb) {
                tma_load_3d(sK + kb * (BN * 128), &tmap_qkv, &ctl->kv_full, E + h * D + kb * 64, n, k0);
                tma_load_3d(sV + kb * (BN * 128), &tmap_qkv, &ctl->kv_full, 2 * E + h * D + kb * 64, n, k0);
            }
            for (int it = 0; it < n_it; ++it) {
                const int s = it % STAGES;
                const int q0 = (first + it) * kFaBM;
                mbar_wait(&ctl->q_empty[s], ((it / STAGES) & 1) ^ 1);
                mbar_expect_tx(&ctl->q_full[s], 2 * kTile);
                for (int kb = 0; kb < KB; ++kb) {
                    tma_load_3d(sQ + s * kTile + kb * (kFaBM * 128), &tmap_qkv, &ctl->q_full[s], h * D + kb * 64, n, q0);
                    tma_load_3d(sdO + s * kTile + kb * (kFaBM * 128), &tmap_do, &ctl->q_full[s], h * D + kb * 64, n, q0);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        const uint32_t idesc_s = umma_idesc_bf16_major(128, 128, 0, 0);      // S, dP: K-major x K-major
        const uint32_t idesc_kv = umma_idesc_bf16_major(128, D, 1, 1);       // dV, dK: MN-major x MN-major
        const uint32_t idesc_q = umma_idesc_bf16_major(128, D, 0, 1);        // dQ: K-major x MN-major
        mbar_wait(&ctl->kv_full, 0);
        for (int it = 0; it < n_it; ++it) {
            const int s = it % STAGES;
            mbar_wait(&ctl->q_full[s], (it / STAGES) & 1);
            if (it > 0) mbar_wait(&ctl->dq_free, (it - 1) & 1);       // dQ of the previous block was read out of the S columns
            tc_fence_after();
            const uint32_t q_base = smem_u32(sQ + s * kTile), do_base = smem_u32(sdO + s * kTile);
            if (elect_one()) {
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t dq_ = umma_desc_k_sw128(q_base + kb * (kFaBM * 128));
                    const uint64_t dk_ = umma_desc_k_sw128(smem_u32(sK + kb * (BN * 128)));
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem, dq_ + 2 * k, dk_ + 2 * k, idesc_s, (kb | k) ? 1u : 0u);
                }
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t ddo = umma_desc_k_sw128(do_base + kb * (kFaBM * 128));
                    const uint64_t dv_ = umma_desc_k_sw128(smem_u32(sV + kb * (BN * 128)));
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem + 128, ddo + 2 * k, dv_ + 2 * k, idesc_s, (kb | k) ? 1u : 0u);
                }
                umma_commit(&ctl->sdp_full);
            }
            __syncwarp();
            mbar_wait(&ctl->pds_full, it & 1);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t p_base = smem_u32(sP), ds_base = smem_u32(sdS);
#pragma unroll
                for (int t = 0; t < 8; ++t) {            // 16 query rows per MMA: K = q
                    // A = P^T / dS^T: [q rows][keys] tiles read MN-major (keys contiguous), key atoms 16 KB apart
                    const uint64_t ap = umma_desc_mn_sw128(p_base + t * 2048, kFaBM * 128);
                    const uint64_t ads = umma_desc_mn_sw128(ds_base + t * 2048, kFaBM * 128);
                    // B = dO / Q: [q rows][D] tiles read MN-major (d contiguous), d atoms 16 KB apart
                    const uint64_t bdo = umma_desc_mn_sw128(do_base + t * 2048, kFaBM * 128);
                    const uint64_t bq = umma_desc_mn_sw128(q_base + t * 2048, kFaBM * 128);
                    umma_bf16(tmem + kDvCol, ap, bdo, idesc_kv, (it | t) ? 1u : 0u);
                    umma_bf16(tmem + kDkCol, ads, bq, idesc_kv, (it | t) ? 1u : 0u);
                }
#pragma unroll
                for (int t = 0; t < 8; ++t) {            // 16 keys per MMA: K = keys
                    const uint64_t ads = umma_desc_k_sw128(ds_base + (t >> 2) * (kFaBM * 128)) + 2 * (t & 3);
                    const uint64_t bk = umma_desc_mn_sw128(smem_u32(sK) + t * 2048, BN * 128);
                    umma_bf16(tmem, ads, bk, idesc_q, t ? 1u : 0u);
                }
                umma_commit(&ctl->q_empty[s]);
                umma_commit(&ctl->dq_full);
            }
            __syncwarp();
        }
    } else {
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
        const CounterRng rng(a.seed);
        const uint32_t thr16 = drop_thr16(a.drop_p);
        const bool drop = a.drop_p > 0.f;
        const float inv_keep = drop ? 65536.f / (65536.f - static_cast<float>(thr16)) : 1.f;
        for (int it = 0; it < n_it; ++it) {
            const int q = (first + it) * kFaBM + row;
            const bool q_ok = q < a.S;
            const size_t stat = static_cast<size_t>(n * a.H + h) * a.S + (q_ok ? q : 0);
            const float lse2 = q_ok ? a.lse[stat] * 1.4426950408889634f : 0.f;
            const float delta = q_ok ? a.delta[stat] : 0.f;
            const uint64_t row_base = (static_cast<uint64_t>(n * a.H + h) * a.S + (q_ok ? q : 0)) * static_cast<uint64_t>((a.S + 7) >> 3);
            mbar_wait(&ctl->sdp_full, it & 1);
            tc_fence_after();
            if (it > 0) {
                // P / dS of the previous block are still operands of its dV / dK / dQ MMAs until dq_full
                // (waited below in the previous iteration), so they may be overwritten now
            }
            const int q_first = (first + it) * kFaBM;
            const bool need_mask = (q_first + kFaBM > a.S) || (k0 + BN > a.S) || (a.causal && k0 + BN - 1 > q_first);
            const float neg_lse = -lse2;
            const float ds_scale = a.scale;
#pragma unroll
            for (int c = 0; c < BN; c += 16) {
                uint32_t rs[16], rp[16];
                tmem_ld_32x16(tmem + lane_addr + c, rs);
                tmem_ld_32x16(tmem + lane_addr + 128 + c, rp);
                tmem_ld_wait();
                float p[16], dp[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    p[i] = fast_exp2(fmaf(__uint_as_float(rs[i]), a.scale_log2, neg_lse));
                    dp[i] = __uint_as_float(rp[i]);
                }
                if (need_mask) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int key = k0 + c + i;
                        if (!(q_ok && key < a.S && (!a.causal || key <= q))) p[i] = 0.f;
                    }
                }
                float pd[16];                                    // what multiplied V in the forward
                if (drop) {
#pragma unroll
                    for (int i = 0; i < 16; i += 8) {
                        const uint32_t keep = drop_keep8(rng, a.stream, row_base + static_cast<uint64_t>((k0 + c + i) >> 3), thr16);
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const bool k1 = (keep >> e) & 1u;
                            pd[i + e] = k1 ? p[i + e] * inv_keep : 0.f;
                            dp[i + e] = k1 ? dp[i + e] * inv_keep : 0.f;       // back through the dropout
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) pd[i] = p[i];
                }
                uint32_t pk_p[8], pk_ds[8];
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    __nv_bfloat162 hp = __floats2bfloat162_rn(pd[i], pd[i + 1]);
                    __nv_bfloat162 hd = __floats2bfloat162_rn(p[i] * (dp[i] - delta) * ds_scale, p[i + 1] * (dp[i + 1] - delta) * ds_scale);
                    pk_p[i >> 1] = *reinterpret_cast<uint32_t*>(&hp);
                    pk_ds[i >> 1] = *reinterpret_cast<uint32_t*>(&hd);
                }
                const uint32_t tile_off = (c >> 6) * (kFaBM * 128);
                const uint32_t col = c & 63;
                *reinterpret_cast<uint4*>(sP + tile_off + sw128_offset(row, col)) = make_uint4(pk_p[0], pk_p[1], pk_p[2], pk_p[3]);
                *reinterpret_cast<uint4*>(sP + tile_off + sw128_offset(row, col + 8)) = make_uint4(pk_p[4], pk_p[5], pk_p[6], pk_p[7]);
                *reinterpret_cast<uint4*>(sdS + tile_off + sw128_offset(row, col)) = make_uint4(pk_ds[0], pk_ds[1], pk_ds[2], pk_ds[3]);
                *reinterpret_cast<uint4*>(sdS + tile_off + sw128_offset(row, col + 8)) = make_uint4(pk_ds[4], pk_ds[5], pk_ds[6], pk_ds[7]);
            }
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&ctl->pds_full);
            // dQ of this query block: out of TMEM, into the fp32 accumulator other key blocks add to as well
            mbar_wait(&ctl->dq_full, it & 1);
            tc_fence_after();
            float* dq = a.dq_acc + (static_cast<size_t>(q_ok ? q : 0) * a.NB + n) * E + h * D;
#pragma unroll
            for (int c = 0; c < D; c += 16) {
                uint32_t r[16];
                tmem_ld_32x16(tmem + lane_addr + c, r);
                tmem_ld_wait();
                if (q_ok) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        red_add_f32x4(dq + c + i, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), __uint_as_float(r[i + 2]),
                                      __uint_as_float(r[i + 3]));
                }
            }
            tc_fence_before();
            mbar_arrive(&ctl->dq_free);
        }
        // dK, dV of this key block: one writer, straight into the packed gradient
        const int key = k0 + row;
        if (n_it > 0) {
            __nv_bfloat16* dk = a.dqkv + (static_cast<size_t>(key < a.S ? key : 0) * a.NB + n) * a.ld_dqkv + E + h * D;
            __nv_bfloat16* dv = dk + E;
#pragma unroll
            for (int c = 0; c < D; c += 16) {
                uint32_t rk[16], rv[16];
                tmem_ld_32x16(tmem + lane_addr + kDkCol + c, rk);
                tmem_ld_32x16(tmem + lane_addr + kDvCol + c, rv);
                tmem_ld_wait();
                if (key < a.S) {
                    uint4 pk[2], pv[2];
                    __nv_bfloat162* hk = reinterpret_cast<__nv_bfloat162*>(pk);
                    __nv_bfloat162* hv = reinterpret_cast<__nv_bfloat162*>(pv);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        hk[i] = __floats2bfloat162_rn(__uint_as_float(rk[2 * i]), __uint_as_float(rk[2 * i + 1]));
                        hv[i] = __floats2bfloat162_rn(__uint_as_float(rv[2 * i]), __uint_as_float(rv[2 * i + 1]));
                    }
                    reinterpret_cast<uint4*>(dk + c)[0] = pk[0];
                    reinterpret_cast<uint4*>(dk + c)[1] = pk[1];
                    reinterpret_cast<uint4*>(dv + c)[0] = pv[0];
                    reinterpret_cast<uint4*>(dv + c)[1] = pv[1];
                }
            }
        }
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
}

// delta[(n, h), s] = sum_d dO[s, n, h, d] * O[s, n, h, d]: one warp per (token, head)
__global__ void flash_delta_kernel(const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                                   float* __restrict__ delta, int NB, int S, int H, int D) {
    pdl_launch_dependents();
    pdl_wait();
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int total = NB * S * H;
    if (w >= total) return;
    const int h = w % H, tok = w / H;                      // tok = s * NB + n
    const int s = tok / NB, n = tok % NB;
    const size_t base = static_cast<size_t>(tok) * H * D + static_cast<size_t>(h) * D;
    float acc = 0.f;
    for (int d = lane * 2; d < D; d += 64) {
        const __nv_bfloat162 o = *reinterpret_cast<const __nv_bfloat162*>(out + base + d);
        const __nv_bfloat162 g = *reinterpret_cast<const __nv_bfloat162*>(dout + base + d);
        acc += __bfloat162float(o.x) * __bfloat162float(g.x) + __bfloat162float(o.y) * __bfloat162float(g.y);
    }
    acc = warp_sum(acc);
    if (lane == 0) delta[static_cast<size_t>(n * H + h) * S + s] = acc;
}

// dqkv[:, 0:E] = bf16(dq_acc)
__global__ void flash_dq_cast_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, size_t T, int E) {
    pdl_launch_dependents();
    pdl_wait();
    const size_t i4 = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (i4 >= T * E) return;
    const size_t t = i4 / E, c = i4 % E;
    const float4 v = *reinterpret_cast<const float4*>(acc + i4);
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(dqkv + t * 3 * E + c);
    o[0] = __floats2bfloat162_rn(v.x, v.y);
    o[1] = __floats2bfloat162_rn(v.z, v.w);
}

}  // namespace pb2

using namespace pb2;

namespace {

// [T, cols] bf16 with row = s * NB + n, seen as (column, n, s); box = 64 columns x 1 x box_s rows, 128-byte swizzle
int make_tmap_seq3d(CUtensorMap* out, const void* ptr, int cols, int NB, int S, int box_s) {
    struct Key { const void* p; int cols, NB, S, box; };
    struct Slot { Key k; CUtensorMap m; bool valid; };
    thread_local Slot cache[64];
    const size_t slot_i = (reinterpret_cast<uintptr_t>(ptr) >> 8 ^ static_cast<size_t>(S) * 131 ^ static_cast<size_t>(box_s)) % 64;
    Slot& sl = cache[slot_i];
    if (sl.valid && sl.k.p == ptr && sl.k.cols == cols && sl.k.NB == NB && sl.k.S == S && sl.k.box == box_s) {
        *out = sl.m;
        return 0;
    }
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return -1;
    cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(NB), static_cast<cuuint64_t>(S)};
    cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols) * 2, static_cast<cuuint64_t>(cols) * NB * 2};
    cuuint32_t box[3] = {64, 1, static_cast<cuuint32_t>(box_s)};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (3-D) failed (%d) for cols %d NB %d S %d box %d", static_cast<int>(r), cols, NB, S, box_s);
        return -1;
    }
    sl.k = Key{ptr, cols, NB, S, box_s};
    sl.m = *out;
    sl.valid = true;
    return 0;
}

template <int D, int BN>
int launch_fwd(const void* qkv, const FaArgs& a, cudaStream_t stream) {
    CUtensorMap tq, tkv;
    if (make_tmap_seq3d(&tq, qkv, 3 * a.H * D, a.NB, a.S, kFaBM)) return -1;
    if (make_tmap_seq3d(&tkv, qkv, 3 * a.H * D, a.NB, a.S, BN)) return -1;
    const size_t smem = static_cast<size_t>(kFaBM) * D * 2 + 4 * static_cast<size_t>(BN) * D * 2 + static_cast<size_t>(kFaBM) * BN * 2 +
                        sizeof(FaCtl) + 1024;
    if (ensure_max_smem(reinterpret_cast<const void*>(flash_fwd_kernel<D, BN>), 227 * 1024)) return -1;
    dim3 grid((a.S + kFaBM - 1) / kFaBM, a.H, a.NB);
    PB2_CUDA(launch_pdl(flash_fwd_kernel<D, BN>, grid, dim3(kFaThreads), smem, stream, tq, tkv, a));
    PB2_LAUNCHED("flash_fwd_kernel");
    return 0;
}

}  // namespace

extern "C" int pb2_attention_uses_flash(int S, int D) {
    // Measured on B200 (profiles/attention_bench_r2.txt): at S = 128, D = 64 the flash pair takes 10 + 28 us against
    // 28 + 84 us for the single-tile kernels; below one 64-key block the single-tile kernels win (one small tile per
    // head, no TMEM pipeline to fill).  PB2_ATTN_FLASH=0 keeps the single-tile kernels wherever they apply.
    static int off = -1;
    if (off < 0) {
        const char* e = getenv("PB2_ATTN_FLASH");
        off = (e && e[0] == '0') ? 1 : 0;
    }
    if (D != 64 && D != 128) return 0;
    if (S > 128) return 1;
    return (S >= 64 && !off) ? 1 : 0;
}

extern "C" int pb2_flash_attention_fwd(const void* qkv, void* out, float* lse, int NB, int S, int H, int D, int causal,
                                       float drop_p, unsigned long long seed, unsigned long long rng_stream,
                                       cudaStream_t stream) {
    if (NB <= 0 || S <= 0 || H <= 0) { set_error("pb2_flash_attention_fwd: empty problem"); return -1; }
    if (D != 64 && D != 128 && D != 256) { set_error("pb2_flash_attention_fwd: head_dim %d not in {64, 128, 256}", D); return -1; }
    if (reinterpret_cast<uintptr_t>(qkv) & 15) { set_error("pb2_flash_attention_fwd: qkv must be 16-byte aligned"); return -1; }
    FaArgs a;
    memset(&a, 0, sizeof(a));
    a.NB = NB; a.S = S; a.H = H; a.D = D; a.causal = causal;
    a.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    a.drop_p = drop_p; a.seed = seed; a.stream = rng_stream;
    a.out = static_cast<__nv_bfloat16*>(out); a.lse = lse; a.ld_out = static_cast<long long>(H) * D;
    if (D == 64) return launch_fwd<64, 128>(qkv, a, stream);
    if (D == 128) return launch_fwd<128, 128>(qkv, a, stream);
    return launch_fwd<256, 64>(qkv, a, stream);
}

namespace {

template <int D, int STAGES>
int launch_bwd(const void* qkv, const void* dout, const FaBwdArgs& a, cudaStream_t stream) {
    CUtensorMap tq, tdo;
    if (make_tmap_seq3d(&tq, qkv, 3 * a.H * D, a.NB, a.S, 128)) return -1;
    if (make_tmap_seq3d(&tdo, dout, a.H * D, a.NB, a.S, 128)) return -1;
    const size_t tile = static_cast<size_t>(128) * D * 2;
    const size_t smem = 2 * tile + 2 * STAGES * tile + 2 * (128 * 128 * 2) + sizeof(FaBwdCtl) + 1024;
    if (ensure_max_smem(reinterpret_cast<const void*>(flash_bwd_kernel<D, STAGES>), 227 * 1024)) return -1;
    dim3 grid((a.S + 127) / 128, a.H, a.NB);
    PB2_CUDA(launch_pdl(flash_bwd_kernel<D, STAGES>, grid, dim3(kFaThreads), smem, stream, tq, tdo, a));
    PB2_LAUNCHED("flash_bwd_kernel");
    return 0;
}

}  // namespace

extern "C" size_t pb2_flash_attention_bwd_scratch_bytes(int NB, int S, int H, int D) {
    const size_t T = static_cast<size_t>(NB) * S, E = static_cast<size_t>(H) * D;
    return ((static_cast<size_t>(NB) * H * S * 4 + 255) / 256) * 256 + T * E * 4;
}

extern "C" int pb2_flash_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                       void* scratch, size_t scratch_bytes, int NB, int S, int H, int D, int causal,
                                       float drop_p, unsigned long long seed, unsigned long long rng_stream,
                                       cudaStream_t stream) {
    if (NB <= 0 || S <= 0 || H <= 0) { set_error("pb2_flash_attention_bwd: empty problem"); return -1; }
    if (D != 64 && D != 128) { set_error("pb2_flash_attention_bwd: head_dim %d not in {64, 128}", D); return -1; }
    const size_t need = pb2_flash_attention_bwd_scratch_bytes(NB, S, H, D);
    if (!scratch || scratch_bytes < need) { set_error("pb2_flash_attention_bwd: scratch of %zu bytes needed, %zu given", need, scratch_bytes); return -1; }
    if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(scratch)) & 15) {
        set_error("pb2_flash_attention_bwd: qkv, dout and scratch must be 16-byte aligned");
        return -1;
    }
    const size_t T = static_cast<size_t>(NB) * S;
    const int E = H * D;
    float* delta = static_cast<float*>(scratch);
    float* dq_acc = reinterpret_cast<float*>(static_cast<char*>(scratch) + ((static_cast<size_t>(NB) * H * S * 4 + 255) / 256) * 256);
    PB2_CUDA(cudaMemsetAsync(dq_acc, 0, T * E * 4, stream));
    {
        const int warps = NB * S * H;
        const int blocks = (warps * 32 + 255) / 256;
        PB2_CUDA(launch_pdl(flash_delta_kernel, dim3(blocks), dim3(256), 0, stream, static_cast<const __nv_bfloat16*>(out),
                            static_cast<const __nv_bfloat16*>(dout), delta, NB, S, H, D));
        PB2_LAUNCHED("flash_delta_kernel");
    }
    FaBwdArgs a;
    memset(&a, 0, sizeof(a));
    a.NB = NB; a.S = S; a.H = H; a.D = D; a.causal = causal;
    a.scale = 1.f / sqrtf(static_cast<float>(D));
    a.scale_log2 = a.scale * 1.4426950408889634f;
    a.drop_p = drop_p; a.seed = seed; a.stream = rng_stream;
    a.lse = lse; a.delta = delta; a.dq_acc = dq_acc;
    a.dqkv = static_cast<__nv_bfloat16*>(dqkv); a.ld_dqkv = 3ll * E;
    const int rc = D == 64 ? launch_bwd<64, 2>(qkv, dout, a, stream) : launch_bwd<128, 1>(qkv, dout, a, stream);
    if (rc) return rc;
    {
        const size_t n4 = T * E / 4;
        const unsigned blocks = static_cast<unsigned>((n4 + 255) / 256);
        PB2_CUDA(launch_pdl(flash_dq_cast_kernel, dim3(blocks), dim3(256), 0, stream, static_cast<const float*>(dq_acc),
                            static_cast<__nv_bfloat16*>(dqkv), T, E));
        PB2_LAUNCHED("flash_dq_cast_kernel");
    }
    return 0;
}
