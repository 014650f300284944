// tcgen05 / TMEM / TMA GEMM for sm_100a:  D[M,N] = A[M,K] * B[N,K]^T  (+ fused epilogue)
//
// Both operands are bf16, K-major (row-major with K contiguous) — the layout of
// activations [tokens, features] and of nn.Linear weights [out, in] — so a
// Linear forward, its dgrad (against a pre-transposed weight copy) and its
// wgrad (against transposed activations) are all this one kernel.
//
// Shape of the problem in this framework: micro-batches carry
// 25-128 tokens against 4096-8192-wide weights, so the GEMMs are weight-streaming
// (HBM) bound.  The kernel is therefore built "swap-AB" first: the *weight* rows
// ride on the 128-wide UMMA M axis, the tokens on the UMMA N axis (any multiple
// of 16 up to 256, so 75 tokens cost an N=80 instruction instead of a padded
// 128-row tile), K is split across CTAs (split-K) until the grid covers the 148
// SMs, and a deep TMA ring (up to 8 x 16-48 KB stages) keeps enough bytes in
// flight per SM to saturate HBM3e.  The epilogue can store the tile transposed,
// which turns D^T[features, tokens] back into row-major [tokens, features].
//
// Pipeline inside a CTA (320 threads, one output tile, one K-split):
//   warp 0      TMA producer: cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier complete_tx
//   warp 1      allocates TMEM; one elected lane issues tcgen05.mma (M=128, N=block_n, K=16) x 4 per stage,
//               tcgen05.commit releases the smem stage / signals the accumulator
//   warps 2-9   epilogue: tcgen05.ld accumulator -> registers -> (split-K fix-up) -> bias / ReLU /
//               counter-RNG dropout / mask-multiply / residual add / accumulate -> global
//
// Split-K fix-up: every split writes its fp32 partial tile to a scratch workspace,
// bumps a per-tile counter, and the CTA that arrives last adds the other splits'
// partials to its own accumulator (still in TMEM) and runs the epilogue.  The
// counter is reset by that CTA, so the workspace is reusable by the next launch
// on the same stream without a memset.
//
// No counterpart in the reference: its GEMMs are cuBLAS calls behind nn.Linear /
// nn.MultiheadAttention (main.py:45,148); SURVEY.md K2/K4/K6/K7, W5.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <thread>

#include "gemm_epilogue.cuh"
#include "pipe_b200.h"

namespace pb2 {

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define PB2_TS(slot)                                                                                   \
    do {                                                                                               \
        if (g.debug_ts)                                                                                \
            g.debug_ts[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 16 + (slot)] = gtime(); \
    } while (0)

struct __align__(8) SmemCtl {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t tmem_full;
    uint64_t recv;          // cluster split-K: the peers' partials of this CTA's chunks have landed (transaction bytes)
    uint32_t tmem_base;
    uint32_t is_last;
};


// Row-major store of one 16-column chunk through a per-warp shared-memory transpose.
// In the TMEM register layout a thread owns one accumulator ROW, so a direct store makes every
// warp-wide instruction touch 32 rows x 16 bytes (half-filled sectors, 32 different lines).  Staging
// the 32 x 16 block in shared memory lets each instruction write 8 rows x 64 contiguous bytes instead
// (whole sectors).  Used for the plain / accumulate / alpha / bias / sumsq epilogue of large
// row-major outputs — the weight-gradient GEMMs, whose 64 MB fp32 results are their main cost.
__device__ __forceinline__ void epilogue_chunk_staged(const GemmArgs& g, float* st, float (&v)[16], int m_tile0,
                                                      int quarter, int lane, int n_first, float bias_m, float& ss) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = v[j] * g.alpha + bias_m;
#pragma unroll
    for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<float4*>(st + lane * 20 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    __syncwarp();
    const int r8 = lane >> 2, c4 = (lane & 3) * 4;
    const int n = n_first + c4;
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
        const int row = rr * 8 + r8;
        const int m = m_tile0 + quarter * 32 + row;
        float4 o = *reinterpret_cast<const float4*>(st + row * 20 + c4);
        if (g.bias && !g.bias_along_m) {
            if (n < g.N) o.x += g.bias[n];
            if (n + 1 < g.N) o.y += g.bias[n + 1];
            if (n + 2 < g.N) o.z += g.bias[n + 2];
            if (n + 3 < g.N) o.w += g.bias[n + 3];
        }
        if (m >= g.M || n >= g.N) continue;
        const long long idx = static_cast<long long>(m) * g.ld_out + n;
        if (g.out_f32) {
            float* dst = static_cast<float*>(g.out) + idx;
            if (n + 3 < g.N && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                if (g.accumulate) {
                    const float4 old = *reinterpret_cast<const float4*>(dst);
                    o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                *reinterpret_cast<float4*>(dst) = o;
                ss += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
            } else {
                float e[4] = {o.x, o.y, o.z, o.w};
                for (int j = 0; j < 4 && n + j < g.N; ++j) {
                    if (g.accumulate) e[j] += dst[j];
                    dst[j] = e[j];
                    ss += e[j] * e[j];
                }
            }
        } else {
            __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(g.out) + idx;
            if (n + 3 < g.N && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0)) {
                uint2 pk;
                *reinterpret_cast<__nv_bfloat162*>(&pk.x) = __floats2bfloat162_rn(o.x, o.y);
                *reinterpret_cast<__nv_bfloat162*>(&pk.y) = __floats2bfloat162_rn(o.z, o.w);
                *reinterpret_cast<uint2*>(dst) = pk;
            } else {
                const float e[4] = {o.x, o.y, o.z, o.w};
                for (int j = 0; j < 4 && n + j < g.N; ++j) dst[j] = __float2bfloat16(e[j]);
            }
            ss += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const GemmArgs g) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages][A 16 KB][B block_n*128 B] then control block
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t a_bytes = kBlockM * kBlockK * 2;
    const uint32_t b_bytes = g.block_n * kBlockK * 2;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem + g.ring_bytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBlockM;
    const int n0 = blockIdx.y * g.block_n;
    const int split = blockIdx.z;
    const int kb_begin = split * g.k_blocks_per_split;
    int kb_count = g.k_blocks_total - kb_begin;
    if (kb_count > g.k_blocks_per_split) kb_count = g.k_blocks_per_split;

    if (threadIdx.x == 0) PB2_TS(0);
    pdl_launch_dependents();          // let the next kernel's prologue overlap this kernel
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < g.stages; ++s) {
            mbar_init(&ctl->full[s], 1);
            mbar_init(&ctl->empty[s], 1);
        }
        mbar_init(&ctl->tmem_full, 1);
        mbar_init(&ctl->recv, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&ctl->tmem_base, g.tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    if (threadIdx.x == 0) PB2_TS(1);

    const bool is_epi = warp >= 2;
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int half = is_epi ? (warp - 2) >> 2 : 0;  // which of the two warps of that quarter
    const int m_local = quarter * 32 + lane;
    const int m_glob = m0 + m_local;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const int ep_tid = threadIdx.x - 64;          // 0..255 for epilogue threads
    const int tile = blockIdx.y * gridDim.x + blockIdx.x;
    const size_t tile_elems = static_cast<size_t>(g.block_n) * kBlockM;
    const int n_chunks = g.block_n >> 4;
    EpiCtx e{CounterRng(g.seed), static_cast<uint32_t>(fminf(g.drop_p, 0.999999f) * 4294967296.0f),
             g.drop_p > 0.f ? 1.f / (1.f - g.drop_p) : 1.f, 0.f, m_glob < g.M};
    ChunkPre pre;
    pre.have = false;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            // Programmatic dependent launch: this kernel may have started while its predecessor is
            // still running.  Weights do not depend on it, so the first ring of A tiles is requested
            // right away; the token tiles (and everything downstream of them) wait for it to finish.
            int pre = 0;
            if (g.a_static) {
                pre = kb_count < g.stages ? kb_count : g.stages;
                for (int it = 0; it < pre; ++it) {
                    uint8_t* sa = smem + static_cast<size_t>(it) * stage_bytes;
                    mbar_expect_tx(&ctl->full[it], stage_bytes);
                    load_a_stage(sa, &tmap_a, &ctl->full[it], (kb_begin + it) * kBlockK, m0, kEvictFirst, g.a_mn);
                }
            }
            pdl_wait();
            for (int it = 0; it < pre; ++it) {
                uint8_t* sb = smem + static_cast<size_t>(it) * stage_bytes + a_bytes;
                tma_load_2d(sb, &tmap_b, &ctl->full[it], (kb_begin + it) * kBlockK, n0, kEvictLast);
            }
            for (int it = pre; it < kb_count; ++it) {
                const int s = it % g.stages;
                const uint32_t round = it / g.stages;
                mbar_wait(&ctl->empty[s], (round & 1) ^ 1);
                uint8_t* sa = smem + static_cast<size_t>(s) * stage_bytes;
                uint8_t* sb = sa + a_bytes;
                mbar_expect_tx(&ctl->full[s], stage_bytes);
                const int k_elem = (kb_begin + it) * kBlockK;
                // weights (A) are streamed once: evict-first; tokens (B) are re-read by every M tile: evict-last
                load_a_stage(sa, &tmap_a, &ctl->full[s], k_elem, m0, kEvictFirst, g.a_mn);
                tma_load_2d(sb, &tmap_b, &ctl->full[s], k_elem, n0, kEvictLast);
            }
        }
        __syncwarp();       // reconverge before any warp-aligned barrier below
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t idesc = umma_idesc_bf16_major(kBlockM, g.block_n, g.a_mn ? 1u : 0u, 0u);
        const uint64_t a_step = g.a_mn ? 128 : 2;       // 16 k: two 8-row groups (2048 B) MN-major, 32 bytes K-major
        for (int it = 0; it < kb_count; ++it) {
            const int s = it % g.stages;
            const uint32_t round = it / g.stages;
            mbar_wait(&ctl->full[s], round & 1);
            tc_fence_after();
            if (it == 0 && lane == 0) PB2_TS(2);
            if (elect_one()) {
                const uint32_t sa = smem_u32(smem + static_cast<size_t>(s) * stage_bytes);
                const uint64_t da = a_desc(sa, g.a_mn);
                const uint64_t db = umma_desc_k_sw128(sa + a_bytes);
#pragma unroll
                for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                    // advance 16 elements = 32 bytes along K inside the swizzle row: +2 in 16-byte units
                    umma_bf16(tmem_base, da + a_step * k, db + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&ctl->empty[s]);                      // frees the smem stage when the MMAs retire
                if (it == kb_count - 1) { umma_commit(&ctl->tmem_full); PB2_TS(3); }
            }
            __syncwarp();
        }
    } else {
        // ===================== epilogue (warps 2..9) =====================
        // Cluster split-K: while the main loop runs, everything of this thread's first owned chunk that does not depend
        // on the accumulator — dropout bits, mask / residual operands (the predecessor's output: wait for it first), bias.
        if (g.cluster_reduce) {
            pdl_wait();
            const uint32_t rank_e = cluster_ctarank();
            if (g.bias && g.bias_along_m && e.row_ok) e.bias_m = g.bias[m_glob];
            const int k_first = half * g.splits + static_cast<int>(rank_e);
            if (half < g.local_chunks && k_first < n_chunks && n0 + k_first * 16 < g.N &&
                (g.drop_p > 0.f || g.mask_src || g.addend))
                epilogue_prefetch(g, e, m_glob, n0 + k_first * 16, pre);
        }
        mbar_wait(&ctl->tmem_full, 0);
        tc_fence_after();
        if (threadIdx.x == 64) PB2_TS(4);
    }

    if (g.cluster_reduce) {
        // ---- split-K reduction inside the cluster, through distributed shared memory ----
        // Chunk k (16 accumulator columns) is owned by cluster rank k % splits.  Every CTA sends its
        // partial of each foreign chunk into the owner's staging area (the now idle TMA ring):
        //   staging[src rank][owner-local chunk][column quad 0..3][row 0..127][4 columns]  fp32
        // (a thread owns an accumulator row and ships its 16 columns as four 16-byte DSMEM stores; rows are
        // contiguous within a column quad, so each store instruction of a warp is one 512-byte run)
        // and each owner then finishes only its own chunks, so the fix-up and the epilogue are
        // spread over all CTAs of the tile and no partial tile ever goes to global memory.
        const uint32_t rank = cluster_ctarank();
        if (ep_tid == 0) {
            // bytes about to arrive here: this CTA's chunks from every other rank
            int own = 0;
            for (int k = static_cast<int>(rank); k < n_chunks; k += g.splits) ++own;
            mbar_expect_tx(&ctl->recv, static_cast<uint32_t>(own) * (g.splits - 1) * kBlockM * 64u);
        }
        cluster_sync_relaxed();                               // (1) every CTA's ring is idle (and its barrier armed)
        if (ep_tid == 0) PB2_TS(7);
        if (is_epi) {
            const uint32_t stage_base = smem_u32(smem);
            for (int k = half; k < n_chunks; k += 2) {
                const uint32_t owner = static_cast<uint32_t>(k % g.splits);
                if (owner == rank) continue;
                uint32_t r[16];
                tmem_ld_32x16(t_lane + k * 16, r);
                tmem_ld_wait();
                // block of (source rank, chunk): [4 column quads][128 rows][16 B], so one store instruction of a warp
                // covers 512 contiguous bytes of the owner's shared memory.  (Row-major rows of 80 bytes made every
                // instruction 32 separate 16-byte remote writes: 1.8 us to ship 30 KB instead of 0.6,
                // profiles/chain_kernel_r2.txt.)
                const uint32_t local = stage_base + (rank * g.local_chunks + (k / g.splits)) * (kBlockM * 64u) + m_local * 16u;
                const uint32_t remote = dsmem_addr(local, owner);
                const uint32_t recv_remote = dsmem_addr(smem_u32(&ctl->recv), owner);
                // data and completion signal travel together (st.async): no release / acquire cluster barrier, which
                // would put a MEMBAR.GPU and an L1 invalidation into every thread's path
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    dsmem_st_async_f32x4(remote + (j / 4) * (kBlockM * 16u), recv_remote, __uint_as_float(r[j]),
                                         __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
            }
        }
        if (ep_tid == 0) PB2_TS(8);
        if (ep_tid == 0) PB2_TS(9);
        if (is_epi) {
            mbar_wait(&ctl->recv, 0);                         // (2) all partials of my chunks delivered
            if (ep_tid == 0) PB2_TS(10);
            float ss = 0.f;
            const float* staging = reinterpret_cast<const float*>(smem);
            for (int kl = half; kl < g.local_chunks; kl += 2) {
                const int k = kl * g.splits + static_cast<int>(rank);
                if (k >= n_chunks) break;
                uint32_t r[16];
                tmem_ld_32x16(t_lane + k * 16, r);
                tmem_ld_wait();
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                for (int src = 0; src < g.splits; ++src) {
                    if (src == static_cast<int>(rank)) continue;
                    const float4* p = reinterpret_cast<const float4*>(staging) +
                                      (static_cast<size_t>(src) * g.local_chunks + kl) * (kBlockM * 4) + m_local;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 t = p[q * kBlockM];
                        v[4 * q] += t.x; v[4 * q + 1] += t.y; v[4 * q + 2] += t.z; v[4 * q + 3] += t.w;
                    }
                }
                if (n0 + k * 16 < g.N)
                    epilogue_chunk(g, e, v, m_glob, n0 + k * 16, ss, (kl == half && pre.have) ? &pre : nullptr);
            }
            if (g.sumsq) {
                ss = warp_sum(ss);
                if (lane == 0) atomicAdd(g.sumsq, ss);
            }
            if (ep_tid == 0) PB2_TS(5);
            tc_fence_before();
        }
    } else if (is_epi) {
        bool last = true;
        if (g.splits > 1) {
            // partial tile layout [column][row]: a warp writes 32 consecutive floats per column
            float* mine = g.partials + (static_cast<size_t>(tile) * g.splits + split) * tile_elems + m_local;
            for (int c = half * 16; c < g.block_n; c += 32) {
                uint32_t r[16];
                tmem_ld_32x16(t_lane + c, r);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; ++j) __stcg(mine + static_cast<size_t>(c + j) * kBlockM, __uint_as_float(r[j]));
            }
            __threadfence();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (ep_tid == 0) {
                unsigned int prev = atomicAdd(&g.counters[tile], 1u);
                ctl->is_last = (prev == static_cast<unsigned int>(g.splits - 1)) ? 1u : 0u;
                if (prev == static_cast<unsigned int>(g.splits - 1)) g.counters[tile] = 0u;   // re-arm for the next launch
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            last = ctl->is_last != 0;
            if (last) __threadfence();
        }

        if (last) {
            EpiCtx e{CounterRng(g.seed), static_cast<uint32_t>(fminf(g.drop_p, 0.999999f) * 4294967296.0f),
                     g.drop_p > 0.f ? 1.f / (1.f - g.drop_p) : 1.f, 0.f, m_glob < g.M};
            if (g.bias && g.bias_along_m && e.row_ok) e.bias_m = g.bias[m_glob];
            float ss = 0.f;
            // plain row-major epilogues go through the shared-memory transpose (the TMA ring is idle
            // once the accumulator is complete); everything fancier takes the per-row path
            const bool staged = !g.transpose_out && !g.relu && g.drop_p <= 0.f && !g.mask_src && !g.addend;
            if (!(g.bias && g.bias_along_m)) e.bias_m = 0.f;
            const float* others = g.partials + static_cast<size_t>(tile) * g.splits * tile_elems + m_local;
            for (int c = half * 16; c < g.block_n; c += 32) {
                uint32_t r[16];
                tmem_ld_32x16(t_lane + c, r);
                tmem_ld_wait();
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                for (int sp = 0; sp < g.splits; ++sp) {
                    if (sp == split) continue;
                    const float* p = others + static_cast<size_t>(sp) * tile_elems + static_cast<size_t>(c) * kBlockM;
                    float t[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) t[j] = __ldcg(p + j * kBlockM);
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += t[j];
                }
                if (n0 + c < g.N) {
                    if (staged)
                        epilogue_chunk_staged(g, reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 20), v, m0, quarter, lane,
                                              n0 + c, e.bias_m, ss);
                    else
                        epilogue_chunk(g, e, v, m_glob, n0 + c, ss);
                }
            }
            if (g.sumsq) {
                ss = warp_sum(ss);
                if (lane == 0) atomicAdd(g.sumsq, ss);
            }
        }
        if (ep_tid == 0) PB2_TS(5);
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, g.tmem_cols);
    }
    if (threadIdx.x == 32) PB2_TS(6);
}

// ---------------------------------------------------------------------------------------
// Persistent variant for problems with more output tiles than SMs (no split-K): the weight-gradient
// GEMMs (4096 x 4096 fp32 outputs, 512 tiles) and the large-token GEMMs of big micro-batches.
//
// One CTA per SM walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...  The TMA ring runs straight
// across tile boundaries and the accumulator is double-buffered in TMEM (2 x block_n columns), so
// the tensor core works on tile i+1 while the eight epilogue warps drain tile i: for a K=512 wgrad
// tile the 128 KB fp32 store takes longer than the MMAs, and without the overlap the two add up.
// Tiles are walked M-fastest, so CTAs that run side by side share the same B tile in L2.
//   acc_full[b]  : MMA warp -> epilogue   (tcgen05.commit)
//   acc_empty[b] : epilogue -> MMA warp   (one arrival per epilogue warp)
// ---- TMA store (shared -> global through a tensor map), bulk-group completion ----
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

constexpr int kStoreBufBytes = 32 * 128;            // one 32-row x 32-column fp32 block, 128-byte swizzled
constexpr int kStoreBytes = kEpilogueWarps * 2 * kStoreBufBytes;     // two blocks per epilogue warp

struct __align__(8) SmemCtlP {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
};

__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_persistent_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                            const __grid_constant__ CUtensorMap tmap_out, const GemmArgs g, const int m_tiles,
                            const int n_tiles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages][A 16 KB][B block_n*128 B], then 20 KB of epilogue staging, then the control block
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t a_bytes = kBlockM * kBlockK * 2;
    const uint32_t b_bytes = g.block_n * kBlockK * 2;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    // after the ring: the epilogue's staging area — 20 KB of padded rows for the register-store paths, or (tma_store)
    // 64 KB of swizzled blocks that TMA reads; ring_bytes is a multiple of 1024, so the blocks are 1024-aligned
    float* staging = reinterpret_cast<float*>(smem + g.ring_bytes);
    const uint32_t staging_bytes = g.tma_store ? kStoreBytes : kEpilogueWarps * 32 * kStagePitch * 4;
    SmemCtlP* ctl = reinterpret_cast<SmemCtlP*>(smem + g.ring_bytes + staging_bytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int total_tiles = m_tiles * n_tiles;
    const int my_tiles = (total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const int kbt = g.k_blocks_total;
    const int total_it = my_tiles * kbt;
    const uint32_t acc_stride = static_cast<uint32_t>(g.tmem_cols) >> 1;
    // An operand tile that a whole row / column of output tiles re-reads stays in L2; with a single tile in
    // the other direction (the vocabulary projection: 225 weight tiles x one token tile) it is streamed once.
    const uint64_t a_hint = n_tiles > 1 ? kEvictLast : kEvictFirst;
    const uint64_t b_hint = m_tiles > 1 ? kEvictLast : kEvictFirst;

    pdl_launch_dependents();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < g.stages; ++s) {
            mbar_init(&ctl->full[s], 1);
            mbar_init(&ctl->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&ctl->acc_full[b], 1);
            mbar_init(&ctl->acc_empty[b], kEpilogueWarps);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&ctl->tmem_base, g.tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            int pre = 0;
            if (g.a_static) {
                // weights do not depend on the preceding kernel: request the first ring of A right away
                pre = total_it < g.stages ? total_it : g.stages;
                for (int it = 0; it < pre; ++it) {
                    const int tile = blockIdx.x + (it / kbt) * gridDim.x;
                    mbar_expect_tx(&ctl->full[it], stage_bytes);
                    load_a_stage(smem + static_cast<size_t>(it) * stage_bytes, &tmap_a, &ctl->full[it],
                                 (it % kbt) * kBlockK, (tile % m_tiles) * kBlockM, a_hint, g.a_mn);
                }
            }
            pdl_wait();
            for (int it = 0; it < pre; ++it) {
                const int tile = blockIdx.x + (it / kbt) * gridDim.x;
                tma_load_2d(smem + static_cast<size_t>(it) * stage_bytes + a_bytes, &tmap_b, &ctl->full[it],
                            (it % kbt) * kBlockK, (tile / m_tiles) * g.block_n, b_hint);
            }
            int s = pre % g.stages;
            uint32_t round = pre / g.stages;
            int kb = pre % kbt;
            int tile = blockIdx.x + (pre / kbt) * gridDim.x;
            for (int it = pre; it < total_it; ++it) {
                mbar_wait(&ctl->empty[s], (round & 1) ^ 1);
                uint8_t* sa = smem + static_cast<size_t>(s) * stage_bytes;
                mbar_expect_tx(&ctl->full[s], stage_bytes);
                load_a_stage(sa, &tmap_a, &ctl->full[s], kb * kBlockK, (tile % m_tiles) * kBlockM, a_hint, g.a_mn);
                tma_load_2d(sa + a_bytes, &tmap_b, &ctl->full[s], kb * kBlockK, (tile / m_tiles) * g.block_n, b_hint);
                if (++s == g.stages) { s = 0; ++round; }
                if (++kb == kbt) { kb = 0; tile += gridDim.x; }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t idesc = umma_idesc_bf16_major(kBlockM, g.block_n, g.a_mn ? 1u : 0u, 0u);
        const uint64_t a_step = g.a_mn ? 128 : 2;
        int s = 0;
        uint32_t round = 0;
        for (int t = 0; t < my_tiles; ++t) {
            const uint32_t acc = t & 1;
            mbar_wait(&ctl->acc_empty[acc], ((t >> 1) & 1) ^ 1);      // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * acc_stride;
            for (int kb = 0; kb < kbt; ++kb) {
                mbar_wait(&ctl->full[s], round & 1);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t sa = smem_u32(smem + static_cast<size_t>(s) * stage_bytes);
                    const uint64_t da = a_desc(sa, g.a_mn);
                    const uint64_t db = umma_desc_k_sw128(sa + a_bytes);
#pragma unroll
                    for (int k = 0; k < kBlockK / kUmmaK; ++k)
                        umma_bf16(d_tmem, da + a_step * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&ctl->empty[s]);
                    if (kb == kbt - 1) umma_commit(&ctl->acc_full[acc]);
                }
                __syncwarp();
                if (++s == g.stages) { s = 0; ++round; }
            }
        }
    } else {
        // ===================== epilogue (warps 2..9) =====================
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int m_local = quarter * 32 + lane;
        const bool staged = !g.transpose_out && !g.relu && g.drop_p <= 0.f && !g.mask_src && !g.addend;
        float* st = staging + (warp - 2) * (32 * kStagePitch);
        float ss = 0.f;
        int nb = 0;              // TMA-store blocks issued by this warp so far (alternates its two buffers across tiles too)
        for (int t = 0; t < my_tiles; ++t) {
            const int tile = blockIdx.x + t * gridDim.x;
            const int m0 = (tile % m_tiles) * kBlockM;
            const int n0 = (tile / m_tiles) * g.block_n;
            const int m_glob = m0 + m_local;
            const uint32_t acc = t & 1;
            EpiCtx e{CounterRng(g.seed), static_cast<uint32_t>(fminf(g.drop_p, 0.999999f) * 4294967296.0f),
                     g.drop_p > 0.f ? 1.f / (1.f - g.drop_p) : 1.f, 0.f, m_glob < g.M};
            if (g.bias && g.bias_along_m && e.row_ok) e.bias_m = g.bias[m_glob];
            mbar_wait(&ctl->acc_full[acc], (t >> 1) & 1);
            tc_fence_after();
            const uint32_t t_lane = tmem_base + acc * acc_stride + (static_cast<uint32_t>(quarter * 32) << 16);
            if (g.tma_store) {
                // Plain fp32 row-major result (the weight gradients: 128 KB per tile, the kernel's main cost): the warp
                // parks a 32-row x 32-column block in shared memory (128-byte rows, TMA's swizzle) and one lane hands it
                // to the TMA unit, which writes whole lines and clips at the matrix edge; two blocks per warp in flight.
                // ~10x fewer instructions than register stores: the epilogue no longer sets the pace of the tile.
                uint8_t* bufs = reinterpret_cast<uint8_t*>(staging) + (warp - 2) * (2 * kStoreBufBytes);
                for (int c = half * 32; c < g.block_n; c += 64, ++nb) {
                    uint32_t r[32];
                    tmem_ld_32x32(t_lane + c, r);
                    tmem_ld_wait();
                    uint8_t* buf = bufs + (nb & 1) * kStoreBufBytes;
                    if (lane == 0) tma_store_wait_read<1>();        // the block stored two rounds ago has been read
                    __syncwarp();
                    if (n0 + c < g.N) {
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * g.alpha;
                        if (g.sumsq && m_glob < g.M) {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (n0 + c + j < g.N) ss += v[j] * v[j];
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                        fence_proxy_async();
                        __syncwarp();
                        if (lane == 0) {
                            tma_store_2d(&tmap_out, buf, n0 + c, m0 + quarter * 32);
                            tma_store_commit();
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->acc_empty[acc]);
                continue;
            }
            for (int c = half * 16; c < g.block_n; c += 32) {
                uint32_t r[16];
                tmem_ld_32x16(t_lane + c, r);
                tmem_ld_wait();
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                if (n0 + c < g.N) {
                    if (staged)
                        epilogue_chunk_staged(g, st, v, m0, quarter, lane, n0 + c, e.bias_m, ss);
                    else
                        epilogue_chunk(g, e, v, m_glob, n0 + c, ss);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->acc_empty[acc]);      // the MMA warp may overwrite this accumulator
        }
        if (g.sumsq) {
            ss = warp_sum(ss);
            if (lane == 0) atomicAdd(g.sumsq, ss);
        }
        if (g.tma_store && lane == 0) tma_store_wait_all();          // the bulk stores must land before the CTA exits
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, g.tmem_cols);
    }
}

static int next_pow2_cols(int n) {
    int c = 32;
    while (c < n) c <<= 1;
    return c;
}

}  // namespace pb2

using namespace pb2;

extern "C" size_t pb2_gemm_workspace_bytes(void) {
    // partials: up to 148*4 CTAs' worth of 128 x 256 fp32 tiles is never needed at once; the launcher
    // bounds splits so that tiles*splits*block_n*128*4 <= this.  Counters live at the end.
    return PB2_GEMM_WORKSPACE_BYTES;
}

static unsigned long long* g_debug_ts = nullptr;
// Debug hook (not part of the supported ABI): phase timestamps of subsequent launches go to `buf`
// (8 x u64 per CTA, CTA-linear order); pass NULL to switch off.
extern "C" void pb2_debug_gemm_timestamps(unsigned long long* buf) { g_debug_ts = buf; }

// SM count of the current device (cached per device ordinal).
static int sm_count() {
    static int cache[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cache[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

static int gemm_launch(const pb2_gemm_desc* d, float* partials, unsigned int* counters, size_t partial_bytes,
                       size_t counter_slots, cudaStream_t stream, bool force_scratch = false) {
    if (d->M <= 0 || d->N <= 0 || d->K <= 0) {
        set_error("pb2_gemm: non-positive shape M=%d N=%d K=%d", d->M, d->N, d->K);
        return -1;
    }
    if ((d->lda % 8) || (d->ldb % 8)) {
        set_error("pb2_gemm: lda/ldb must be multiples of 8 elements (16-byte TMA rows), got %lld / %lld", d->lda, d->ldb);
        return -1;
    }
    if ((reinterpret_cast<uintptr_t>(d->a) & 15) || (reinterpret_cast<uintptr_t>(d->b) & 15)) {
        set_error("pb2_gemm: A and B must be 16-byte aligned");
        return -1;
    }
    GemmArgs g{};
    g.M = d->M; g.N = d->N; g.K = d->K;
    int bn = d->block_n;
    if (bn <= 0) {
        bn = ((d->N + 15) / 16) * 16;
        if (bn > 256) bn = 256;
        if (d->N > 256) {
            // Pick the number of N tiles.  Widest tiles (256) are the most efficient per tile, but with one
            // persistent CTA per SM the last round of tiles can leave most SMs idle: 48 x 4 = 192 tiles are two
            // rounds on 148 SMs for the work of 1.3.  Cost model fitted to the 1024-token shapes of the
            // reference-as-shipped configuration (scripts/bench_gemm.py, SWEEP=shipped): rounds x (tile width + 96);
            // a narrower tile must promise > 8 % before it is taken (QKV forward 6144 x 1024 x 2048: 4 -> 6 N tiles,
            // 29.1 -> 25.8 us; every other measured shape keeps 256).  Small problems with a short reduction
            // (K <= 2048) that would otherwise run as <= 74 tiles x split-K take enough N tiles to fill the SMs
            // instead (2048 x 1024 x 2048: 16.3 -> 14.7 us).
            const int m_t = (d->M + kBlockM - 1) / kBlockM;
            const int sms = sm_count();
            const int nt_min = (d->N + 255) / 256, nt_max = (d->N + 127) / 128;
            auto width = [&](int nt) { return (((d->N + nt - 1) / nt) + 15) / 16 * 16; };
            int best_nt = nt_min;
            if (static_cast<long long>(m_t) * nt_min > sms) {
                auto cost = [&](int nt) {
                    const long long tiles = static_cast<long long>(m_t) * nt;
                    return static_cast<double>((tiles + sms - 1) / sms) * (width(nt) + 96);
                };
                const double base = cost(nt_min);
                double best = base;
                for (int nt = nt_min + 1; nt <= nt_max; ++nt) {
                    if (width(nt) % 16) continue;
                    const double c = cost(nt);
                    if (c < best && c < 0.92 * base) { best = c; best_nt = nt; }
                }
            } else if (d->K <= 2048 && static_cast<long long>(m_t) * nt_min * 2 <= sms && d->splits <= 0) {
                while (best_nt < nt_max && static_cast<long long>(m_t) * (best_nt + 1) <= sms) ++best_nt;
            }
            bn = width(best_nt);
        }
    }
    if (bn % 16 || bn < 16 || bn > 256) {
        set_error("pb2_gemm: block_n must be a multiple of 16 in [16, 256], got %d", bn);
        return -1;
    }
    g.block_n = bn;
    g.k_blocks_total = (d->K + kBlockK - 1) / kBlockK;
    const int m_tiles = (d->M + kBlockM - 1) / kBlockM;
    const int n_tiles = (d->N + bn - 1) / bn;
    const long long tiles = static_cast<long long>(m_tiles) * n_tiles;

    int splits = d->splits;
    if (splits <= 0) {
        // aim for roughly one CTA per SM (a CTA streams ~130 GB/s, so ~64 already saturate HBM), keep
        // at least 4 k-blocks per split, and stay within one cluster (<= 8) so the reduction can
        // run through distributed shared memory
        splits = 1;
        while (tiles * splits * 2 <= 160 && g.k_blocks_total / (splits * 2) >= 4 && splits < 8) splits *= 2;
    }
    if (splits > g.k_blocks_total) splits = g.k_blocks_total;
    g.k_blocks_per_split = (g.k_blocks_total + splits - 1) / splits;
    splits = (g.k_blocks_total + g.k_blocks_per_split - 1) / g.k_blocks_per_split;
    g.splits = splits;

    const int stage_bytes = (kBlockM + bn) * kBlockK * 2;
    int stages = (200 * 1024) / stage_bytes;
    {
        // With more CTAs than SMs (and no persistent schedule) the ring is halved.  Measured, for the
        // record: letting two GEMM CTAs share an SM so that the next GEMM's weight tiles arrive during
        // this one's epilogue (__launch_bounds__(320, 2) = 96 registers, PB2_GEMM_RING_KB=108) made the
        // training step 9 % slower, 84 KB 28 % slower: with ~3 us of loaded HBM latency a CTA needs its
        // full ~180 KB of loads in flight to hold its share of the bandwidth, and an L2 prefetch of the next
        // GEMM's weights from this kernel's tail changed nothing (12.3 -> 12.0 us per chained E x E GEMM); neither
        // did a look-ahead of 4-16 k-blocks of this kernel's own weight stream into L2 with
        // cp.async.bulk.prefetch.tensor (12.7 -> 12.6-13.6 us, and no rescue for a 108 KB ring).
        // Also measured and dropped: (a) half-height tiles instead of split-K — UMMA M = 64 (accumulator row r
        // sits in TMEM lane r % 16 + 32 * (r / 16); numerics verified), 64 x 2 token halves = 128 CTAs, no
        // exchange after the main loop — runs the E x E shape in 18.3 us instead of 12.7: the main loop then
        // needs ~217 ns per 64-wide k-block per CTA whatever the CTA count, the L2 policy of the shared
        // weight tile or the number of independent accumulators (1/2/4), i.e. SS-mode MMAs at M = 64 cost
        // ~100 cycles each (cuBLAS's 64-row nvjet kernel for this shape is a cta_group::2 kernel);
        // (b) issuing the A and B tensor copies from two different threads: no change.
        static int cap_kb = -1;
        if (cap_kb < 0) {
            const char* e = getenv("PB2_GEMM_RING_KB");
            cap_kb = e ? atoi(e) : 200;
        }
        const int kb = (tiles * splits > 148 && cap_kb > 110) ? 110 : cap_kb;   // > 1 wave: always two per SM
        const int two = (kb * 1024) / stage_bytes;
        if (two >= 2 && two < stages) stages = two;
    }
    if (stages > kMaxStages) stages = kMaxStages;
    if (d->stages > 0) stages = d->stages < (200 * 1024) / stage_bytes ? d->stages : (200 * 1024) / stage_bytes;   // explicit request wins
    if (stages > g.k_blocks_per_split) stages = g.k_blocks_per_split;
    if (stages < 1) stages = 1;
    g.stages = stages;
    g.tmem_cols = next_pow2_cols(bn);

    // split-K reduction: inside a cluster through DSMEM when the splits fit one (2..8 CTAs) and the
    // staging area fits the idle TMA ring; otherwise through the global scratch workspace
    const int n_chunks = bn / 16;
    g.local_chunks = (n_chunks + splits - 1) / splits;
    const size_t staging = static_cast<size_t>(splits) * g.local_chunks * kBlockM * 64;
    const bool pow2 = (splits & (splits - 1)) == 0;
    g.cluster_reduce = (splits > 1 && splits <= 8 && pow2 && !force_scratch && d->reduce_mode != 1 &&
                        staging <= static_cast<size_t>(stages) * stage_bytes) ? 1 : 0;
    if (d->reduce_mode == 2 && !g.cluster_reduce && splits > 1) {
        set_error("pb2_gemm: cluster reduction requested but not possible (splits=%d, staging %zu B, ring %d B)",
                  splits, staging, stages * stage_bytes);
        return -1;
    }
    if (splits > 1 && !g.cluster_reduce) {
        const size_t need = static_cast<size_t>(tiles) * splits * bn * kBlockM * sizeof(float);
        if (need > partial_bytes || static_cast<size_t>(tiles) > counter_slots) {
            set_error("pb2_gemm: split-K scratch too small (need %zu B partials for %lld tiles x %d splits, have %zu B / %zu counters)",
                      need, tiles, splits, partial_bytes, counter_slots);
            return -1;
        }
        g.partials = partials;
        g.counters = counters;
    }

    g.out = d->out; g.ld_out = d->ld_out; g.out_f32 = d->out_f32; g.transpose_out = d->transpose_out;
    g.bias = d->bias; g.bias_along_m = d->bias_along_m; g.relu = d->relu;
    g.drop_p = d->drop_p; g.seed = d->seed; g.stream = d->rng_stream;
    g.mask_src = static_cast<const __nv_bfloat16*>(d->mask_src); g.mask_scale = d->mask_scale;
    g.addend = static_cast<const __nv_bfloat16*>(d->addend);
    g.accumulate = d->accumulate; g.alpha = d->alpha == 0.f ? 1.f : d->alpha;
    g.sumsq = d->sumsq;
    g.a_static = d->a_static;
    g.a_mn = d->a_mn_major ? 1 : 0;
    g.debug_ts = g_debug_ts;

    CUtensorMap ta, tb;
    if (g.a_mn) {
        // A is [K, M] with M contiguous: boxes of 64 k-rows x 64 m-columns
        if (make_tmap_bf16_2d(&ta, d->a, d->K, d->M, d->lda, kBlockK, 64)) return -1;
    } else if (make_tmap_bf16_2d(&ta, d->a, d->M, d->K, d->lda, kBlockM, kBlockK)) return -1;
    if (make_tmap_bf16_2d(&tb, d->b, d->N, d->K, d->ldb, bn, kBlockK)) return -1;

    // more tiles than SMs and no split-K: one persistent CTA per SM, epilogue overlapped with the next tile
    if (splits == 1 && d->persistent != 2 && (d->persistent == 1 || tiles > sm_count())) {
        int pst = (204 * 1024) / stage_bytes;
        if (pst > kMaxStages) pst = kMaxStages;
        if (d->stages > 0 && d->stages < pst) pst = d->stages;
        // short reductions (the K = 500-token weight gradients: 8 k-blocks per tile) measured 4-5 % faster with
        // a 3-stage ring than with 4 (27.3 vs 28.7 us, 67.2 vs 70.4 us); long ones want the deepest ring
        if (d->stages <= 0 && g.k_blocks_total <= 8 && pst > 3) pst = 3;
        if (pst >= 2 && 2 * bn <= 512) {
            // plain fp32 row-major results (the weight gradients) leave through TMA stores
            static int tma_store_off = -1;
            if (tma_store_off < 0) {
                const char* e = getenv("PB2_GEMM_TMA_STORE");
                tma_store_off = (e && e[0] == '0') ? 1 : 0;
            }
            CUtensorMap tout = ta;          // placeholder when unused
            g.tma_store = (!tma_store_off && d->out_f32 && !d->transpose_out && !d->accumulate && !d->relu && d->drop_p <= 0.f &&
                           !d->mask_src && !d->addend && !d->bias && (d->ld_out % 4) == 0 && bn % 32 == 0 &&
                           (reinterpret_cast<uintptr_t>(d->out) & 15) == 0) ? 1 : 0;
            size_t staging_bytes = kEpilogueWarps * 32 * kStagePitch * 4;
            if (g.tma_store) {
                staging_bytes = kStoreBytes;
                while (pst > 2 && static_cast<size_t>(pst) * stage_bytes + staging_bytes + sizeof(SmemCtlP) + 1024 > 227 * 1024) --pst;
                if (static_cast<size_t>(pst) * stage_bytes + staging_bytes + sizeof(SmemCtlP) + 1024 > 227 * 1024) g.tma_store = 0;
            }
            if (g.tma_store) {
                if (make_tmap_f32_2d(&tout, d->out, d->M, d->N, d->ld_out, 32, 32)) return -1;
            } else {
                staging_bytes = kEpilogueWarps * 32 * kStagePitch * 4;
            }
            g.stages = pst;
            g.ring_bytes = pst * stage_bytes;
            g.tmem_cols = next_pow2_cols(2 * bn);
            g.debug_ts = nullptr;
            const size_t smem_p = static_cast<size_t>(g.ring_bytes) + staging_bytes + sizeof(SmemCtlP) + 1024;
            if (ensure_max_smem(reinterpret_cast<const void*>(gemm_bf16_persistent_kernel), 227 * 1024)) return -1;
            cudaLaunchConfig_t pc;
            memset(&pc, 0, sizeof(pc));
            pc.gridDim = dim3(static_cast<unsigned>(tiles < sm_count() ? tiles : sm_count()));
            pc.blockDim = dim3(kThreads);
            pc.dynamicSmemBytes = smem_p;
            pc.stream = stream;
            cudaLaunchAttribute pa[1];
            pa[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            pa[0].val.programmaticStreamSerializationAllowed = 1;
            pc.attrs = pa;
            pc.numAttrs = 1;
            PB2_CUDA(cudaLaunchKernelEx(&pc, gemm_bf16_persistent_kernel, ta, tb, tout, g, m_tiles, n_tiles));
            PB2_LAUNCHED("gemm_bf16_persistent_kernel");
            return 0;
        }
    }

    g.ring_bytes = stages * stage_bytes;
    if (g.ring_bytes < kEpilogueWarps * 32 * 20 * 4) g.ring_bytes = kEpilogueWarps * 32 * 20 * 4;   // staging needs 20 KB
    const size_t smem = static_cast<size_t>(g.ring_bytes) + sizeof(SmemCtl) + 1024;
    if (ensure_max_smem(reinterpret_cast<const void*>(gemm_bf16_tn_kernel), 227 * 1024)) return -1;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(m_tiles, n_tiles, splits);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n_attr = 0;
    // programmatic dependent launch: the kernel's prologue (and, for static weights, its first
    // weight tiles) may overlap the tail of the preceding kernel; griddepcontrol.wait orders the rest
    attr[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n_attr].val.programmaticStreamSerializationAllowed = 1;
    ++n_attr;
    if (g.cluster_reduce) {
        attr[n_attr].id = cudaLaunchAttributeClusterDimension;
        attr[n_attr].val.clusterDim.x = 1;
        attr[n_attr].val.clusterDim.y = 1;
        attr[n_attr].val.clusterDim.z = splits;
        ++n_attr;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n_attr;
    PB2_CUDA(cudaLaunchKernelEx(&cfg, gemm_bf16_tn_kernel, ta, tb, g));
    PB2_LAUNCHED("gemm_bf16_tn_kernel");
    return 0;
}

extern "C" int pb2_gemm_bf16(const pb2_gemm_desc* d, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (!d) { set_error("pb2_gemm_bf16: null descriptor"); return -1; }
    float* partials = nullptr;
    unsigned int* counters = nullptr;
    size_t partial_bytes = 0, slots = 0;
    if (workspace && workspace_bytes > PB2_GEMM_COUNTER_BYTES) {
        partial_bytes = workspace_bytes - PB2_GEMM_COUNTER_BYTES;
        partials = static_cast<float*>(workspace);
        counters = reinterpret_cast<unsigned int*>(static_cast<char*>(workspace) + partial_bytes);
        slots = (PB2_GEMM_COUNTER_BYTES - PB2_CHAIN_BARRIER_BYTES) / sizeof(unsigned int);   // the tail belongs to pb2_chain_run
    }
    return gemm_launch(d, partials, counters, partial_bytes, slots, stream);
}

// ---------------------------------------------------------------------------------------
// pb2_gemm_bf16_noalloc: the same GEMM for callers that cannot obtain scratch memory.
//
// Some hosts of this library have no way to hand over a workspace: a plug-in running inside
// a server that reserved all of HBM for its own pool at start-up, a control loop that is not
// allowed to touch cudaMalloc because of the device-wide synchronisation it implies, a C
// caller with no tensor library to allocate through.  For them the split-K scratch lives
// in the library image itself: a file-scope __device__ array of PB2_GEMM_STATIC_SCRATCH_BYTES
// (+ the arrival counters).  Every GPU that loads the library carries this array until the
// process exits, which is why it is kept modest; problems whose partial tiles do not fit
// are rejected, not silently shrunk.
//
// The array's name is only a host-side handle.  Its device address is looked up once per
// device with cudaGetSymbolAddress while that device is current, and it is that address
// the kernel receives.  The look-up happens inside the first call on each device, so make
// that first call during start-up, outside any stream capture.
//
// All calls share the one array and it keeps whatever the previous call left behind.  The
// arrival counters — the only part that must start at zero — are cleared on the caller's
// stream ahead of every launch; the partial tiles are always written before they are read.
// Two calls that can overlap on the GPU (different streams, or different host threads on
// different streams) would race on the array: serialise them.
//
// The array is looked up for the device that is current at the time of the call, so `stream` must
// belong to the current device (a stream of another GPU would be given this GPU's scratch address).
//
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[PB2_GEMM_STATIC_SCRATCH_BYTES + PB2_GEMM_COUNTER_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* s_scratch_addr[kMaxDevices];            // device address of g_static_scratch, per device
std::atomic<int> s_scratch_state[kMaxDevices];  // 0 = not looked up, 1 = ready
std::atomic_flag s_scratch_lock = ATOMIC_FLAG_INIT;
}  // namespace

extern "C" size_t pb2_gemm_noalloc_max_scratch_bytes(void) { return PB2_GEMM_STATIC_SCRATCH_BYTES; }

extern "C" int pb2_gemm_bf16_noalloc(const pb2_gemm_desc* d, cudaStream_t stream) {
    if (!d) { set_error("pb2_gemm_bf16_noalloc: null descriptor"); return -1; }
    int dev = -1;
    PB2_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("pb2_gemm_bf16_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
        return -1;
    }
    if (s_scratch_state[dev].load(std::memory_order_acquire) == 0) {
        // one-time set-up for this device (the current one): resolve the symbol to a device address
        while (s_scratch_lock.test_and_set(std::memory_order_acquire)) std::this_thread::yield();
        int rc = 0;
        if (s_scratch_state[dev].load(std::memory_order_relaxed) == 0) {
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || p == nullptr) {
                set_error("pb2_gemm_bf16_noalloc: cudaGetSymbolAddress(static scratch) on device %d: %s", dev,
                          cudaGetErrorString(e));
                rc = -1;
            } else {
                s_scratch_addr[dev] = p;
                s_scratch_state[dev].store(1, std::memory_order_release);
            }
        }
        s_scratch_lock.clear(std::memory_order_release);
        if (rc) return rc;
    }
    char* base = static_cast<char*>(s_scratch_addr[dev]);
    float* partials = reinterpret_cast<float*>(base);
    unsigned int* counters = reinterpret_cast<unsigned int*>(base + PB2_GEMM_STATIC_SCRATCH_BYTES);
    // the counters must be zero when the kernel starts, whatever the last call (or a failed one) left
    PB2_CUDA(cudaMemsetAsync(counters, 0, PB2_GEMM_COUNTER_BYTES, stream));
    int rc = gemm_launch(d, partials, counters, PB2_GEMM_STATIC_SCRATCH_BYTES,
                         PB2_GEMM_COUNTER_BYTES / sizeof(unsigned int), stream, /*force_scratch=*/true);
    if (rc) {
        // make the limit explicit for the one failure mode that is specific to this entry point
        const char* why = pb2_last_error();
        if (why && strstr(why, "scratch too small")) {
            char buf[400];
            snprintf(buf, sizeof(buf), "%s", why);
            set_error("pb2_gemm_bf16_noalloc: the built-in scratch is limited to %u bytes of split-K partials; "
                      "lower `splits`/`block_n` or use pb2_gemm_bf16 with your own workspace [%s]",
                      static_cast<unsigned>(PB2_GEMM_STATIC_SCRATCH_BYTES), buf);
        }
    }
    return rc;
}
