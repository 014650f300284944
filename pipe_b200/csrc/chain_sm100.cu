// A run of dependent skinny GEMMs (and the LayerNorms between them) as ONE persistent kernel.
//
// Why: with 25-128 tokens per micro-batch every Linear of a Transformer layer is a weight-streaming GEMM of
// 5-16 us of HBM time, and as separate kernels each of them pays a launch, a TMEM allocation, a cold TMA ring,
// an epilogue during which nothing is streamed and a tear-down: 12.7 us measured for the 33 MB E x E weight
// whose bytes take 5.2 us (profiles/gemm_bench_r2.txt).  Every boundary is a full dependency (LayerNorm needs
// whole rows, the next GEMM reduces over all features), so the kernels cannot simply be merged tile by tile —
// but the WEIGHTS of the next GEMM depend on nothing.  Here one CTA per SM stays resident for the whole run:
//
//   warp 0   weight producer: walks every GEMM of the chain ahead of time and keeps the TMA ring full of
//            weight tiles, across op boundaries — HBM keeps streaming while the epilogue of the previous op,
//            the grid barrier and a LayerNorm run
//   warp 10  token producer: waits for the grid barrier of the op that produces its input, then loads the
//            activation tiles (L2-resident, 10-16 KB per stage) into the same ring slots
//   warp 1   MMA issuer: tcgen05.mma into one of two TMEM accumulators (so the next tile's MMAs overlap the
//            epilogue of this one)
//   warps 2-9 epilogue: split-K reduction across the 4-CTA cluster through distributed shared memory, then
//            the same fused epilogue as the one-GEMM kernel (bias / ReLU / dropout / mask / residual), then
//            the arrival on the grid barrier; they also execute the add + LayerNorm ops (one row per CTA)
//
// A tile (128 output features x all tokens) is owned by a cluster of four CTAs that split K four ways — the
// decomposition the one-GEMM kernel picks for E x E at 75 tokens, kept for every GEMM of the chain so that
// the reduction order (own partial, then cluster ranks in ascending order) does not depend on the shape.
// The cluster handshake uses mbarriers instead of barrier.cluster because the producer warps run ahead of the
// epilogue by design and cannot take part in a cluster-wide barrier at tile boundaries.
//
// Grid barrier: one 64-bit arrival counter in the caller's workspace; op i is complete when it has reached
// (i + 1) x gridDim.x.  The CTA that leaves the kernel last puts it back to zero.  All CTAs are co-resident:
// the grid is sized with cudaOccupancyMaxActiveClusters and a dependent (programmatically launched) kernel
// is only scheduled once every CTA of this grid has started.
//
// Reference counterpart: none — these are separate cuBLAS / ATen calls under nn.TransformerEncoderLayer
// (main.py:148).  SURVEY.md K2/K4/K5, §6.2 (weight-streaming regime).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "gemm_epilogue.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kChainCluster = 4;
constexpr int kChainThreads = kThreads + 32;        // + the token-producer warp
constexpr int kChainMaxGemms = 6;
constexpr int kChainMaxLns = 4;
constexpr int kLnVecs = 4;                          // 4 x 8 bf16 x 256 threads = 8192 columns (as norm_embed.cu)

struct ChainLn {
    const __nv_bfloat16* x;
    const __nv_bfloat16* h;
    const float* gamma;
    const float* beta;
    __nv_bfloat16* y;
    float* mean;
    float* rstd;
    int rows, E;
    float eps;
};

struct ChainParams {
    CUtensorMap ta[kChainMaxGemms];
    CUtensorMap tb[kChainMaxGemms];
    GemmArgs g[kChainMaxGemms];
    int m_tiles[kChainMaxGemms];
    ChainLn ln[kChainMaxLns];
    int kind[PB2_CHAIN_MAX_OPS];      // 0 = GEMM, 1 = add + LayerNorm
    int idx[PB2_CHAIN_MAX_OPS];       // index into g / ln
    int n_ops;
    int stages;
    int bn;                           // token tile (all GEMMs of a chain see the same tokens)
    int acc_cols;                     // TMEM columns per accumulator
    int local_chunks;                 // 16-column chunks a cluster rank owns
    int pf_ahead;                     // k-blocks the L2 prefetch cursor runs ahead of the loads
    unsigned long long* bar;          // [0] arrivals, [1] CTAs that left
    unsigned long long* ts;           // debug timestamps or null
};

struct __align__(8) ChainCtl {
    uint64_t full_a[kMaxStages];
    uint64_t full_b[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint64_t recv;          // every peer's partials of my chunks have landed (remote arrivals)
    uint64_t sent_free;     // every peer has consumed what I sent for the previous tile (remote arrivals)
    uint32_t tmem_base;
    uint32_t pad;
    float red[32];
};

// debug: per-CTA phase timestamps (globaltimer ns), [cta][op 0..8][8]; row 8 = {kernel start, kernel end}
static unsigned long long* g_chain_ts = nullptr;
__device__ __forceinline__ void chain_ts(unsigned long long* ts, int op, int slot) {
    if (ts) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        ts[(static_cast<size_t>(blockIdx.x) * 9 + op) * 8 + slot] = t;
    }
}

__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* m, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void ep_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// Block until `count` ops of this launch are complete on every CTA.
__device__ __forceinline__ void grid_wait(const unsigned long long* bar, int count) {
    const unsigned long long want = static_cast<unsigned long long>(count) * gridDim.x;
    while (ld_acquire_u64(bar) < want) {
    }
}

// sum over the 256 epilogue threads; same tree as block_sum() of norm_embed.cu so the statistics match bit for bit
__device__ __forceinline__ float ep_sum(float v, float* red, int ep_tid) {
    v = warp_sum(v);
    const int w = ep_tid >> 5, lane = ep_tid & 31;
    ep_bar();
    if (lane == 0) red[w] = v;
    ep_bar();
    float t = ep_tid < 8 ? red[ep_tid] : 0.f;
    if (w == 0) {
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    ep_bar();
    return red[0];
}

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&f)[8]) {
    const uint4 raw = __ldcg(reinterpret_cast<const uint4*>(p));      // written by other SMs during this launch: L2, not L1
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h2[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&f)[8]) {
    uint4 raw;
    __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int j = 0; j < 4; ++j) h2[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(p) = raw;
}

// y = LayerNorm(x + h) for one row, by the 256 epilogue threads (the arithmetic of add_ln_fwd_kernel)
__device__ __forceinline__ void ln_row(const ChainLn& L, size_t row, float* red, int ep_tid) {
    const int E = L.E;
    const __nv_bfloat16* xr = L.x + row * E;
    const __nv_bfloat16* hr = L.h ? L.h + row * E : nullptr;
    float z[kLnVecs][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kLnVecs; ++i) {
        const int c = (i * 256 + ep_tid) * 8;
        if (c < E) {
            ld8(xr + c, z[i]);
            if (hr) {
                float t[8];
                ld8(hr + c, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[i][j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) s += z[i][j];
        }
    }
    const float mean = ep_sum(s, red, ep_tid) / E;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kLnVecs; ++i) {
        const int c = (i * 256 + ep_tid) * 8;
        if (c < E) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = z[i][j] - mean;
                q += d * d;
            }
        }
    }
    const float rstd = rsqrtf(ep_sum(q, red, ep_tid) / E + L.eps);
    if (ep_tid == 0) {
        L.mean[row] = mean;
        L.rstd[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < kLnVecs; ++i) {
        const int c = (i * 256 + ep_tid) * 8;
        if (c < E) {
            float o[8];
            const float4 g0 = *reinterpret_cast<const float4*>(L.gamma + c), g1 = *reinterpret_cast<const float4*>(L.gamma + c + 4);
            const float4 b0 = *reinterpret_cast<const float4*>(L.beta + c), b1 = *reinterpret_cast<const float4*>(L.beta + c + 4);
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (z[i][j] - mean) * rstd * gg[j] + bb[j];
            st8(L.y + row * E + c, o);
        }
    }
}

__global__ void __launch_bounds__(kChainThreads, 1)
chain_kernel(const __grid_constant__ ChainParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages][A 16 KB][B bn*128 B] | split-K staging (3 sources x local_chunks x 128 rows x 64 B) | control
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t a_bytes = kBlockM * kBlockK * 2;
    const uint32_t b_bytes = P.bn * kBlockK * 2;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    const uint32_t ring_bytes = P.stages * stage_bytes;
    const uint32_t staging_bytes = (kChainCluster - 1) * P.local_chunks * kBlockM * 64;
    float* staging = reinterpret_cast<float*>(smem + ring_bytes);
    ChainCtl* ctl = reinterpret_cast<ChainCtl*>(smem + ring_bytes + staging_bytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int cluster_id = blockIdx.x / kChainCluster;
    const int n_clusters = gridDim.x / kChainCluster;

    pdl_launch_dependents();
    if (threadIdx.x == 0) chain_ts(P.ts, 8, 0);
    if (warp == 0 && lane == 0) {
        for (int s = 0; s < P.stages; ++s) {
            mbar_init(&ctl->full_a[s], 1);
            mbar_init(&ctl->full_b[s], 1);
            mbar_init(&ctl->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&ctl->acc_full[b], 1);
            mbar_init(&ctl->acc_empty[b], kEpilogueWarps);
        }
        mbar_init(&ctl->recv, 1);
        mbar_init(&ctl->sent_free, (kChainCluster - 1) * kEpilogueWarps);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&ctl->tmem_base, 2 * P.acc_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cluster_sync_all();             // peers' barriers are initialised before anyone arrives on them remotely
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        // ===================== weight producer: never waits for anything but ring space =====================
        if (elect_one()) {
            for (int gi = 0; gi < kChainMaxGemms; ++gi)
                if (P.m_tiles[gi] > 0) tma_prefetch_desc(&P.ta[gi]);
            // A second cursor runs `P.pf_ahead` k-blocks ahead of the loads and asks L2 for those tiles: when the ring is
            // full (epilogue of the previous op, grid barrier, LayerNorm) DRAM keeps streaming into L2, and the loads
            // that follow the barrier are L2 hits.
            struct Cursor {
                int op, t, kb, kbn, kb0, gi;
                bool done;
            };
            auto seek = [&](Cursor& c) {            // position on the first k-block of the next GEMM with work, from c.op
                for (; c.op < P.n_ops; ++c.op) {
                    if (P.kind[c.op] != 0) continue;
                    c.gi = P.idx[c.op];
                    const GemmArgs& g = P.g[c.gi];
                    c.kb0 = rank * g.k_blocks_per_split;
                    c.kbn = g.k_blocks_total - c.kb0;
                    if (c.kbn > g.k_blocks_per_split) c.kbn = g.k_blocks_per_split;
                    c.t = cluster_id;
                    c.kb = 0;
                    if (c.t < P.m_tiles[c.gi] && c.kbn > 0) return;
                }
                c.done = true;
            };
            auto advance = [&](Cursor& c) {
                if (++c.kb < c.kbn) return;
                c.kb = 0;
                c.t += n_clusters;
                if (c.t < P.m_tiles[c.gi]) return;
                ++c.op;
                seek(c);
            };
            Cursor ld{0, 0, 0, 0, 0, 0, false}, pf{0, 0, 0, 0, 0, 0, false};
            seek(ld);
            seek(pf);
            uint32_t it = 0, pf_it = 0;
            while (!ld.done) {
                while (!pf.done && pf_it < it + static_cast<uint32_t>(P.pf_ahead)) {
                    if (pf_it >= it + static_cast<uint32_t>(P.stages)) {        // tiles the ring is about to load need no hint
                        const GemmArgs& g = P.g[pf.gi];
                        const int k_elem = (pf.kb0 + pf.kb) * kBlockK, m0 = pf.t * kBlockM;
                        if (g.a_mn) {
                            tma_prefetch_l2_2d(&P.ta[pf.gi], m0, k_elem);
                            tma_prefetch_l2_2d(&P.ta[pf.gi], m0 + 64, k_elem);
                        } else {
                            tma_prefetch_l2_2d(&P.ta[pf.gi], k_elem, m0);
                        }
                    }
                    ++pf_it;
                    advance(pf);
                }
                const int s = it % P.stages;
                mbar_wait(&ctl->empty[s], ((it / P.stages) & 1) ^ 1);
                uint8_t* sa = smem + static_cast<size_t>(s) * stage_bytes;
                mbar_expect_tx(&ctl->full_a[s], a_bytes);
                load_a_stage(sa, &P.ta[ld.gi], &ctl->full_a[s], (ld.kb0 + ld.kb) * kBlockK, ld.t * kBlockM, kEvictFirst,
                             P.g[ld.gi].a_mn);
                ++it;
                advance(ld);
            }
        }
        __syncwarp();
    } else if (warp == 10) {
        // ===================== token producer: waits for the op that produces its operand =====================
        if (elect_one()) {
            for (int gi = 0; gi < kChainMaxGemms; ++gi)
                if (P.m_tiles[gi] > 0) tma_prefetch_desc(&P.tb[gi]);
            pdl_wait();
            uint32_t it = 0;
            for (int op = 0; op < P.n_ops; ++op) {
                if (P.kind[op] != 0) continue;
                const int gi = P.idx[op];
                const GemmArgs& g = P.g[gi];
                const int kb0 = rank * g.k_blocks_per_split;
                int kbn = g.k_blocks_total - kb0;
                if (kbn > g.k_blocks_per_split) kbn = g.k_blocks_per_split;
                if (op > 0) {
                    grid_wait(P.bar, op);
                    asm volatile("fence.proxy.async;" ::: "memory");    // generic-proxy writes of other SMs -> TMA reads
                }
                chain_ts(P.ts, op, 0);
                for (int t = cluster_id; t < P.m_tiles[gi]; t += n_clusters) {
                    for (int kb = 0; kb < kbn; ++kb, ++it) {
                        const int s = it % P.stages;
                        mbar_wait(&ctl->empty[s], ((it / P.stages) & 1) ^ 1);
                        uint8_t* sb = smem + static_cast<size_t>(s) * stage_bytes + a_bytes;
                        mbar_expect_tx(&ctl->full_b[s], b_bytes);
                        tma_load_2d(sb, &P.tb[gi], &ctl->full_b[s], (kb0 + kb) * kBlockK, 0, kEvictLast);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        uint32_t it = 0, acc_it = 0;
        for (int op = 0; op < P.n_ops; ++op) {
            if (P.kind[op] != 0) continue;
            const int gi = P.idx[op];
            const GemmArgs& g = P.g[gi];
            const int kb0 = rank * g.k_blocks_per_split;
            int kbn = g.k_blocks_total - kb0;
            if (kbn > g.k_blocks_per_split) kbn = g.k_blocks_per_split;
            const uint32_t idesc = umma_idesc_bf16_major(kBlockM, P.bn, g.a_mn ? 1u : 0u, 0u);
            const uint64_t a_step = g.a_mn ? 128 : 2;
            for (int t = cluster_id; t < P.m_tiles[gi]; t += n_clusters, ++acc_it) {
                const uint32_t buf = acc_it & 1;
                mbar_wait(&ctl->acc_empty[buf], ((acc_it >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * P.acc_cols;
                for (int kb = 0; kb < kbn; ++kb, ++it) {
                    const int s = it % P.stages;
                    const uint32_t par = (it / P.stages) & 1;
                    mbar_wait(&ctl->full_a[s], par);
                    mbar_wait(&ctl->full_b[s], par);
                    tc_fence_after();
                    if (kb == 0 && t == cluster_id && lane == 0) chain_ts(P.ts, op, 3);
                    if (elect_one()) {
                        const uint32_t sa = smem_u32(smem + static_cast<size_t>(s) * stage_bytes);
                        const uint64_t da = a_desc(sa, g.a_mn);
                        const uint64_t db = umma_desc_k_sw128(sa + a_bytes);
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k)
                            umma_bf16(d_tmem, da + a_step * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        umma_commit(&ctl->empty[s]);
                        if (kb == kbn - 1) umma_commit(&ctl->acc_full[buf]);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ===================== epilogue warps 2..9 =====================
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int m_local = quarter * 32 + lane;
        const int ep_tid = threadIdx.x - 64;
        const int n_chunks = P.bn >> 4;
        const uint32_t staging_u32 = smem_u32(staging);
        uint32_t free_addr[kChainCluster];
#pragma unroll
        for (int q = 0; q < kChainCluster; ++q) free_addr[q] = dsmem_addr(smem_u32(&ctl->sent_free), q);
        pdl_wait();
        uint32_t acc_it = 0;
        for (int op = 0; op < P.n_ops; ++op) {
            // One counter serves every barrier of the launch, so no CTA may arrive for op before all have arrived for
            // op - 1 (a cluster without a tile in two consecutive GEMMs would otherwise count twice and let a consumer start
            // early).  For CTAs with work this wait is long over by the time their accumulator is ready.
            if (op > 0) {
                if (ep_tid == 0) grid_wait(P.bar, op);
                ep_bar();
            }
            if (P.kind[op] == 0) {
                const int gi = P.idx[op];
                const GemmArgs& g = P.g[gi];
                EpiCtx e{CounterRng(g.seed), static_cast<uint32_t>(fminf(g.drop_p, 0.999999f) * 4294967296.0f),
                         g.drop_p > 0.f ? 1.f / (1.f - g.drop_p) : 1.f, 0.f, false};
                for (int t = cluster_id; t < P.m_tiles[gi]; t += n_clusters, ++acc_it) {
                    const uint32_t buf = acc_it & 1;
                    const int m_glob = t * kBlockM + m_local;
                    e.row_ok = m_glob < g.M;
                    e.bias_m = (g.bias && g.bias_along_m && e.row_ok) ? g.bias[m_glob] : 0.f;
                    const uint32_t t_lane = tmem_base + buf * P.acc_cols + (static_cast<uint32_t>(quarter * 32) << 16);
                    // while the MMAs run: what the first owned chunk needs besides the accumulator (dropout bits, mask and
                    // residual operands — their producers are complete, see the wait at the top of the op)
                    ChunkPre pre;
                    pre.have = false;
                    {
                        const int k_first = half * kChainCluster + static_cast<int>(rank);
                        if (half < P.local_chunks && k_first < n_chunks && k_first * 16 < g.N &&
                            (g.drop_p > 0.f || g.mask_src || g.addend))
                            epilogue_prefetch(g, e, m_glob, k_first * 16, pre);
                    }
                    mbar_wait(&ctl->acc_full[buf], (acc_it >> 1) & 1);
                    tc_fence_after();
                    if (ep_tid == 0 && t == cluster_id) chain_ts(P.ts, op, 1);
                    // the peers have read what this CTA sent them for the previous tile
                    if (acc_it > 0) mbar_wait(&ctl->sent_free, (acc_it - 1) & 1);
                    if (ep_tid == 0) {
                        // bytes this CTA is about to receive: its chunks of this tile from the three peers
                        int own = 0;
                        for (int k = static_cast<int>(rank); k < n_chunks; k += kChainCluster) ++own;
                        mbar_expect_tx(&ctl->recv, static_cast<uint32_t>(own) * (kChainCluster - 1) * kBlockM * 64u);
                    }
                    for (int k = half; k < n_chunks; k += 2) {
                        const uint32_t owner = static_cast<uint32_t>(k % kChainCluster);
                        if (owner == rank) continue;
                        uint32_t r[16];
                        tmem_ld_32x16(t_lane + k * 16, r);
                        tmem_ld_wait();
                        const uint32_t si = rank < owner ? rank : rank - 1;           // my slot among the owner's sources
                        // staging block of (source, chunk): [4 column quads][128 rows][16 B] — a warp's store instruction
                        // covers 512 contiguous bytes (the row-major form with one 16-byte piece per row every 80 bytes
                        // moved the same data as 32 separate remote writes per instruction)
                        const uint32_t local = staging_u32 + (si * P.local_chunks + (k / kChainCluster)) * (kBlockM * 64u) + m_local * 16u;
                        const uint32_t remote = dsmem_addr(local, owner);
                        const uint32_t recv_remote = dsmem_addr(smem_u32(&ctl->recv), owner);
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            dsmem_st_async_f32x4(remote + (j / 4) * (kBlockM * 16u), recv_remote, __uint_as_float(r[j]),
                                                 __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                    }
                    if (ep_tid == 0 && t == cluster_id) chain_ts(P.ts, op, 4);
                    if (ep_tid == 0 && t == cluster_id) chain_ts(P.ts, op, 5);
                    mbar_wait(&ctl->recv, acc_it & 1);
                    if (ep_tid == 0 && t == cluster_id) chain_ts(P.ts, op, 6);
                    float ss = 0.f;
                    for (int kl = half; kl < P.local_chunks; kl += 2) {
                        const int k = kl * kChainCluster + static_cast<int>(rank);
                        if (k >= n_chunks) break;
                        uint32_t r[16];
                        tmem_ld_32x16(t_lane + k * 16, r);
                        tmem_ld_wait();
                        float v[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
#pragma unroll
                        for (int si = 0; si < kChainCluster - 1; ++si) {
                            const float4* p = reinterpret_cast<const float4*>(staging) +
                                              (static_cast<size_t>(si) * P.local_chunks + kl) * (kBlockM * 4) + m_local;
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float4 x4 = p[q * kBlockM];
                                v[4 * q] += x4.x; v[4 * q + 1] += x4.y; v[4 * q + 2] += x4.z; v[4 * q + 3] += x4.w;
                            }
                        }
                        if (k * 16 < g.N) epilogue_chunk(g, e, v, m_glob, k * 16, ss, (kl == half && pre.have) ? &pre : nullptr);
                    }
                    if (ep_tid == 0 && t == cluster_id) chain_ts(P.ts, op, 7);
                    // accumulator and staging of this tile are free again
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(&ctl->acc_empty[buf]);
#pragma unroll
                        for (int q = 0; q < kChainCluster; ++q)
                            if (q != static_cast<int>(rank)) mbar_arrive_remote_relaxed(free_addr[q]);
                    }
                }
            } else {
                const ChainLn& L = P.ln[P.idx[op]];
                if (ep_tid == 0) chain_ts(P.ts, op, 0);
                for (int row = blockIdx.x; row < L.rows; row += gridDim.x) ln_row(L, row, ctl->red, ep_tid);
            }
            if (ep_tid == 0) chain_ts(P.ts, op, 2);
            if (op + 1 < P.n_ops) {
                // this CTA's share of op is stored: publish it and count the arrival (the CTA barrier orders every thread's
                // stores before the one release below)
                ep_bar();
                if (ep_tid == 0) {
                    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(P.bar), "l"(1ull) : "memory");
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();             // no CTA leaves while a peer may still write into its shared memory
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * P.acc_cols);
    }
    if (threadIdx.x == 0) chain_ts(P.ts, 8, 1);
    if (threadIdx.x == 0 && P.n_ops > 1) {
        // every thread of this CTA is past its last grid_wait; the CTA that leaves last re-arms the counters
        const unsigned long long prev = atomicAdd(P.bar + 1, 1ull);
        if (prev == gridDim.x - 1) {
            P.bar[0] = 0ull;
            P.bar[1] = 0ull;
            __threadfence();
        }
    }
}

static uint32_t chain_pow2_cols(int n) {
    uint32_t c = 32;
    while (static_cast<int>(c) < n) c <<= 1;
    return c;
}

// Why a chain cannot run as one kernel (nullptr: it can).
static const char* chain_reject(const pb2_chain_op* ops, int n_ops) {
    if (!ops || n_ops <= 0 || n_ops > PB2_CHAIN_MAX_OPS) return "op count out of range";
    int n_gemm = 0, n_ln = 0, tokens = -1;
    for (int i = 0; i < n_ops; ++i) {
        if (ops[i].kind == 0) {
            const pb2_gemm_desc& d = ops[i].gemm;
            if (++n_gemm > kChainMaxGemms) return "too many GEMMs";
            if (d.M <= 0 || d.N <= 0 || d.K <= 0) return "non-positive GEMM shape";
            if (tokens >= 0 && d.N != tokens) return "GEMMs of a chain must share the token count";
            tokens = d.N;
            if (d.N > 128) return "more than 128 tokens";
            if ((d.lda % 8) || (d.ldb % 8)) return "lda / ldb not multiples of 8";
            if ((reinterpret_cast<uintptr_t>(d.a) & 15) || (reinterpret_cast<uintptr_t>(d.b) & 15)) return "unaligned operand";
            if (d.accumulate || d.sumsq || (d.bias && !d.bias_along_m)) return "epilogue not available in a chain";
            const int kb = (d.K + kBlockK - 1) / kBlockK, per = (kb + kChainCluster - 1) / kChainCluster;
            if (kb - (kChainCluster - 1) * per < 1) return "reduction too short for a four-way split";
        } else if (ops[i].kind == 1) {
            const pb2_chain_ln& l = ops[i].ln;
            if (++n_ln > kChainMaxLns) return "too many LayerNorms";
            if (l.rows <= 0 || l.E <= 0 || (l.E % 8) || l.E > 8 * 256 * kLnVecs) return "LayerNorm width";
        } else {
            return "unknown op kind";
        }
    }
    if (n_gemm == 0) return "no GEMM in the chain";
    return nullptr;
}

extern "C" void pb2_debug_chain_timestamps(unsigned long long* buf) { g_chain_ts = buf; }

extern "C" int pb2_chain_supported(const pb2_chain_op* ops, int n_ops) { return chain_reject(ops, n_ops) == nullptr ? 1 : 0; }

extern "C" int pb2_chain_run(const pb2_chain_op* ops, int n_ops, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (const char* why = chain_reject(ops, n_ops)) { set_error("pb2_chain_run: %s", why); return -1; }
    if (!workspace || workspace_bytes < PB2_GEMM_COUNTER_BYTES) {
        set_error("pb2_chain_run: needs the GEMM workspace (its last %u bytes hold the grid-barrier counters)", PB2_CHAIN_BARRIER_BYTES);
        return -1;
    }
    ChainParams P;
    memset(&P, 0, sizeof(P));
    P.n_ops = n_ops;
    P.bar = reinterpret_cast<unsigned long long*>(static_cast<char*>(workspace) + workspace_bytes - PB2_CHAIN_BARRIER_BYTES);
    P.ts = g_chain_ts;
    int ng = 0, nl = 0, tokens = 0;
    for (int i = 0; i < n_ops; ++i)
        if (ops[i].kind == 0) tokens = ops[i].gemm.N;
    const int bn = ((tokens + 15) / 16) * 16;
    P.bn = bn;
    P.acc_cols = static_cast<int>(chain_pow2_cols(bn));
    const int n_chunks = bn / 16;
    P.local_chunks = (n_chunks + kChainCluster - 1) / kChainCluster;
    for (int i = 0; i < n_ops; ++i) {
        P.kind[i] = ops[i].kind;
        if (ops[i].kind == 0) {
            const pb2_gemm_desc* d = &ops[i].gemm;
            GemmArgs& g = P.g[ng];
            g.M = d->M; g.N = d->N; g.K = d->K;
            g.block_n = bn;
            g.k_blocks_total = (d->K + kBlockK - 1) / kBlockK;
            g.k_blocks_per_split = (g.k_blocks_total + kChainCluster - 1) / kChainCluster;
            g.splits = kChainCluster;
            g.local_chunks = P.local_chunks;
            g.cluster_reduce = 1;
            g.out = d->out; g.ld_out = d->ld_out; g.out_f32 = d->out_f32; g.transpose_out = d->transpose_out;
            g.bias = d->bias; g.bias_along_m = d->bias_along_m; g.relu = d->relu;
            g.drop_p = d->drop_p; g.seed = d->seed; g.stream = d->rng_stream;
            g.mask_src = static_cast<const __nv_bfloat16*>(d->mask_src); g.mask_scale = d->mask_scale;
            g.addend = static_cast<const __nv_bfloat16*>(d->addend);
            g.alpha = d->alpha == 0.f ? 1.f : d->alpha;
            g.a_static = 1;
            g.a_mn = d->a_mn_major ? 1 : 0;
            if (g.a_mn) {
                if (make_tmap_bf16_2d(&P.ta[ng], d->a, d->K, d->M, d->lda, kBlockK, 64)) return -1;
            } else if (make_tmap_bf16_2d(&P.ta[ng], d->a, d->M, d->K, d->lda, kBlockM, kBlockK)) return -1;
            if (make_tmap_bf16_2d(&P.tb[ng], d->b, d->N, d->K, d->ldb, bn, kBlockK)) return -1;
            P.m_tiles[ng] = (d->M + kBlockM - 1) / kBlockM;
            P.idx[i] = ng++;
        } else {
            const pb2_chain_ln& l = ops[i].ln;
            ChainLn& L = P.ln[nl];
            L.x = static_cast<const __nv_bfloat16*>(l.x); L.h = static_cast<const __nv_bfloat16*>(l.h);
            L.gamma = l.gamma; L.beta = l.beta; L.y = static_cast<__nv_bfloat16*>(l.y);
            L.mean = l.mean; L.rstd = l.rstd; L.rows = l.rows; L.E = l.E; L.eps = l.eps;
            P.idx[i] = nl++;
        }
    }
    const size_t stage_bytes = static_cast<size_t>(kBlockM + bn) * kBlockK * 2;
    const size_t staging = static_cast<size_t>(kChainCluster - 1) * P.local_chunks * kBlockM * 64;
    const size_t fixed = staging + sizeof(ChainCtl) + 1024;
    int stages = static_cast<int>((227 * 1024 - fixed) / stage_bytes);
    if (stages > kMaxStages) stages = kMaxStages;
    {
        static int cap = -1;
        if (cap < 0) { const char* e = getenv("PB2_CHAIN_STAGES"); cap = e ? atoi(e) : 0; }
        if (cap > 0 && cap < stages) stages = cap;
    }
    if (stages < 2) { set_error("pb2_chain_run: shared memory holds fewer than two stages"); return -1; }
    P.stages = stages;
    {
        static int pf = -1;
        if (pf < 0) { const char* e = getenv("PB2_CHAIN_PREFETCH"); pf = e ? atoi(e) : 16; }
        P.pf_ahead = pf;
    }
    const size_t smem = stages * stage_bytes + fixed;
    if (ensure_max_smem(reinterpret_cast<const void*>(chain_kernel), 227 * 1024)) return -1;

    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.blockDim = dim3(kChainThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kChainCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;

    // clusters that are resident at once (cached per device): every CTA must be, the grid barrier spins
    static int max_clusters[64] = {0};
    int dev = 0;
    PB2_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("pb2_chain_run: device ordinal %d", dev); return -1; }
    if (max_clusters[dev] == 0) {
        cfg.gridDim = dim3(kChainCluster * 37);
        cfg.dynamicSmemBytes = 227 * 1024 - 1024;
        cfg.numAttrs = 1;
        int n = 0;
        PB2_CUDA(cudaOccupancyMaxActiveClusters(&n, chain_kernel, &cfg));
        if (n < 1) { set_error("pb2_chain_run: no resident cluster of %d CTAs", kChainCluster); return -1; }
        static int cap = -1;
        if (cap < 0) { const char* e = getenv("PB2_CHAIN_CLUSTERS"); cap = e ? atoi(e) : 0; }
        if (cap > 0 && cap < n) n = cap;
        max_clusters[dev] = n;
        cfg.dynamicSmemBytes = smem;
    }
    int most_tiles = 1;
    for (int gi = 0; gi < ng; ++gi) most_tiles = P.m_tiles[gi] > most_tiles ? P.m_tiles[gi] : most_tiles;
    int clusters = max_clusters[dev] < most_tiles ? max_clusters[dev] : most_tiles;
    // LayerNorm rows are spread over CTAs: keep enough of them
    for (int li = 0; li < nl; ++li) {
        const int want = (P.ln[li].rows + kChainCluster - 1) / kChainCluster;
        const int c = want < max_clusters[dev] ? want : max_clusters[dev];
        if (c > clusters) clusters = c;
    }
    cfg.gridDim = dim3(kChainCluster * clusters);
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    PB2_CUDA(cudaLaunchKernelEx(&cfg, chain_kernel, P));
    PB2_LAUNCHED("chain_kernel");
    return 0;
}

}  // namespace pb2
