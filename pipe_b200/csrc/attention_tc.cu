// Tensor-core attention core for short sequences (S <= 128 while the tiles fit shared memory): the five small matrix products of
// attention forward + backward on warp-level MMA (bf16 in, fp32 accumulate), everything else —
// softmax, dropout mask, the dS recurrence — on the scores held in shared memory.
//
// One (sequence, head) pair is owned by a thread-block cluster of 1, 2 or 4 CTAs for the whole pass.
// With more than one CTA the head dimension is split: every CTA gathers its D / cs columns of Q, K, V
// (and dO) from the packed [T, 3E] projection output into zero-padded shared tiles, computes the
// partial S x S score products over those columns, and pushes them into its peers' shared memory
// (DSMEM); after one cluster barrier every CTA holds all partials, sums them in the same order (so
// the softmax is bit-identical across the cluster), and finishes the products whose outputs are its
// own D / cs columns.  The S x S score matrix never leaves the SMs, and results are scattered
// straight back into the packed layouts the neighbouring GEMMs read.
//
// Why split a head at all: a micro-batch of the tutorial configuration is 3 sequences x 16 heads =
// 48 independent problems on 148 SMs, and each problem is bound by instruction issue of its one CTA
// (ncu: issue slot busy ~100 %, 7.8 k instructions per warp), not by math or memory.  Four CTAs per
// head put 192 CTAs on the machine and cut every D-proportional phase by four.
//
// Why warp-level MMA and not tcgen05 here: at S = 25 the whole problem is a 32 x 32 x 256 product
// per head; a UMMA tile is 128 rows and needs TMA-swizzled operands, TMEM allocation and an mbarrier
// pipeline — fixed costs that exceed the work.  These products are latency-, not throughput-bound
// (attention is <= 0.2 % of the model's FLOPs).
// The SIMT kernels in attention.cu remain the fallback when the tiles do not fit (S = 128 with
// head_dim > 64).
//
// Reference counterpart: nn.MultiheadAttention core (main.py:148).  SURVEY.md K3.
#include <mma.h>
#include <stdlib.h>
#include <string.h>

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
    int cs;                          // CTAs per (sequence, head): the cluster size
    int Dc;                          // head-dim columns per CTA (D / cs)
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
    const int dpad = D + 8, vec_per_row = D / 8;              // 2 .. 32 vectors per row
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

// One or two score-shaped products at once, the tiles of both spread over all warps:
//   C0[SP x SP] = A0[SP x D] * B0[SP x D]^T   and (if A1)   C1 = A1 * B1^T
// fp32 results with ld = SP + 8; operands row-major bf16 with ld = D + 8.
__device__ __forceinline__ void tc_abt(const __nv_bfloat16* A0, const __nv_bfloat16* B0, float* C0,
                                       const __nv_bfloat16* A1, const __nv_bfloat16* B1, float* C1, int SP, int D,
                                       int rows_a = 0) {
    // rows_a: rows of A (and of C) to compute, default all SP; B always has SP rows (C has SP columns)
    const int warp = threadIdx.x >> 5, tiles = SP / 16, dpad = D + 8, cpad = SP + 8;
    const int per = (rows_a ? rows_a / 16 : tiles) * tiles, total = A1 ? 2 * per : per;
    for (int t = warp; t < total; t += kTcWarps) {
        const int second = t >= per;
        const int r = t - second * per, ti = r / tiles, tj = r - ti * tiles;
        const __nv_bfloat16* A = (second ? A1 : A0) + ti * 16 * dpad;
        const __nv_bfloat16* B = (second ? B1 : B0) + tj * 16 * dpad;
        // two accumulators break the MMA dependency chain along K
        FragC acc0, acc1;
        wmma::fill_fragment(acc0, 0.f);
        wmma::fill_fragment(acc1, 0.f);
        for (int k = 0; k < D; k += 32) {
            FragA a0, a1;
            FragBT b0, b1;
            wmma::load_matrix_sync(a0, A + k, dpad);
            wmma::load_matrix_sync(b0, B + k, dpad);
            if (k + 16 < D) {
                wmma::load_matrix_sync(a1, A + k + 16, dpad);
                wmma::load_matrix_sync(b1, B + k + 16, dpad);
            }
            wmma::mma_sync(acc0, a0, b0, acc0);
            if (k + 16 < D) wmma::mma_sync(acc1, a1, b1, acc1);
        }
#pragma unroll
        for (int e = 0; e < acc0.num_elements; ++e) acc0.x[e] += acc1.x[e];
        wmma::store_matrix_sync((second ? C1 : C0) + ti * 16 * cpad + tj * 16, acc0, cpad, wmma::mem_row_major);
    }
}

// Push `count` floats (a multiple of 4, 16-byte aligned) of this CTA's shared memory to the same
// location in every other CTA of the cluster.  Each 16-byte store completes 16 bytes of the transaction the
// receiver armed on its `recv` barrier (st.async): the receiver waits for its bytes, nobody executes a
// release / acquire cluster barrier (a MEMBAR.GPU per thread with the remote stores outstanding).
__device__ __forceinline__ void tc_push(const float* local, int count, int cs, uint32_t rank, uint64_t* recv) {
    const uint32_t base = smem_u32(local), bar = smem_u32(recv);
    for (int v = threadIdx.x * 4; v < count; v += kTcThreads * 4) {
        const float4 x = *reinterpret_cast<const float4*>(local + v);
#pragma unroll 4
        for (int peer = 0; peer < cs; ++peer)
            if (peer != static_cast<int>(rank))
                dsmem_st_async_f32x4(dsmem_addr(base + v * 4u, peer), dsmem_addr(bar, peer), x.x, x.y, x.z, x.w);
    }
}
__device__ __forceinline__ void cluster_arrive_relaxed() {
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}

// Softmax over the first S columns of the first S rows of the scores (scaled by `scale`).  The scores
// arrive as `nslot` partial matrices `slot` floats apart (one per cluster rank) that are summed on
// the way in; the probabilities are written to the first one.  Eight lanes share a row, so 32 rows
// are in flight at once (a warp-per-row loop would walk the rows one latency chain after another).
// SP <= 128: at most 16 columns per lane.
template <int NSLOT>
__device__ __forceinline__ void tc_softmax(float* sc, int slot, int S, int SP, float scale, int causal) {
    const int sub = threadIdx.x & 7, cpad = SP + 8;
    for (int i = threadIdx.x >> 3; i < SP; i += kTcThreads / 8) {
        float* r = sc + i * cpad;
        const int lim = i < S ? (causal ? i + 1 : S) : 0;
        float x[16];
        float mx = -INFINITY;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = sub + u * 8;
            x[u] = -INFINITY;
            if (j < lim) {
                float t = r[j];
#pragma unroll
                for (int s = 1; s < NSLOT; ++s) t += r[s * slot + j];
                x[u] = t * scale;
            }
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

// Forward: softmax of `rows` score rows starting at global row `row0`, times the dropout mask, written
// as bf16 *over the fp32 rows they came from* (row pitch bpad = 2 * cpad bf16 elements, the same bytes as
// the fp32 pitch).  Eight lanes share a row and hold all of it in registers, so once they have all read
// (one __syncwarp) the row can be overwritten; no second pass over shared memory and no extra buffer.
template <int NSLOT>
__device__ __forceinline__ void tc_softmax_drop(float* sc, int slot, int row0, int rows, const AttnTc& a, int n, int h) {
    const int S = a.S, SP = a.SP, sub = threadIdx.x & 7, cpad = SP + 8, bpad = 2 * cpad;
    const CounterRng rng(a.seed);
    const uint32_t thr = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
    const float inv_keep = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
    __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(sc);
    for (int rl = threadIdx.x >> 3; rl < rows; rl += kTcThreads / 8) {
        const int i = row0 + rl;
        const float* r = sc + rl * cpad;
        const int lim = i < S ? (a.causal ? i + 1 : S) : 0;
        float x[16];
        float mx = -INFINITY;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = sub + u * 8;
            x[u] = -INFINITY;
            if (j < lim) {
                float t = r[j];
#pragma unroll
                for (int q = 1; q < NSLOT; ++q) t += r[q * slot + j];
                x[u] = t * a.scale;
            }
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
        __syncwarp();              // every lane of the row has its fp32 scores: the row may be overwritten
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = sub + u * 8;
            if (j < SP) {
                float w = x[u] * inv;
                if (j < lim) w *= tc_keep(rng, a, thr, inv_keep, n, h, i, j);
                pb[rl * bpad + j] = __float2bfloat16(w);
            }
        }
    }
}

// out rows (< S) of a [16 x 16] fp32 tile held in this warp's staging area -> bf16 global;
// `row_stride` is the distance between consecutive sequence positions in elements
__device__ __forceinline__ void tc_emit_tile(const float* stage, __nv_bfloat16* dst_row0, size_t row_stride, int row0, int S) {
    const int lane = threadIdx.x & 31;
    // 16 rows x 8 column pairs = 128 pairs, 4 per lane
    const int c2 = lane & 7, r_lo = lane >> 3;
    __nv_bfloat16* dst = dst_row0 + r_lo * row_stride + c2 * 2;
    const float* src = stage + r_lo * 24 + c2 * 2;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (row0 + r_lo + q * 4 < S) {
            const float2 v = *reinterpret_cast<const float2*>(src + q * 4 * 24);
            *reinterpret_cast<__nv_bfloat162*>(dst + q * 4 * row_stride) = __floats2bfloat162_rn(v.x, v.y);
        }
    }
}

template <int CS>
__global__ void __launch_bounds__(kTcThreads)
attn_tc_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, const AttnTc a) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    attn_ts(0);
    constexpr int cs = CS;
    const uint32_t rank = cs > 1 ? cluster_ctarank() : 0u;
    extern __shared__ __align__(128) uint8_t smem[];
    // Query rows are processed in blocks of RB <= 64: the score block [RB x SP] is the only S^2-sized
    // buffer, so S = 128 needs ~100 KB instead of ~170 KB and two CTAs share an SM (256 CTAs = one wave).
    const int S = a.S, SP = a.SP, Dc = a.Dc, dpad = Dc + 8, cpad = SP + 8, bpad = 2 * cpad;
    const int RB = SP < 64 ? SP : 64, slot = RB * cpad;
    __nv_bfloat16* q = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* k = q + SP * dpad;
    __nv_bfloat16* v = k + SP * dpad;
    float* sc = reinterpret_cast<float*>(v + SP * dpad);                     // cs partial score blocks
    const __nv_bfloat16* pb = reinterpret_cast<const __nv_bfloat16*>(sc);    // probabilities, written over slot 0
    float* stage = sc + cs * slot + (threadIdx.x >> 5) * (16 * 24);
    // peers' partial scores of the current row block have landed (transaction bytes); lives behind the staging tiles
    uint64_t& recv = *reinterpret_cast<uint64_t*>(sc + cs * slot + kTcWarps * (16 * 24));
    if (cs > 1) {
        if (threadIdx.x == 0) {
            mbar_init(&recv, 1);
            fence_mbar_init();
            mbar_expect_tx(&recv, static_cast<uint32_t>(cs - 1) * slot * 4u);
        }
        cluster_arrive_relaxed();            // waited for right before the first DSMEM store: every peer is running (and armed) by then
    }
    const int pair = blockIdx.x / cs;
    const int n = pair / a.H, h = pair - n * a.H;
    const long long ld = 3LL * a.E;
    const int col0 = h * a.D + static_cast<int>(rank) * Dc;
    const int warp = threadIdx.x >> 5, ct = Dc / 16;
    const size_t row_stride = static_cast<size_t>(a.NB) * a.E;

    tc_gather(qkv + col0, qkv + a.E + col0, qkv + 2 * a.E + col0, nullptr, ld, 0, q, k, v, nullptr, 3, a.NB, S, SP, Dc, n);
    __syncthreads();
    attn_ts(1);
    for (int rb0 = 0; rb0 < SP; rb0 += RB) {
        const int rows = SP - rb0 < RB ? SP - rb0 : RB;          // the last block may be short (SP = 80, 96, 112)
        tc_abt(q + rb0 * dpad, k, sc + rank * slot, nullptr, nullptr, nullptr, SP, Dc, rows);
        __syncthreads();
        if (cs > 1) {
            if (rb0 == 0) cluster_wait();
            tc_push(sc + rank * slot, slot, cs, rank, &recv);
            mbar_wait(&recv, (rb0 / RB) & 1);
            __syncthreads();                 // slower threads are still reading this CTA's own partial (it is overwritten next)
        }
        attn_ts(2);
        tc_softmax_drop<CS>(sc, slot, rb0, rows, a, n, h);
        __syncthreads();
        attn_ts(4);

        // O[block rows, own columns] = P V: each warp keeps up to four output tiles in flight
        const int total = (rows / 16) * ct;
        for (int t0 = warp; t0 < total; t0 += kTcWarps * 4) {
            FragC acc[4];
            int ti[4], tc[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                wmma::fill_fragment(acc[u], 0.f);
                const int t = t0 + u * kTcWarps;
                ti[u] = t / ct;
                tc[u] = t - ti[u] * ct;
            }
            for (int kk = 0; kk < SP; kk += 16) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (t0 + u * kTcWarps < total) {
                        FragA pa;
                        FragB vb;
                        wmma::load_matrix_sync(pa, pb + ti[u] * 16 * bpad + kk, bpad);
                        wmma::load_matrix_sync(vb, v + kk * dpad + tc[u] * 16, dpad);
                        wmma::mma_sync(acc[u], pa, vb, acc[u]);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (t0 + u * kTcWarps < total) {
                    const int row = rb0 + ti[u] * 16;
                    wmma::store_matrix_sync(stage, acc[u], 24, wmma::mem_row_major);
                    __syncwarp();
                    tc_emit_tile(stage, out + (static_cast<size_t>(row) * a.NB + n) * a.E + col0 + tc[u] * 16, row_stride, row, S);
                    __syncwarp();
                }
            }
        }
        if (rb0 + RB < SP) {
            if (cs > 1) {
                // the peers push the next block's scores into sc: everybody is done with this block first, and the
                // barrier is armed again before that (execution barrier only: the data is published by the transaction)
                if (threadIdx.x == 0) mbar_expect_tx(&recv, static_cast<uint32_t>(cs - 1) * slot * 4u);
                cluster_arrive_relaxed();
                cluster_wait();
            } else {
                __syncthreads();                 // the next block's scores overwrite the probabilities
            }
        }
    }
    attn_ts(5);
}

template <int CS>
__global__ void __launch_bounds__(kTcThreads)
attn_tc_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dout,
                   __nv_bfloat16* __restrict__ dqkv, const AttnTc a) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    attn_ts(8);
    constexpr int cs = CS;
    const uint32_t rank = cs > 1 ? cluster_ctarank() : 0u;
    extern __shared__ __align__(128) uint8_t smem[];
    const int S = a.S, SP = a.SP, Dc = a.Dc, dpad = Dc + 8, cpad = SP + 8, slot = SP * cpad;
    __nv_bfloat16* q = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* k = q + SP * dpad;
    __nv_bfloat16* v = k + SP * dpad;
    __nv_bfloat16* dO = v + SP * dpad;
    float* P = reinterpret_cast<float*>(dO + SP * dpad);       // cs partial score matrices, then the probabilities in the first
    float* dP = P + cs * slot;                                 // cs partial dO V^T matrices
    // The bf16 matrices the last three products read are written over the fp32 rows they are derived
    // from (same row pitch in bytes, so bpad = 2 * cpad bf16 elements): S = 128 with D = 64 then
    // fits the 227 KB of shared memory.
    __nv_bfloat16* Pd = reinterpret_cast<__nv_bfloat16*>(P);                  // dropped-out probabilities
    __nv_bfloat16* dS = reinterpret_cast<__nv_bfloat16*>(dP);                 // score gradients
    const int bpad = 2 * cpad;
    float* stage = dP + cs * slot + (threadIdx.x >> 5) * (16 * 24);
    // peers' partial P and dP have landed (transaction bytes); lives behind the staging tiles
    uint64_t& recv = *reinterpret_cast<uint64_t*>(dP + cs * slot + kTcWarps * (16 * 24));
    if (cs > 1) {
        if (threadIdx.x == 0) {
            mbar_init(&recv, 1);
            fence_mbar_init();
            mbar_expect_tx(&recv, 2u * static_cast<uint32_t>(cs - 1) * slot * 4u);
        }
        cluster_arrive_relaxed();
    }
    const int pair = blockIdx.x / cs;
    const int n = pair / a.H, h = pair - n * a.H;
    const long long ld = 3LL * a.E;
    const int col0 = h * a.D + static_cast<int>(rank) * Dc;

    tc_gather(qkv + col0, qkv + a.E + col0, qkv + 2 * a.E + col0, dout + col0, ld, a.E, q, k, v, dO, 4, a.NB, S, SP, Dc, n);
    __syncthreads();
    attn_ts(9);
    // raw scores and the gradient w.r.t. the dropped-out probabilities (partial over this CTA's columns)
    tc_abt(q, k, P + rank * slot, dO, v, dP + rank * slot, SP, Dc);
    __syncthreads();
    if (cs > 1) {
        cluster_wait();
        tc_push(P + rank * slot, slot, cs, rank, &recv);
        tc_push(dP + rank * slot, slot, cs, rank, &recv);
        mbar_wait(&recv, 0);
        __syncthreads();                     // slower threads are still reading this CTA's own partials (overwritten next)
    }
    attn_ts(10);
    tc_softmax<CS>(P, slot, S, SP, a.scale, a.causal);
    __syncthreads();
    attn_ts(11);

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
                    float g = dP[i * cpad + j];
#pragma unroll
                    for (int s = 1; s < CS; ++s) g += dP[s * slot + i * cpad + j];
                    dpm[u] = g * ks;
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
    attn_ts(12);

    // dV = Pd^T dO,  dQ = dS K,  dK = dS^T Q  (own columns) — 3 x (SP/16) x (Dc/16) output tiles, four in flight per warp
    const int warp = threadIdx.x >> 5, rt = SP / 16, ct = Dc / 16, per = rt * ct, total = 3 * per;
    const size_t row_stride = static_cast<size_t>(a.NB) * ld;
    for (int t0 = warp; t0 < total; t0 += kTcWarps * 4) {
        FragC acc[4];
        int which[4], ti[4], tc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            wmma::fill_fragment(acc[u], 0.f);
            const int t = t0 + u * kTcWarps;
            which[u] = (t >= per) + (t >= 2 * per);
            const int r = t - which[u] * per;
            ti[u] = r / ct;
            tc[u] = r - ti[u] * ct;
        }
        for (int kk = 0; kk < SP; kk += 16) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (t0 + u * kTcWarps < total) {
                    FragB bm;
                    if (which[u] == 1) {         // dQ: A = dS (row-major), B = K
                        FragA am;
                        wmma::load_matrix_sync(am, dS + ti[u] * 16 * bpad + kk, bpad);
                        wmma::load_matrix_sync(bm, k + kk * dpad + tc[u] * 16, dpad);
                        wmma::mma_sync(acc[u], am, bm, acc[u]);
                    } else {                     // dV: A = Pd^T, B = dO;  dK: A = dS^T, B = Q
                        FragAT at;
                        wmma::load_matrix_sync(at, (which[u] == 0 ? Pd : dS) + kk * bpad + ti[u] * 16, bpad);
                        wmma::load_matrix_sync(bm, (which[u] == 0 ? dO : q) + kk * dpad + tc[u] * 16, dpad);
                        wmma::mma_sync(acc[u], at, bm, acc[u]);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (t0 + u * kTcWarps < total) {
                wmma::store_matrix_sync(stage, acc[u], 24, wmma::mem_row_major);
                __syncwarp();
                const int c = (which[u] == 0 ? 2 * a.E : which[u] == 1 ? 0 : a.E) + col0 + tc[u] * 16;
                tc_emit_tile(stage, dqkv + (static_cast<size_t>(ti[u] * 16) * a.NB + n) * ld + c, row_stride, ti[u] * 16, S);
                __syncwarp();
            }
        }
    }
    attn_ts(13);
}

size_t attn_tc_smem(int SP, int Dc, int cs, bool bwd) {
    const size_t dpad = Dc + 8, cpad = SP + 8;
    size_t b = (bwd ? 4 : 3) * SP * dpad * 2;             // Q K V (dO), this CTA's columns
    const size_t rows = bwd ? SP : (SP < 64 ? SP : 64);    // forward walks the query rows in blocks of <= 64
    b += (bwd ? 2 : 1) * cs * rows * cpad * 4;             // fp32 partial score matrices, one per cluster rank
                                                           // (the bf16 copies alias the first)
    b += kTcWarps * 16 * 24 * 4;                           // per-warp output staging
    b += 16;                                               // the exchange's mbarrier
    return b + 128;
}

bool attn_tc_supported(int S, int D) {
    if (S > 128 || D % 16 != 0 || D > 256) return false;
    const int SP = (S + 15) / 16 * 16;
    return attn_tc_smem(SP, D, 1, true) <= 227 * 1024;     // e.g. S = 128 needs D <= 64; S <= 64 any D <= 256
}

// CTAs per (sequence, head).  Splitting pays while the launch is smaller than the machine: the
// per-head work is issue-bound, and the extra cost is one DSMEM exchange of the S x S partials.
// PB2_ATTN_CLUSTER=1|2|4 overrides (for tests and measurements).
static int g_attn_cluster_forced = -1;
static int attn_tc_cluster(int NB, int H, int SP, int D, bool bwd) {
    int& forced = g_attn_cluster_forced;
    if (forced < 0) {
        const char* e = getenv("PB2_ATTN_CLUSTER");
        forced = e ? atoi(e) : 0;
    }
    if (forced == 1) return 1;
    const int cands[2] = {4, 2};
    for (int c : cands) {
        if (forced && c != forced) continue;
        if (D % (16 * c) != 0 || D / c < 32) continue;
        if (attn_tc_smem(SP, D / c, c, bwd) > 96 * 1024) continue;
        if (!forced && static_cast<long long>(NB) * H * c > 2 * 148) continue;
        return c;
    }
    return 1;
}

static int attn_tc_launch(const void* kernel, size_t smem, int ctas, int cs, void** args, cudaStream_t stream) {
    if (ensure_max_smem(kernel, 227 * 1024)) return -1;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n_attr = 0;
    if (pdl_enabled()) {
        attr[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n_attr].val.programmaticStreamSerializationAllowed = 1;
        ++n_attr;
    }
    if (cs > 1) {
        attr[n_attr].id = cudaLaunchAttributeClusterDimension;
        attr[n_attr].val.clusterDim.x = cs;
        attr[n_attr].val.clusterDim.y = 1;
        attr[n_attr].val.clusterDim.z = 1;
        ++n_attr;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n_attr;
    PB2_CUDA(cudaLaunchKernelExC(&cfg, kernel, args));
    return 0;
}

int attn_tc_fwd(const void* qkv, void* out, int NB, int S, int H, int D, int causal, float drop_p,
                unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream) {
    const int SP = (S + 15) / 16 * 16;
    const int cs = attn_tc_cluster(NB, H, SP, D, false);
    AttnTc a{NB, S, H, D, H * D, SP, 1.f / sqrtf(static_cast<float>(D)), causal, drop_p, seed, rng_stream, cs, D / cs};
    const __nv_bfloat16* qp = static_cast<const __nv_bfloat16*>(qkv);
    __nv_bfloat16* op = static_cast<__nv_bfloat16*>(out);
    void* args[] = {&qp, &op, &a};
    const void* kern = cs == 4 ? reinterpret_cast<const void*>(attn_tc_fwd_kernel<4>)
                     : cs == 2 ? reinterpret_cast<const void*>(attn_tc_fwd_kernel<2>)
                               : reinterpret_cast<const void*>(attn_tc_fwd_kernel<1>);
    if (attn_tc_launch(kern, attn_tc_smem(SP, D / cs, cs, false), NB * H * cs, cs,
                       args, stream))
        return -1;
    PB2_LAUNCHED("attn_tc_fwd_kernel");
    return 0;
}

int attn_tc_bwd(const void* qkv, const void* dout, void* dqkv, int NB, int S, int H, int D, int causal, float drop_p,
                unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream) {
    const int SP = (S + 15) / 16 * 16;
    const int cs = attn_tc_cluster(NB, H, SP, D, true);
    AttnTc a{NB, S, H, D, H * D, SP, 1.f / sqrtf(static_cast<float>(D)), causal, drop_p, seed, rng_stream, cs, D / cs};
    const __nv_bfloat16* qp = static_cast<const __nv_bfloat16*>(qkv);
    const __nv_bfloat16* dp = static_cast<const __nv_bfloat16*>(dout);
    __nv_bfloat16* gp = static_cast<__nv_bfloat16*>(dqkv);
    void* args[] = {&qp, &dp, &gp, &a};
    const void* kern = cs == 4 ? reinterpret_cast<const void*>(attn_tc_bwd_kernel<4>)
                     : cs == 2 ? reinterpret_cast<const void*>(attn_tc_bwd_kernel<2>)
                               : reinterpret_cast<const void*>(attn_tc_bwd_kernel<1>);
    if (attn_tc_launch(kern, attn_tc_smem(SP, D / cs, cs, true), NB * H * cs, cs,
                       args, stream))
        return -1;
    PB2_LAUNCHED("attn_tc_bwd_kernel");
    return 0;
}

}  // namespace pb2

// Debug hook (not part of the supported ABI): force the CTAs per head (1, 2, 4; 0 = choose).
extern "C" void pb2_debug_attn_cluster(int cs) { pb2::g_attn_cluster_forced = cs; }

extern "C" int pb2_debug_attn_timestamps(unsigned long long* out16) {
    return cudaMemcpyFromSymbol(out16, pb2::g_attn_ts, sizeof(unsigned long long) * 16) == cudaSuccess ? 0 : -1;
}
