// Row-wise fused kernels of the Transformer block that are not GEMMs:
//   * residual add + LayerNorm forward / backward (post-norm: y = LN(x + h)), with the
//     dropout that produced h re-generated from the counter RNG in the backward
//   * token embedding * sqrt(E) + sinusoidal position + dropout forward, scatter-add backward
//   * softmax cross-entropy over (padded) bf16 logits, forward + dlogits in one pass
//   * bf16 transpose with fused column sums (wgrad operands + bias gradients)
//   * sum of squares (gradient-norm partials)
//
// All are HBM-bound streaming kernels: one CTA per row, 16-byte vector loads, the row
// held in registers between the statistics pass and the normalise pass (E <= 8192).
//
// Reference counterparts: ATen dropout + add + native_layer_norm (nn.TransformerEncoderLayer,
// main.py:148), nn.Embedding + PositionalEncoding (main.py:24-40,57-73), nn.CrossEntropyLoss
// (main.py:182,217).  SURVEY.md K1, K5, K9.
#include <stdlib.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kLnThreads = 256;
constexpr int kMaxVecPerThread = 4;     // 4 x 8 bf16 x 256 threads = 8192 columns

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
    if (warp == 0) {
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    return red[0];
}
__device__ __forceinline__ float block_max(float v, float* red) {
    v = warp_max(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : -INFINITY;
    if (warp == 0) {
        t = warp_max(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    return red[0];
}

struct alignas(16) Bf16x8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
    Bf16x8 raw = *reinterpret_cast<const Bf16x8*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(raw.v[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
    Bf16x8 raw;
#pragma unroll
    for (int i = 0; i < 4; ++i) raw.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    *reinterpret_cast<Bf16x8*>(p) = raw;
}

// ---------------------------------------------------------------------------------
// y = LayerNorm(x + h) * gamma + beta         (h may be null: plain LayerNorm)
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kLnThreads)
add_ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ h,
                  const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                  float* __restrict__ mean_out, float* __restrict__ rstd_out, int E, float eps) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const __nv_bfloat16* xr = x + row * E;
    const __nv_bfloat16* hr = h ? h + row * E : nullptr;
    float z[kMaxVecPerThread][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVecPerThread; ++i) {
        const int c = (i * kLnThreads + threadIdx.x) * 8;
        if (c < E) {
            load8(xr + c, z[i]);
            if (hr) {
                float t[8];
                load8(hr + c, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[i][j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) s += z[i][j];
        }
    }
    const float mean = block_sum(s, red) / E;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVecPerThread; ++i) {
        const int c = (i * kLnThreads + threadIdx.x) * 8;
        if (c < E) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = z[i][j] - mean;
                q += d * d;
            }
        }
    }
    const float rstd = rsqrtf(block_sum(q, red) / E + eps);
    if (threadIdx.x == 0) {
        mean_out[row] = mean;
        rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < kMaxVecPerThread; ++i) {
        const int c = (i * kLnThreads + threadIdx.x) * 8;
        if (c < E) {
            float o[8];
            const float4 g0 = *reinterpret_cast<const float4*>(gamma + c), g1 = *reinterpret_cast<const float4*>(gamma + c + 4);
            const float4 b0 = *reinterpret_cast<const float4*>(beta + c), b1 = *reinterpret_cast<const float4*>(beta + c + 4);
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (z[i][j] - mean) * rstd * gg[j] + bb[j];
            store8(y + row * E + c, o);
        }
    }
}

// ---------------------------------------------------------------------------------
// backward of y = LN(x + h):   dz = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma
//   dres = dz (+ extra)                 gradient for the residual branch (x)
//   dh   = dz * dropout_mask / (1 - p)  gradient for the sub-layer output *before* its dropout
//   dgamma += dy * xhat, dbeta += dy    (fp32 atomics)
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kLnThreads)
add_ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                  const __nv_bfloat16* __restrict__ h, const float* __restrict__ gamma,
                  const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                  __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dh, float* __restrict__ dgamma,
                  float* __restrict__ dbeta, int E, float drop_p, unsigned long long seed,
                  unsigned long long rng_stream) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[kMaxVecPerThread][8], g[kMaxVecPerThread][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVecPerThread; ++i) {
        const int c = (i * kLnThreads + threadIdx.x) * 8;
        if (c < E) {
            float a[8], d[8];
            load8(x + row * E + c, a);
            if (h) {
                float t[8];
                load8(h + row * E + c, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] += t[j];
            }
            load8(dy + row * E + c, d);
            const float4 g0 = *reinterpret_cast<const float4*>(gamma + c), g1 = *reinterpret_cast<const float4*>(gamma + c + 4);
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            float dgv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xh[i][j] = (a[j] - mean) * rstd;
                g[i][j] = d[j] * gg[j];
                s1 += g[i][j];
                s2 += g[i][j] * xh[i][j];
                dgv[j] = d[j] * xh[i][j];
            }
            // 16-byte vector reductions (red.global.add.v4.f32): a quarter of the atomic operations
            red_add_f32x4(dgamma + c, dgv[0], dgv[1], dgv[2], dgv[3]);
            red_add_f32x4(dgamma + c + 4, dgv[4], dgv[5], dgv[6], dgv[7]);
            red_add_f32x4(dbeta + c, d[0], d[1], d[2], d[3]);
            red_add_f32x4(dbeta + c + 4, d[4], d[5], d[6], d[7]);
        }
    }
    const float m1 = block_sum(s1, red) / E;
    const float m2 = block_sum(s2, red) / E;
    const CounterRng ph(seed);
    const uint32_t thr = static_cast<uint32_t>(fminf(drop_p, 0.999999f) * 4294967296.0f);
    const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
#pragma unroll
    for (int i = 0; i < kMaxVecPerThread; ++i) {
        const int c = (i * kLnThreads + threadIdx.x) * 8;
        if (c < E) {
            float dz[8], dd[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) dz[j] = rstd * (g[i][j] - m1 - xh[i][j] * m2);
            store8(dres + row * E + c, dz);
            if (dh) {
                if (drop_p > 0.f) {
                    // 8 consecutive elements = two generator calls
                    const uint64_t idx = row * static_cast<uint64_t>(E) + c;
                    const uint4 r0 = ph(idx >> 2, rng_stream), r1 = ph((idx >> 2) + 1, rng_stream);
                    const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) dd[j] = rr[j] >= thr ? dz[j] * inv_keep : 0.f;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) dd[j] = dz[j];
                }
                store8(dh + row * E + c, dd);
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// embedding: out[s*N + n, :] = dropout(W[tok[n, s], :] * sqrt(E) + pe[s, :])
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
embed_fwd_kernel(const long long* __restrict__ tokens, const float* __restrict__ weight,
                 const float* __restrict__ pe, __nv_bfloat16* __restrict__ out, int N, int S, int E, float scale,
                 float drop_p, unsigned long long seed, unsigned long long rng_stream) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    const int row = blockIdx.x;            // row = s * N + n
    const int s = row / N, n = row % N;
    const long long tok = tokens[static_cast<size_t>(n) * S + s];
    const float* w = weight + static_cast<size_t>(tok) * E;
    const float* p = pe + static_cast<size_t>(s) * E;
    const CounterRng ph(seed);
    const uint32_t thr = static_cast<uint32_t>(fminf(drop_p, 0.999999f) * 4294967296.0f);
    const float inv_keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    for (int c = threadIdx.x * 4; c < E; c += blockDim.x * 4) {
        const float4 a = *reinterpret_cast<const float4*>(w + c);
        const float4 b = *reinterpret_cast<const float4*>(p + c);
        float v[4] = {a.x * scale + b.x, a.y * scale + b.y, a.z * scale + b.z, a.w * scale + b.w};
        if (drop_p > 0.f) {
            const uint64_t idx = static_cast<uint64_t>(row) * E + c;
            const uint4 r = ph(idx >> 2, rng_stream);
            const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = rr[j] >= thr ? v[j] * inv_keep : 0.f;
        }
        __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(out + static_cast<size_t>(row) * E + c);
        o[0] = __floats2bfloat162_rn(v[0], v[1]);
        o[1] = __floats2bfloat162_rn(v[2], v[3]);
    }
}

__global__ void __launch_bounds__(256)
embed_bwd_kernel(const long long* __restrict__ tokens, const __nv_bfloat16* __restrict__ dout,
                 float* __restrict__ dweight, int N, int S, int E, float scale, float drop_p,
                 unsigned long long seed, unsigned long long rng_stream) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    const int row = blockIdx.x;
    const int s = row / N, n = row % N;
    const long long tok = tokens[static_cast<size_t>(n) * S + s];
    float* dw = dweight + static_cast<size_t>(tok) * E;
    const CounterRng ph(seed);
    const uint32_t thr = static_cast<uint32_t>(fminf(drop_p, 0.999999f) * 4294967296.0f);
    const float k = scale * (drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f);
    for (int c = threadIdx.x * 4; c < E; c += blockDim.x * 4) {
        const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(dout + static_cast<size_t>(row) * E + c);
        const float2 d0 = __bfloat1622float2(d[0]), d1 = __bfloat1622float2(d[1]);
        float v[4] = {d0.x, d0.y, d1.x, d1.y};
        uint32_t rr[4] = {~0u, ~0u, ~0u, ~0u};
        if (drop_p > 0.f) {
            const uint64_t idx = static_cast<uint64_t>(row) * E + c;
            const uint4 r = ph(idx >> 2, rng_stream);
            rr[0] = r.x; rr[1] = r.y; rr[2] = r.z; rr[3] = r.w;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (drop_p <= 0.f || rr[j] >= thr) atomicAdd(dw + c + j, v[j] * k);
    }
}

// ---------------------------------------------------------------------------------
// Embedding gradient of a data-parallel hybrid, exchanged as operands.  The dense gradient is V x E fp32 (472 MB
// at V = 28782, E = 4096) of which at most `tokens` rows are non-zero, so the replicas do not all-reduce it:
//   embed_pack:    row r of this micro-batch = dout[r] * scale * dropout mask, as fp32, plus its token id, stored into
//                  this replica's region of a stash in EVERY replica's memory (peer-mapped stores, like the wgrad stash)
//   embed_scatter: after the group barrier every replica holds all replicas' rows at the same slots; one CTA per slot
//                  that is the FIRST occurrence of its token adds up all slots with that token in slot order and writes
//                  alpha * sum into dweight[token].  One writer per token, fixed order: no atomics, and every replica
//                  computes bit-identical gradients.
// ---------------------------------------------------------------------------------
struct EmbedPackArgs {
    const long long* tokens;       // [N, S]
    const __nv_bfloat16* dout;     // [S * N, E]
    float* rows;                   // stash rows of this micro-batch: [S * N, E] fp32 (local address)
    long long* toks;               // stash token ids of this micro-batch (local address)
    int N, S, E;
    float scale, drop_p;
    unsigned long long seed, rng_stream;
    int n_peers;
    long long peer_delta[PB2_MAX_PEERS];
};

__global__ void __launch_bounds__(256)
embed_pack_kernel(const EmbedPackArgs a) {
    pdl_launch_dependents();
    pdl_wait();
    const int row = blockIdx.x;
    const int s = row / a.N, n = row % a.N;
    const long long tok = a.tokens[static_cast<size_t>(n) * a.S + s];
    const CounterRng ph(a.seed);
    const uint32_t thr = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
    const float k = a.scale * (a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f);
    if (threadIdx.x == 0) {
        a.toks[row] = tok;
        for (int q = 0; q < a.n_peers; ++q)
            *reinterpret_cast<long long*>(reinterpret_cast<char*>(a.toks + row) + a.peer_delta[q]) = tok;
    }
    for (int c = threadIdx.x * 4; c < a.E; c += blockDim.x * 4) {
        const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(a.dout + static_cast<size_t>(row) * a.E + c);
        const float2 d0 = __bfloat1622float2(d[0]), d1 = __bfloat1622float2(d[1]);
        float v[4] = {d0.x * k, d0.y * k, d1.x * k, d1.y * k};
        if (a.drop_p > 0.f) {
            const uint64_t idx = static_cast<uint64_t>(row) * a.E + c;
            const uint4 r = ph(idx >> 2, a.rng_stream);
            if (r.x < thr) v[0] = 0.f;
            if (r.y < thr) v[1] = 0.f;
            if (r.z < thr) v[2] = 0.f;
            if (r.w < thr) v[3] = 0.f;
        }
        float* dst = a.rows + static_cast<size_t>(row) * a.E + c;
        const float4 out = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(dst) = out;
        for (int q = 0; q < a.n_peers; ++q)
            *reinterpret_cast<float4*>(reinterpret_cast<char*>(dst) + a.peer_delta[q]) = out;
    }
    // see transpose_multi_kernel: the peer stores must have been performed when this grid retires
    if (a.n_peers > 0) {
        __syncthreads();
        if (threadIdx.x == 0) __threadfence_system();
    }
}

constexpr int kScatterMaxSlots = 4096;           // token ids of the whole stash are staged in shared memory (32 KB)

__global__ void __launch_bounds__(256)
embed_scatter_kernel(const long long* __restrict__ toks, const float* __restrict__ rows, float* __restrict__ dweight,
                     int R, int E, float alpha) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ long long s_tok[kScatterMaxSlots];
    __shared__ int earlier;
    const int r = blockIdx.x;
    const long long tok = toks[r];
    if (tok < 0) return;                                   // unused slot (uniform for the CTA)
    if (threadIdx.x == 0) earlier = 0;
    for (int i = threadIdx.x; i < R; i += blockDim.x) s_tok[i] = toks[i];
    __syncthreads();
    // not the first slot with this token: an earlier CTA owns the sum
    for (int i = threadIdx.x; i < r; i += blockDim.x)
        if (s_tok[i] == tok) earlier = 1;
    __syncthreads();
    if (earlier) return;
    for (int c = threadIdx.x * 4; c < E; c += blockDim.x * 4) {
        float4 acc = *reinterpret_cast<const float4*>(rows + static_cast<size_t>(r) * E + c);
        for (int i = r + 1; i < R; ++i) {
            if (s_tok[i] != tok) continue;                 // uniform across the CTA, broadcast read
            const float4 v = *reinterpret_cast<const float4*>(rows + static_cast<size_t>(i) * E + c);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        float4* dw = reinterpret_cast<float4*>(dweight + static_cast<size_t>(tok) * E + c);
        float4 old = *dw;
        old.x += acc.x * alpha; old.y += acc.y * alpha; old.z += acc.z * alpha; old.w += acc.w * alpha;
        *dw = old;
    }
}

// ---------------------------------------------------------------------------------
// softmax cross-entropy over bf16 logits [T, ld] (V valid columns).  One CTA per row:
//   loss[row] = logsumexp(l) - l[target];   logits <- softmax(l) - onehot(target)   (in place)
// ---------------------------------------------------------------------------------
constexpr int kCeThreads = 512;
constexpr int kCeVecs = 8;           // 16-byte vectors per thread held in registers: rows up to 512 * 8 * 8 = 32768 logits

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                  float* __restrict__ loss_rows, const float* __restrict__ row_scale, int V, long long ld,
                  int write_grad) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    __nv_bfloat16* l = logits + row * ld;
    const long long tgt = targets[row];

    if ((ld & 7) == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0 && V <= kCeThreads * kCeVecs * 8) {
        // One trip through memory: the row (57 KB for the tutorial vocabulary) is read once with 16-byte
        // loads, lives in registers for the max / sum / gradient passes, and is written back once.
        uint4 raw[kCeVecs];
        float mx = -INFINITY;
#pragma unroll
        for (int u = 0; u < kCeVecs; ++u) {
            const int c = (u * kCeThreads + threadIdx.x) * 8;
            raw[u] = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);      // bf16 -inf
            if (c < V) raw[u] = *reinterpret_cast<const uint4*>(l + c);
        }
#pragma unroll
        for (int u = 0; u < kCeVecs; ++u) {
            const int c = (u * kCeThreads + threadIdx.x) * 8;
            const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&raw[u]);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (c + j < V) mx = fmaxf(mx, __bfloat162float(e[j]));
        }
        mx = block_max(mx, red);
        float se = 0.f;
#pragma unroll
        for (int u = 0; u < kCeVecs; ++u) {
            const int c = (u * kCeThreads + threadIdx.x) * 8;
            const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&raw[u]);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (c + j < V) se += __expf(__bfloat162float(e[j]) - mx);
        }
        se = block_sum(se, red);
        const float lse = mx + __logf(se);
        if (threadIdx.x == 0 && loss_rows) loss_rows[row] = lse - __bfloat162float(l[tgt]);
        if (write_grad) {
            const float rs = row_scale ? row_scale[row] : 1.f;
            __syncthreads();           // thread 0 has read l[tgt]
#pragma unroll
            for (int u = 0; u < kCeVecs; ++u) {
                const int c = (u * kCeThreads + threadIdx.x) * 8;
                if (c >= ld) continue;
                __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(&raw[u]);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float p = 0.f;               // columns [V, ld) are padding: zero gradient
                    if (c + j < V) {
                        p = __expf(__bfloat162float(e[j]) - lse);
                        if (c + j == tgt) p -= 1.f;
                        p *= rs;
                    }
                    e[j] = __float2bfloat16(p);
                }
                *reinterpret_cast<uint4*>(l + c) = raw[u];
            }
            // padding beyond what the register tile covers (ld much larger than V): never the case for
            // the LM head, handled for completeness
            for (long long c = static_cast<long long>(kCeThreads) * kCeVecs * 8 + threadIdx.x; c < ld; c += blockDim.x)
                l[c] = __float2bfloat16(0.f);
        }
        return;
    }

    float mx = -INFINITY;
    for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, __bfloat162float(l[c]));
    mx = block_max(mx, red);
    float se = 0.f;
    for (int c = threadIdx.x; c < V; c += blockDim.x) se += __expf(__bfloat162float(l[c]) - mx);
    se = block_sum(se, red);
    const float lse = mx + __logf(se);
    if (threadIdx.x == 0 && loss_rows) loss_rows[row] = lse - __bfloat162float(l[tgt]);
    if (write_grad) {
        const float rs = row_scale ? row_scale[row] : 1.f;
        __syncthreads();
        for (int c = threadIdx.x; c < V; c += blockDim.x) {
            float p = __expf(__bfloat162float(l[c]) - lse);
            if (c == tgt) p -= 1.f;
            l[c] = __float2bfloat16(p * rs);
        }
        for (long long c = V + threadIdx.x; c < ld; c += blockDim.x) l[c] = __float2bfloat16(0.f);
    }
}

// ---------------------------------------------------------------------------------
// dst[c, r] = src[r, c] (bf16), dst row stride ld_dst >= R; zero-fills dst columns [R, ld_dst);
// optional colsum[c] += scale * sum_r src[r, c]   (bias gradient)
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
transpose_colsum_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src, __nv_bfloat16* __restrict__ dst,
                        long long ld_dst, int R, int C, float* __restrict__ colsum, float scale) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    __shared__ __nv_bfloat16 tile[32][33];
    const int c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
    float acc = 0.f;
    for (int r0 = 0; r0 < static_cast<int>(ld_dst); r0 += 32) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = r0 + ty + k * 8, c = c0 + tx;
            __nv_bfloat16 v = __float2bfloat16(0.f);
            if (r < R && c < C) v = src[static_cast<size_t>(r) * ld_src + c];
            tile[ty + k * 8][tx] = v;
            acc += __bfloat162float(v);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int c = c0 + ty + k * 8, r = r0 + tx;
            if (c < C && r < ld_dst) dst[static_cast<size_t>(c) * ld_dst + r] = tile[tx][ty + k * 8];
        }
        __syncthreads();
    }
    if (colsum) {
        // threads with equal tx hold partial sums of column c0 + tx
        __shared__ float part[8][32];
        part[ty][tx] = acc;
        __syncthreads();
        if (ty == 0 && c0 + tx < C) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) t += part[k][tx];
            atomicAdd(colsum + c0 + tx, t * scale);
        }
    }
}

// Up to PB2_MAX_TRANSPOSES independent transposes in one launch (blockIdx.y selects the job):
// a layer's backward stacks all its wgrad operands with a single kernel.
__global__ void __launch_bounds__(256)
transpose_multi_kernel(const pb2_transpose_jobs jobs) {
    pdl_launch_dependents();
    pdl_wait();          // launched programmatically: nothing below may run before the predecessor is complete
    const pb2_transpose_job j = jobs.job[blockIdx.y];
    if (blockIdx.x * 32 >= j.C) return;
    __shared__ __nv_bfloat16 tile[32][33];
    __shared__ float part[8][32];
    const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(j.src);
    __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(j.dst);
    const int c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    float acc = 0.f;
    for (int r0 = 0; r0 < j.R; r0 += 32) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = r0 + ty + k * 8, c = c0 + tx;
            __nv_bfloat16 v = __float2bfloat16(0.f);
            if (r < j.R && c < j.C) v = src[static_cast<size_t>(r) * j.ld_src + c];
            tile[ty + k * 8][tx] = v;
            acc += __bfloat162float(v);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int c = c0 + ty + k * 8, r = r0 + tx;
            if (c < j.C && r < j.R) {
                __nv_bfloat16* at = dst + static_cast<size_t>(c) * j.ld_dst + r;
                const __nv_bfloat16 v = tile[tx][ty + k * 8];
                *at = v;
                // the same slot of every data-parallel replica's stash (peer-mapped memory over NVLink)
                for (int q = 0; q < jobs.n_peers; ++q)
                    *reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(at) + jobs.peer_delta[q]) = v;
            }
        }
        __syncthreads();
    }
    if (j.colsum) {
        part[ty][tx] = acc;
        __syncthreads();
        if (ty == 0 && c0 + tx < j.C) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) t += part[k][tx];
            atomicAdd(j.colsum + c0 + tx, t * j.scale);
        }
    }
    // Stores into a peer's memory are posted: the grid can retire while they are still on their way, and a flag raised
    // by a LATER kernel (pb2_group_barrier) may then overtake them on another NVLink lane.  A system-scope fence by the
    // storing thread itself makes "this kernel has completed" imply "its peer stores have been performed".
    // One fence per CTA is enough: the barrier orders the other threads' stores before thread 0's fence, which is cumulative.
    if (jobs.n_peers > 0) {
        __syncthreads();
        if (threadIdx.x == 0) __threadfence_system();
    }
}

__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ x, size_t n, float* __restrict__ out) {
    pdl_wait();          // launched programmatically; it does not release its own dependents early (the optimizer follows)
    __shared__ float red[32];
    float s = 0.f;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 4;
    for (size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            const float4 v = *reinterpret_cast<const float4*>(x + i);
            s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        } else {
            for (size_t j = i; j < n; ++j) s += x[j] * x[j];
        }
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(out, s);
}

}  // namespace pb2

using namespace pb2;

static int check_e(const char* who, int E) {
    if (E % 8 || E > kLnThreads * 8 * kMaxVecPerThread || E <= 0) {
        set_error("%s: row width must be a multiple of 8 and <= %d, got %d", who, kLnThreads * 8 * kMaxVecPerThread, E);
        return -1;
    }
    return 0;
}

extern "C" int pb2_add_layernorm_fwd(const void* x, const void* h, const float* gamma, const float* beta, void* y,
                                     float* mean, float* rstd, int rows, int E, float eps, cudaStream_t stream) {
    if (check_e("pb2_add_layernorm_fwd", E)) return -1;
    if (rows <= 0) return 0;
    PB2_CUDA(launch_pdl(add_ln_fwd_kernel, dim3(rows), dim3(kLnThreads), 0, stream, static_cast<const __nv_bfloat16*>(x),
                                                       static_cast<const __nv_bfloat16*>(h), gamma, beta,
                                                       static_cast<__nv_bfloat16*>(y), mean, rstd, E, eps));
    PB2_LAUNCHED("add_ln_fwd_kernel");
    return 0;
}

extern "C" int pb2_add_layernorm_bwd(const void* dy, const void* x, const void* h, const float* gamma,
                                     const float* mean, const float* rstd, void* dres, void* dh, float* dgamma,
                                     float* dbeta, int rows, int E, float drop_p, unsigned long long seed,
                                     unsigned long long rng_stream, cudaStream_t stream) {
    if (check_e("pb2_add_layernorm_bwd", E)) return -1;
    if (rows <= 0) return 0;
    PB2_CUDA(launch_pdl(add_ln_bwd_kernel, dim3(rows), dim3(kLnThreads), 0, stream, 
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const __nv_bfloat16*>(h), gamma, mean, rstd, static_cast<__nv_bfloat16*>(dres),
        static_cast<__nv_bfloat16*>(dh), dgamma, dbeta, E, drop_p, seed, rng_stream));
    PB2_LAUNCHED("add_ln_bwd_kernel");
    return 0;
}

extern "C" int pb2_embed_fwd(const long long* tokens, const float* weight, const float* pe, void* out, int N, int S,
                             int E, float scale, float drop_p, unsigned long long seed,
                             unsigned long long rng_stream, cudaStream_t stream) {
    if (E % 4) { set_error("pb2_embed_fwd: E must be a multiple of 4"); return -1; }
    if (N * S <= 0) return 0;
    PB2_CUDA(launch_pdl(embed_fwd_kernel, dim3(N * S), dim3(256), 0, stream, tokens, weight, pe, static_cast<__nv_bfloat16*>(out), N, S, E, scale,
                                                drop_p, seed, rng_stream));
    PB2_LAUNCHED("embed_fwd_kernel");
    return 0;
}

extern "C" int pb2_embed_bwd(const long long* tokens, const void* dout, float* dweight, int N, int S, int E,
                             float scale, float drop_p, unsigned long long seed, unsigned long long rng_stream,
                             cudaStream_t stream) {
    if (E % 4) { set_error("pb2_embed_bwd: E must be a multiple of 4"); return -1; }
    if (N * S <= 0) return 0;
    PB2_CUDA(launch_pdl(embed_bwd_kernel, dim3(N * S), dim3(256), 0, stream, tokens, static_cast<const __nv_bfloat16*>(dout), dweight, N, S, E,
                                                scale, drop_p, seed, rng_stream));
    PB2_LAUNCHED("embed_bwd_kernel");
    return 0;
}

extern "C" int pb2_embed_pack(const long long* tokens, const void* dout, float* rows, long long* toks, int N, int S, int E,
                              float scale, float drop_p, unsigned long long seed, unsigned long long rng_stream, int n_peers,
                              const long long* peer_delta, cudaStream_t stream) {
    if (E % 4) { set_error("pb2_embed_pack: E must be a multiple of 4"); return -1; }
    if (n_peers < 0 || n_peers > PB2_MAX_PEERS) { set_error("pb2_embed_pack: n_peers %d out of range", n_peers); return -1; }
    if (N * S <= 0) return 0;
    EmbedPackArgs a;
    memset(&a, 0, sizeof(a));
    a.tokens = tokens; a.dout = static_cast<const __nv_bfloat16*>(dout); a.rows = rows; a.toks = toks;
    a.N = N; a.S = S; a.E = E; a.scale = scale; a.drop_p = drop_p; a.seed = seed; a.rng_stream = rng_stream;
    a.n_peers = n_peers;
    for (int q = 0; q < n_peers; ++q) a.peer_delta[q] = peer_delta[q];
    PB2_CUDA(launch_pdl(embed_pack_kernel, dim3(N * S), dim3(256), 0, stream, a));
    PB2_LAUNCHED("embed_pack_kernel");
    return 0;
}

extern "C" int pb2_embed_scatter(long long* toks, const float* rows, float* dweight, int R, int E, float alpha,
                                 int reset_slots, cudaStream_t stream) {
    if (E % 4) { set_error("pb2_embed_scatter: E must be a multiple of 4"); return -1; }
    if (R > kScatterMaxSlots) { set_error("pb2_embed_scatter: at most %d stash slots (replicas x tokens per step), got %d", kScatterMaxSlots, R); return -1; }
    if (R <= 0) return 0;
    PB2_CUDA(launch_pdl(embed_scatter_kernel, dim3(R), dim3(256), 0, stream, toks, rows, dweight, R, E, alpha));
    PB2_LAUNCHED("embed_scatter_kernel");
    if (reset_slots) PB2_CUDA(cudaMemsetAsync(toks, 0xFF, static_cast<size_t>(R) * sizeof(long long), stream));   // all slots "empty" (-1)
    return 0;
}

extern "C" int pb2_cross_entropy(void* logits, const long long* targets, float* loss_rows, const float* row_scale,
                                 int rows, int V, long long ld, int write_grad, cudaStream_t stream) {
    if (rows <= 0) return 0;
    if (V <= 0 || ld < V) { set_error("pb2_cross_entropy: bad V/ld (%d / %lld)", V, ld); return -1; }
    PB2_CUDA(launch_pdl(ce_fwd_bwd_kernel, dim3(rows), dim3(kCeThreads), 0, stream, static_cast<__nv_bfloat16*>(logits), targets, loss_rows, row_scale, V,
                                                ld, write_grad));
    PB2_LAUNCHED("ce_fwd_bwd_kernel");
    return 0;
}

extern "C" int pb2_transpose_colsum(const void* src, long long ld_src, void* dst, long long ld_dst, int R, int C,
                                    float* colsum, float scale, cudaStream_t stream) {
    if (R <= 0 || C <= 0) return 0;
    if (ld_dst < R) { set_error("pb2_transpose_colsum: ld_dst (%lld) < rows (%d)", ld_dst, R); return -1; }
    PB2_CUDA(launch_pdl(transpose_colsum_kernel, dim3((C + 31) / 32), dim3(256), 0, stream, static_cast<const __nv_bfloat16*>(src), ld_src,
                                                               static_cast<__nv_bfloat16*>(dst), ld_dst, R, C, colsum,
                                                               scale));
    PB2_LAUNCHED("transpose_colsum_kernel");
    return 0;
}

extern "C" int pb2_transpose_multi(const pb2_transpose_jobs* jobs, cudaStream_t stream) {
    if (!jobs || jobs->count <= 0) return 0;
    if (jobs->count > PB2_MAX_TRANSPOSES) { set_error("pb2_transpose_multi: too many jobs (%d)", jobs->count); return -1; }
    if (jobs->n_peers < 0 || jobs->n_peers > PB2_MAX_PEERS) { set_error("pb2_transpose_multi: n_peers %d out of range", jobs->n_peers); return -1; }
    int max_c = 0;
    for (int i = 0; i < jobs->count; ++i) {
        if (jobs->job[i].C > max_c) max_c = jobs->job[i].C;
        if (jobs->job[i].R <= 0 || jobs->job[i].C <= 0) { set_error("pb2_transpose_multi: empty job %d", i); return -1; }
    }
    dim3 grid((max_c + 31) / 32, jobs->count);
    PB2_CUDA(launch_pdl(transpose_multi_kernel, grid, dim3(256), 0, stream, *jobs));
    PB2_LAUNCHED("transpose_multi_kernel");
    return 0;
}

extern "C" int pb2_sumsq(const float* x, size_t n, float* out, cudaStream_t stream) {
    if (n == 0) return 0;
    if (reinterpret_cast<uintptr_t>(x) & 15) { set_error("pb2_sumsq: input must be 16-byte aligned"); return -1; }
    size_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    PB2_CUDA(launch_pdl(sumsq_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, stream, x, n, out));
    PB2_LAUNCHED("sumsq_kernel");
    return 0;
}
