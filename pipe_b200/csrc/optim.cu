// Multi-tensor fused clip + Adam for a pipeline stage, one launch for every parameter.
//
// Per element:  g *= clip_coef * grad_scale;  Adam moment update;  fp32 master weight update;
// bf16 shadow weight (what the GEMMs read) and, for matrices, the *transposed* bf16 shadow
// (what the dgrad GEMMs read) written in the same pass through a shared-memory tile — so the
// step touches each parameter's state exactly once (4 fp32 reads, 3 fp32 + 2 bf16 writes).
// The clip coefficient is computed on the device from the squared gradient norm that the
// wgrad epilogues / sumsq kernel accumulated, so there is no host round trip between backward
// and the optimizer.
//
// Reference counterpart: torch.nn.utils.clip_grad_norm_(0.5) + torch.optim.Adam (main.py:184,
// 219-220).  SURVEY.md K10, K11.
#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kTileR = 32, kTileC = 64;

struct AdamHyper {
    float lr, beta1, beta2, eps, weight_decay;
    float bc1, bc2;          // 1 - beta^t
    float max_norm;          // <= 0: no clipping
    float grad_scale;        // e.g. 1 / replicas
};

__device__ __forceinline__ float adam_elem(float& w, float g, float& m, float& v, const AdamHyper& h, float coef) {
    g *= coef;
    if (h.weight_decay != 0.f) g = fmaf(h.weight_decay, w, g);
    m = fmaf(h.beta1, m, (1.f - h.beta1) * g);
    v = fmaf(h.beta2, v, (1.f - h.beta2) * g * g);
    const float denom = sqrtf(v) * rsqrtf(h.bc2) + h.eps;
    w -= (h.lr / h.bc1) * (m / denom);
    return w;
}

__global__ void __launch_bounds__(256)
adam_tiles_kernel(const pb2_adam_tensor* __restrict__ table, int n_tensors, const float* __restrict__ normsq,
                  const AdamHyper h) {
    __shared__ __nv_bfloat16 tile[kTileR][kTileC + 2];
    __shared__ int s_tensor;
    // locate the tensor that owns this block's tile
    if (threadIdx.x == 0) {
        int lo = 0, hi = n_tensors - 1;
        const int b = blockIdx.x;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (table[mid].tile_start <= b) lo = mid; else hi = mid - 1;
        }
        s_tensor = lo;
    }
    __syncthreads();
    const pb2_adam_tensor t = table[s_tensor];
    const int tile_id = blockIdx.x - t.tile_start;
    const int tiles_c = (t.cols + kTileC - 1) / kTileC;
    const int r0 = (tile_id / tiles_c) * kTileR, c0 = (tile_id % tiles_c) * kTileC;

    float coef = h.grad_scale;
    if (h.max_norm > 0.f) {
        const float total = sqrtf(*normsq) * h.grad_scale;
        coef *= fminf(1.f, h.max_norm / (total + 1e-6f));
    }

    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;      // 16 x 16, 4 columns each, rows ty and ty+16
    const bool vec = (t.cols & 3) == 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int r = r0 + ty + half * 16, c = c0 + tx * 4;
        float w4[4] = {0.f, 0.f, 0.f, 0.f};
        if (r < t.rows && c < t.cols) {
            const size_t off = static_cast<size_t>(r) * t.cols + c;
            if (vec) {
                float4 w = *reinterpret_cast<float4*>(t.w + off);
                const float4 g = *reinterpret_cast<const float4*>(t.g + off);
                float4 m = *reinterpret_cast<float4*>(t.m + off);
                float4 v = *reinterpret_cast<float4*>(t.v + off);
                adam_elem(w.x, g.x, m.x, v.x, h, coef);
                adam_elem(w.y, g.y, m.y, v.y, h, coef);
                adam_elem(w.z, g.z, m.z, v.z, h, coef);
                adam_elem(w.w, g.w, m.w, v.w, h, coef);
                *reinterpret_cast<float4*>(t.w + off) = w;
                *reinterpret_cast<float4*>(t.m + off) = m;
                *reinterpret_cast<float4*>(t.v + off) = v;
                w4[0] = w.x; w4[1] = w.y; w4[2] = w.z; w4[3] = w.w;
                if (t.w16) {
                    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(static_cast<__nv_bfloat16*>(t.w16) + off);
                    o[0] = __floats2bfloat162_rn(w.x, w.y);
                    o[1] = __floats2bfloat162_rn(w.z, w.w);
                }
            } else {
                for (int j = 0; j < 4 && c + j < t.cols; ++j) {
                    float w = t.w[off + j], m = t.m[off + j], v = t.v[off + j];
                    adam_elem(w, t.g[off + j], m, v, h, coef);
                    t.w[off + j] = w; t.m[off + j] = m; t.v[off + j] = v;
                    w4[j] = w;
                    if (t.w16) static_cast<__nv_bfloat16*>(t.w16)[off + j] = __float2bfloat16(w);
                }
            }
        }
        if (t.w16t) {
#pragma unroll
            for (int j = 0; j < 4; ++j) tile[ty + half * 16][tx * 4 + j] = __float2bfloat16(w4[j]);
        }
    }
    if (t.w16t) {
        __syncthreads();
        // transposed store: thread -> column c (0..63), 8 consecutive rows
        const int c = threadIdx.x >> 2, rg = (threadIdx.x & 3) * 8;
        if (c0 + c < t.cols) {
            __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(t.w16t) + static_cast<size_t>(c0 + c) * t.ld_t + r0 + rg;
            if (r0 + rg + 8 <= t.rows && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                alignas(16) __nv_bfloat16 tmp[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) tmp[j] = tile[rg + j][c];
                *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(tmp);
            } else {
                for (int j = 0; j < 8 && r0 + rg + j < t.rows; ++j) dst[j] = tile[rg + j][c];
            }
        }
    }
}

// fp32 -> bf16 (+ transposed bf16) shadow refresh without an optimizer step (after init / load_state_dict)
__global__ void __launch_bounds__(256)
shadow_tiles_kernel(const pb2_adam_tensor* __restrict__ table, int n_tensors) {
    __shared__ __nv_bfloat16 tile[kTileR][kTileC + 2];
    __shared__ int s_tensor;
    if (threadIdx.x == 0) {
        int lo = 0, hi = n_tensors - 1;
        const int b = blockIdx.x;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (table[mid].tile_start <= b) lo = mid; else hi = mid - 1;
        }
        s_tensor = lo;
    }
    __syncthreads();
    const pb2_adam_tensor t = table[s_tensor];
    if (!t.w16 && !t.w16t) return;
    const int tile_id = blockIdx.x - t.tile_start;
    const int tiles_c = (t.cols + kTileC - 1) / kTileC;
    const int r0 = (tile_id / tiles_c) * kTileR, c0 = (tile_id % tiles_c) * kTileC;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int r = r0 + ty + half * 16;
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx * 4 + j;
            float w = 0.f;
            if (r < t.rows && c < t.cols) {
                w = t.w[static_cast<size_t>(r) * t.cols + c];
                if (t.w16) static_cast<__nv_bfloat16*>(t.w16)[static_cast<size_t>(r) * t.cols + c] = __float2bfloat16(w);
            }
            tile[ty + half * 16][tx * 4 + j] = __float2bfloat16(w);
        }
    }
    if (t.w16t) {
        __syncthreads();
        const int c = threadIdx.x >> 2, rg = (threadIdx.x & 3) * 8;
        if (c0 + c < t.cols) {
            __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(t.w16t) + static_cast<size_t>(c0 + c) * t.ld_t + r0 + rg;
            for (int j = 0; j < 8 && r0 + rg + j < t.rows; ++j) dst[j] = tile[rg + j][c];
        }
    }
}

}  // namespace pb2

using namespace pb2;

extern "C" int pb2_adam_step(const pb2_adam_tensor* table_dev, int n_tensors, int total_tiles, const float* normsq_dev,
                             float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                             float max_norm, float grad_scale, cudaStream_t stream) {
    if (n_tensors <= 0 || total_tiles <= 0) return 0;
    if (max_norm > 0.f && !normsq_dev) { set_error("pb2_adam_step: clipping needs the squared-norm scalar"); return -1; }
    AdamHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.weight_decay = weight_decay;
    h.bc1 = 1.f - powf(beta1, static_cast<float>(step));
    h.bc2 = 1.f - powf(beta2, static_cast<float>(step));
    h.max_norm = max_norm;
    h.grad_scale = grad_scale == 0.f ? 1.f : grad_scale;
    adam_tiles_kernel<<<total_tiles, 256, 0, stream>>>(table_dev, n_tensors, normsq_dev, h);
    PB2_LAUNCHED("adam_tiles_kernel");
    return 0;
}

extern "C" int pb2_refresh_shadows(const pb2_adam_tensor* table_dev, int n_tensors, int total_tiles,
                                   cudaStream_t stream) {
    if (n_tensors <= 0 || total_tiles <= 0) return 0;
    shadow_tiles_kernel<<<total_tiles, 256, 0, stream>>>(table_dev, n_tensors);
    PB2_LAUNCHED("shadow_tiles_kernel");
    return 0;
}
