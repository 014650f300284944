// Gradient all-reduce over NVLink 5 / NVSwitch, fused with scale (1/replicas), squared-norm
// accumulation and an optional bf16 cast-out — the communication step of the Pipe-inside-DDP
// hybrid (k pipeline replicas x n stages; the k GPUs that hold the same stage form a group).
//
// The gradients live in *symmetric memory*: every rank of the group can address every other rank's
// buffer (peer-mapped pointers) and, where the fabric supports it, a multicast alias of all of them.
// One kernel per rank, launched at the same point of each rank's stream, does the whole collective
// with plain loads and stores — no NCCL, no copy engine, no staging buffers:
//
//   two-shot P2P      rank r owns slice r.  It loads slice r from all k buffers over NVLink, adds in
//   (any topology)    fp32, scales, accumulates the squared norm, and stores the result back into
//                     slice r of all k buffers (reduce-scatter by pull, all-gather by push).  Slice r
//                     of every buffer is read and written by rank r only, so the exchange is in
//                     place and needs just two cross-GPU barriers: "inputs ready" and "outputs landed".
//   NVLS multimem     the same ownership, but `multimem.ld_reduce` lets the switch add the k copies
//   (NVSwitch)        and return one value, and `multimem.st` broadcasts the result to all k buffers:
//                     each GPU moves 1/k of the bytes the P2P variant moves.
//
// Cross-GPU barriers use flags in the symmetric buffer's signal pad: release stores to the peers'
// pads, acquire spins on the local pad, a launch epoch instead of resets.  Every CTA handshakes
// with its counterpart CTA on the peers, so there is no intra-GPU grid barrier.  A spin that
// exceeds its budget traps with the rank / phase / peer that failed to show up (the watchdog NCCL
// would otherwise provide).  The kernel is sized small (default 64 of 148 SMs) so it can share the GPU
// with the backward of the pipeline stages that are still running.
//
// Reference counterpart: none in code — the reference only remarks that DDP may wrap Pipe
// (pipe.py:290-293); the baseline is NCCL's all-reduce.  SURVEY.md K18, W11, §5.8.
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kArThreads = 512;
constexpr int kArMaxWorld = 8;
constexpr int kArMaxCtas = 128;

struct ArArgs {
    float* bufs[kArMaxWorld];           // peer-mapped data pointers (index = rank in group)
    uint32_t* pads[kArMaxWorld];        // peer-mapped signal pads: [2 phases][kArMaxCtas][kArMaxWorld] u32
    float* mc;                          // multicast alias of the data (NVLS) or null
    int world, rank;
    size_t n;                           // elements, multiple of 4 * world
    float scale;
    float* sumsq;                       // optional: += sum of squares of this rank's reduced slice
    __nv_bfloat16* cast_out;            // optional local bf16 copy of the full reduced buffer
    uint32_t epoch;
    unsigned long long spin_budget;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Barrier between CTA `blockIdx.x` of every rank.  All threads of the CTA call it.
__device__ __forceinline__ void cross_gpu_barrier(const ArArgs& a, int phase) {
    __syncthreads();
    if (threadIdx.x < a.world) {
        const int peer = threadIdx.x;
        const size_t slot = (static_cast<size_t>(phase) * kArMaxCtas + blockIdx.x) * kArMaxWorld;
        __threadfence_system();
        st_release_sys(a.pads[peer] + slot + a.rank, a.epoch);              // tell `peer` I am here
        const uint32_t* mine = a.pads[a.rank] + slot + peer;                // wait until `peer` told me
        unsigned long long spins = 0;
        while (ld_acquire_sys(mine) != a.epoch) {
            if (++spins > a.spin_budget) {
                printf("pb2 all-reduce watchdog: rank %d CTA %d phase %d never heard from peer %d (epoch %u)\n",
                       a.rank, blockIdx.x, phase, peer, a.epoch);
                __trap();
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st_f4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ void finish(const ArArgs& a, float4& v, size_t idx, float& ss) {
    v.x *= a.scale; v.y *= a.scale; v.z *= a.scale; v.w *= a.scale;
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
}

template <bool kMultimem>
__global__ void __launch_bounds__(kArThreads)
allreduce_kernel(const ArArgs a) {
    asm volatile("griddepcontrol.launch_dependents;");
    cross_gpu_barrier(a, 0);                                   // every peer's gradients are final

    const size_t slice = a.n / a.world;
    const size_t begin = slice * a.rank;
    const size_t stride = static_cast<size_t>(gridDim.x) * kArThreads * 4;
    float ss = 0.f;

    // independent float4 per thread per trip (P2P holds world x kUnroll loads in registers)
    constexpr int kUnroll = kMultimem ? 4 : 2;
    for (size_t off = (static_cast<size_t>(blockIdx.x) * kArThreads + threadIdx.x) * 4; off < slice; off += stride * kUnroll) {
        float4 v[kUnroll];
        bool has[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            has[u] = off + u * stride < slice;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (kMultimem) {
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)
                if (has[u]) v[u] = multimem_ld_reduce_f4(a.mc + begin + off + u * stride);
        } else {
            float4 t[kUnroll][kArMaxWorld];
#pragma unroll
            for (int q = 0; q < kArMaxWorld; ++q) {
                if (q < a.world) {
#pragma unroll
                    for (int u = 0; u < kUnroll; ++u)
                        if (has[u]) t[u][q] = ld_f4(a.bufs[q] + begin + off + u * stride);
                }
            }
#pragma unroll
            for (int q = 0; q < kArMaxWorld; ++q) {
                if (q < a.world) {
#pragma unroll
                    for (int u = 0; u < kUnroll; ++u)
                        if (has[u]) { v[u].x += t[u][q].x; v[u].y += t[u][q].y; v[u].z += t[u][q].z; v[u].w += t[u][q].w; }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (has[u]) finish(a, v[u], begin + off + u * stride, ss);
        if (kMultimem) {
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)
                if (has[u]) multimem_st_f4(a.mc + begin + off + u * stride, v[u]);
        } else {
#pragma unroll
            for (int q = 0; q < kArMaxWorld; ++q) {
                if (q < a.world) {
#pragma unroll
                    for (int u = 0; u < kUnroll; ++u)
                        if (has[u]) st_f4(a.bufs[q] + begin + off + u * stride, v[u]);
                }
            }
        }
    }

    if (a.sumsq) {
        __shared__ float red[kArThreads / 32];
        ss = warp_sum(ss);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        if (threadIdx.x < 32) {
            float t = threadIdx.x < kArThreads / 32 ? red[threadIdx.x] : 0.f;
            t = warp_sum(t);
            if (threadIdx.x == 0) atomicAdd(a.sumsq, t);
        }
    }

    cross_gpu_barrier(a, 1);                                   // every peer's pushes have landed here

    if (a.cast_out) {
        // local epilogue: bf16 copy of the full reduced buffer (what a bf16 consumer would read).
        // The barrier above pairs CTA b here with CTA b on every peer and nothing else, so this CTA may
        // only read what CTA b of some peer pushed: walk every peer's slice with the producers' own
        // element mapping (offset within the slice = (b * threads + t) * 4 + trip * stride).
        const float* mine = a.bufs[a.rank];
        for (int q = 0; q < a.world; ++q) {
            const size_t base = slice * q;
            for (size_t off = (static_cast<size_t>(blockIdx.x) * kArThreads + threadIdx.x) * 4; off < slice; off += stride) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(mine + base + off));
                __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(a.cast_out + base + off);
                o[0] = __floats2bfloat162_rn(v.x, v.y);
                o[1] = __floats2bfloat162_rn(v.z, v.w);
            }
        }
    }
}

// One CTA: the flag exchange of cross_gpu_barrier on its own, for protocols whose data moves in other kernels
// (the replicated wgrad stash: pushes happen in the transpose kernels, the consumer is the wgrad GEMM).
__global__ void __launch_bounds__(32)
group_barrier_kernel(const ArArgs a, const int phase) {
    cross_gpu_barrier(a, phase);
}

}  // namespace pb2

using namespace pb2;

extern "C" int pb2_allreduce_max_ctas(void) { return kArMaxCtas; }
extern "C" size_t pb2_allreduce_pad_bytes(void) {
    return static_cast<size_t>(2) * kArMaxCtas * kArMaxWorld * sizeof(uint32_t);
}

extern "C" int pb2_group_barrier(void* const* peer_pads, int world, int rank, unsigned int epoch, int phase,
                                 cudaStream_t stream) {
    if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world) {
        set_error("pb2_group_barrier: world %d / rank %d out of range (max world %d)", world, rank, kArMaxWorld);
        return -1;
    }
    if (epoch == 0 || phase < 0 || phase > 1) { set_error("pb2_group_barrier: epoch must be non-zero, phase 0 or 1"); return -1; }
    ArArgs a;
    memset(&a, 0, sizeof(a));
    for (int q = 0; q < world; ++q) {
        if (!peer_pads[q]) { set_error("pb2_group_barrier: null pad %d", q); return -1; }
        a.pads[q] = static_cast<uint32_t*>(peer_pads[q]);
    }
    a.world = world; a.rank = rank; a.epoch = epoch;
    a.spin_budget = 4000000000ull;
    // uses CTA slot 0 of the pad: give the barrier a pad of its own, not one an all-reduce is using
    group_barrier_kernel<<<1, 32, 0, stream>>>(a, phase);
    PB2_LAUNCHED("group_barrier_kernel");
    return 0;
}

extern "C" int pb2_allreduce_f32(float* const* peer_bufs, void* const* peer_pads, float* multicast, int world,
                                 int rank, size_t n, float scale, float* sumsq, void* cast_out_bf16,
                                 unsigned int epoch, int ctas, int use_multimem, cudaStream_t stream) {
    if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world) {
        set_error("pb2_allreduce_f32: world %d / rank %d out of range (max world %d)", world, rank, kArMaxWorld);
        return -1;
    }
    if (n % (static_cast<size_t>(4) * world)) {
        set_error("pb2_allreduce_f32: element count %zu must be a multiple of 4 * world (%d)", n, 4 * world);
        return -1;
    }
    if (epoch == 0) { set_error("pb2_allreduce_f32: epoch must be non-zero (pads start zeroed)"); return -1; }
    if (use_multimem && !multicast) { set_error("pb2_allreduce_f32: multimem requested without a multicast pointer"); return -1; }
    if (n == 0) return 0;
    if (ctas <= 0) ctas = 64;
    if (ctas > kArMaxCtas) ctas = kArMaxCtas;
    ArArgs a;
    memset(&a, 0, sizeof(a));
    for (int q = 0; q < world; ++q) {
        if (!peer_bufs[q] || !peer_pads[q]) { set_error("pb2_allreduce_f32: null peer pointer %d", q); return -1; }
        if (reinterpret_cast<uintptr_t>(peer_bufs[q]) & 15) { set_error("pb2_allreduce_f32: peer buffer %d not 16-byte aligned", q); return -1; }
        a.bufs[q] = peer_bufs[q];
        a.pads[q] = static_cast<uint32_t*>(peer_pads[q]);
    }
    a.mc = multicast;
    a.world = world; a.rank = rank; a.n = n; a.scale = scale;
    a.sumsq = sumsq; a.cast_out = static_cast<__nv_bfloat16*>(cast_out_bf16);
    a.epoch = epoch;
    a.spin_budget = 4000000000ull;        // ~ tens of seconds of polling before the watchdog fires
    if (use_multimem)
        allreduce_kernel<true><<<ctas, kArThreads, 0, stream>>>(a);
    else
        allreduce_kernel<false><<<ctas, kArThreads, 0, stream>>>(a);
    PB2_LAUNCHED("allreduce_kernel");
    return 0;
}
