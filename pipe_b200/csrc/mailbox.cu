// Stage-to-stage transfers of the multi-process pipeline over NVLink peer memory, without NCCL.
//
// Every stage boundary and direction has a *mailbox* in symmetric memory (both ranks of the pair map
// each other's allocation): `slots` message slots on the receiver, one flag word per (slot, CTA) next
// to them, and one acknowledgement word per (slot, CTA) on the sender.
//
//   send kernel (sender's stream, right after the kernel that produced the tensor):
//       CTA c waits until the receiver has drained this slot's previous message (ack[c] >= seq - 1),
//       stores its slice of the tensor straight into the receiver's slot over NVLink,
//       fences, and publishes flag[c] = seq with a system-scope release store.
//   recv kernel (receiver's stream, right before the kernels that consume the tensor):
//       CTA c spins on its local flag[c] (acquire) until it reads seq, copies its slice out of the
//       slot into the destination tensor, and stores ack[c] = seq into the sender's memory.
//
// So a transfer costs one small kernel per side, the payload crosses the link exactly once and lands
// while the receiver is still busy with the previous micro-batch, and no host thread, proxy or
// rendezvous is involved (the NCCL send/recv pair this replaces costs ~70 us of pipeline time per hop
// in the 8-stage configuration).  CTA c of the sender only ever talks to CTA c of the
// receiver, so there is no grid-wide barrier on either side; `seq` is the step number, so flags are
// never reset.  A spin that exceeds its budget traps with the rank's role, the slot and the word it
// was waiting for (the watchdog NCCL would otherwise provide).
//
// Reference counterpart: Copy.forward / Copy.backward (cudaMemcpyPeerAsync on copy streams,
// README.md:185-237) in the single-process reference; NCCL p2p is this repo's baseline for the
// multi-process variant.  SURVEY.md K12, W12, §5.8.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

constexpr int kMbThreads = 256;
constexpr int kMbMaxCtas = PB2_MAILBOX_MAX_CTAS;

__device__ __forceinline__ void mb_st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t mb_ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Thread 0 of the CTA spins until *word >= want (sequence numbers only grow); everybody else waits at the barrier.
__device__ __forceinline__ void mb_wait_at_least(const uint32_t* word, uint32_t want, unsigned long long budget,
                                                 const char* what, int slot) {
    if (threadIdx.x == 0) {
        unsigned long long spins = 0;
        while (static_cast<int32_t>(mb_ld_acquire_sys(word) - want) < 0) {
            if (++spins > budget) {
                printf("pb2 mailbox watchdog: %s, slot %d, CTA %d: waited for %u, word holds %u\n", what, slot,
                       blockIdx.x, want, mb_ld_acquire_sys(word));
                __trap();
            }
        }
    }
    __syncthreads();
}

// [begin, end) of this CTA's slice, in 16-byte units
__device__ __forceinline__ void mb_slice(size_t n16, size_t& begin, size_t& end) {
    const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
    begin = per * blockIdx.x;
    end = begin + per < n16 ? begin + per : n16;
    if (begin > n16) begin = n16;
}

__global__ void __launch_bounds__(kMbThreads)
mailbox_send_kernel(const uint4* __restrict__ src, size_t n16, uint4* __restrict__ remote_slot,
                    uint32_t* __restrict__ remote_flags, const uint32_t* __restrict__ local_acks, uint32_t seq,
                    int slot, unsigned long long budget) {
    pdl_launch_dependents();
    pdl_wait();                 // `src` is produced by the preceding kernel on this stream
    mb_wait_at_least(local_acks + blockIdx.x, seq - 1, budget, "sender waiting for the slot to be drained", slot);
    size_t begin, end;
    mb_slice(n16, begin, end);
    constexpr int kUnroll = 4;
    size_t i = begin + threadIdx.x;
    for (; i + (kUnroll - 1) * kMbThreads < end; i += kUnroll * kMbThreads) {
        uint4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) v[u] = src[i + u * kMbThreads];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) remote_slot[i + u * kMbThreads] = v[u];
    }
    for (; i < end; i += kMbThreads) remote_slot[i] = src[i];
    __threadfence_system();     // my stores are performed at the peer before the flag can be seen
    __syncthreads();
    if (threadIdx.x == 0) mb_st_release_sys(remote_flags + blockIdx.x, seq);
}

__global__ void __launch_bounds__(kMbThreads)
mailbox_recv_kernel(uint4* __restrict__ dst, size_t n16, const uint4* __restrict__ local_slot,
                    const uint32_t* __restrict__ local_flags, uint32_t* __restrict__ remote_acks, uint32_t seq,
                    int slot, unsigned long long budget) {
    pdl_launch_dependents();
    pdl_wait();                 // `dst` may still be in use by the preceding kernel (allocator reuse)
    mb_wait_at_least(local_flags + blockIdx.x, seq, budget, "receiver waiting for the message", slot);
    size_t begin, end;
    mb_slice(n16, begin, end);
    constexpr int kUnroll = 4;
    size_t i = begin + threadIdx.x;
    for (; i + (kUnroll - 1) * kMbThreads < end; i += kUnroll * kMbThreads) {
        uint4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) v[u] = __ldcg(local_slot + i + u * kMbThreads);   // written by the peer: bypass L1
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) dst[i + u * kMbThreads] = v[u];
    }
    for (; i < end; i += kMbThreads) dst[i] = __ldcg(local_slot + i);
    __syncthreads();            // every thread has read its part of the slot
    if (threadIdx.x == 0) mb_st_release_sys(remote_acks + blockIdx.x, seq);
}

static unsigned long long mailbox_budget() {
    // ~20 s of spinning at ~50 ns per system-scope poll; PB2_MAILBOX_TIMEOUT_S overrides
    static unsigned long long budget = 0;
    if (!budget) {
        const char* e = getenv("PB2_MAILBOX_TIMEOUT_S");
        const double s = e ? atof(e) : 20.0;
        budget = static_cast<unsigned long long>((s > 0 ? s : 20.0) * 2.0e7);
    }
    return budget;
}

static int mailbox_check(const char* who, const void* a, const void* b, size_t bytes, int ctas) {
    if (bytes == 0 || bytes % 16) {
        set_error("%s: message size must be a positive multiple of 16 bytes, got %zu", who, bytes);
        return -1;
    }
    if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15)) {
        set_error("%s: tensor and slot must be 16-byte aligned", who);
        return -1;
    }
    if (ctas < 1 || ctas > kMbMaxCtas) {
        set_error("%s: ctas must be in [1, %d], got %d", who, kMbMaxCtas, ctas);
        return -1;
    }
    return 0;
}

}  // namespace pb2

using namespace pb2;

extern "C" int pb2_mailbox_send(const void* src, size_t bytes, void* remote_slot, unsigned int* remote_flags,
                                const unsigned int* local_acks, unsigned int seq, int slot, int ctas,
                                cudaStream_t stream) {
    if (mailbox_check("pb2_mailbox_send", src, remote_slot, bytes, ctas)) return -1;
    if (seq == 0) { set_error("pb2_mailbox_send: seq starts at 1"); return -1; }
    PB2_CUDA(launch_pdl(mailbox_send_kernel, dim3(ctas), dim3(kMbThreads), 0, stream, static_cast<const uint4*>(src),
                        bytes / 16, static_cast<uint4*>(remote_slot), remote_flags, local_acks, seq, slot,
                        mailbox_budget()));
    PB2_LAUNCHED("mailbox_send_kernel");
    return 0;
}

extern "C" int pb2_mailbox_recv(void* dst, size_t bytes, const void* local_slot, const unsigned int* local_flags,
                                unsigned int* remote_acks, unsigned int seq, int slot, int ctas, cudaStream_t stream) {
    if (mailbox_check("pb2_mailbox_recv", dst, local_slot, bytes, ctas)) return -1;
    if (seq == 0) { set_error("pb2_mailbox_recv: seq starts at 1"); return -1; }
    PB2_CUDA(launch_pdl(mailbox_recv_kernel, dim3(ctas), dim3(kMbThreads), 0, stream, static_cast<uint4*>(dst),
                        bytes / 16, static_cast<const uint4*>(local_slot), local_flags, remote_acks, seq, slot,
                        mailbox_budget()));
    PB2_LAUNCHED("mailbox_recv_kernel");
    return 0;
}
