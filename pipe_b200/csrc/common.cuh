// Shared device/host helpers for the pipe_b200 sm_100a kernels.
//
// Everything here is inline PTX for Blackwell (tcgen05 / TMEM / TMA / mbarrier)
// plus the counter-based generator used by every fused-dropout epilogue.  No
// CUTLASS/CuTe dependency: descriptor encodings follow the PTX ISA bit layouts.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <atomic>

// ---------------------------------------------------------------------------
// host-side bookkeeping shared by all translation units
// ---------------------------------------------------------------------------
namespace pb2 {

extern std::atomic<long long> g_launches;        // kernels launched by this library
void set_error(const char* fmt, ...);            // remembered for pb2_last_error()
int check_cuda(cudaError_t e, const char* what); // 0 on success, else sets error and returns -1

#define PB2_CUDA(expr)                                   \
    do {                                                 \
        if (pb2::check_cuda((expr), #expr)) return -1;   \
    } while (0)

#define PB2_LAUNCHED(name)                                         \
    do {                                                           \
        pb2::g_launches.fetch_add(1, std::memory_order_relaxed);   \
        if (pb2::check_cuda(cudaGetLastError(), name)) return -1;  \
    } while (0)

// Launch with programmatic stream serialization (programmatic dependent launch): the grid may be
// scheduled while its predecessor on the stream is still running, and must execute
// griddepcontrol.wait (pdl_wait()) before it touches anything the predecessor reads or writes.
// Every kernel of this library that is launched this way waits first thing, so all it saves is the
// launch latency between dependent kernels (1.5 - 2.5 us each, ~900 times per training step).
// PB2_NO_PDL=1 in the environment falls back to plain launches.
bool pdl_enabled();
int record_and_wait(cudaStream_t recorded_on, cudaStream_t waiter);   // event record on one stream, wait on the other
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Driver entry point for tensor-map encoding, fetched through the runtime so the
// library does not link against libcuda directly.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_tiled();

// 2-D bf16 row-major [rows, cols] tensor map with a [box_rows, 64] box and 128-byte swizzle.
int make_tmap_bf16_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols);

// 2-D fp32 row-major [rows, cols] tensor map with a [box_rows, box_cols <= 32] box and 128-byte swizzle (TMA stores).
int make_tmap_f32_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                     uint32_t box_rows, uint32_t box_cols);

// Raise a kernel's dynamic shared-memory limit once per (kernel, device, host thread).
int ensure_max_smem(const void* func, int bytes);

}  // namespace pb2

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
#ifdef __CUDACC__

namespace pb2 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t"
        ".reg .b32 rx;\n\t"
        ".reg .pred px;\n\t"
        "elect.sync rx|px, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, px;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}

// ---- mbarrier ---------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// ---- TMA ----------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tile load: c0 = element index along the contiguous dimension, c1 = row.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int32_t c0,
                                            int32_t c1, uint64_t cache_hint) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
        "l"(cache_hint)
        : "memory");
}
// L2 eviction-priority policies (createpolicy encodings used by TMA cache hints).
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

// ---- programmatic dependent launch ---------------------------------------------------
// launch_dependents: the next kernel in the stream (if launched with programmatic serialization)
// may start its prologue now.  wait: block until every prerequisite grid has completed and its
// writes are visible.  Everything a kernel does before its wait must not touch data that the
// preceding kernel produces.
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- thread-block clusters / distributed shared memory ------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// All threads of all CTAs in the cluster; release/acquire so DSMEM stores before it are visible after it.
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// The two halves of the cluster barrier, for code that has independent work between them.
__device__ __forceinline__ void cluster_arrive() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Address of the same shared-memory location in CTA `rank` of this cluster.
__device__ __forceinline__ uint32_t dsmem_addr(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void dsmem_st_f32x4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void dsmem_st_f32(uint32_t addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// Remote arrival without a fence: used for "I have read what you sent" (the loads it orders have returned their values by
// program order), never to publish data.
__device__ __forceinline__ void mbar_arrive_remote_relaxed(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// 16 bytes into a peer's shared memory, completing 16 bytes of a transaction on the peer's mbarrier: data and signal travel
// together, so neither side needs a fence (a release at cluster scope costs a MEMBAR.GPU per arrival: 3.7 us per tile measured).
__device__ __forceinline__ void dsmem_st_async_f32x4(uint32_t addr, uint32_t mbar, float a, float b, float c, float d) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%2, %3, %4, %5}, [%1];"
                 ::"r"(addr), "r"(mbar), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// Execution-only cluster barrier: no release / acquire of memory, so no MEMBAR.GPU in front of the arrival (the
// release form costs one per thread).  For "everybody has reached this point" when the data that follows is published
// some other way (st.async + mbarrier transaction counts).
__device__ __forceinline__ void cluster_sync_relaxed() {
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}

// ---- tcgen05 / TMEM -----------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
                 "r"(cols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate, one CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive on an mbarrier once every MMA issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread t = lane base + t).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor for a K-major bf16 tile stored as rows of 64
// elements (128 bytes) with the 128-byte swizzle TMA applies: 8-row groups are
// 1024 bytes apart (SBO), LBO is unused for swizzled K-major layouts.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);        // start address  [0,14)
    d |= static_cast<uint64_t>(1) << 16;                           // LBO (ignored)  [16,30)
    d |= static_cast<uint64_t>(1024 >> 4) << 32;                   // SBO = 1024 B   [32,46)
    d |= static_cast<uint64_t>(1) << 46;                           // descriptor version (Blackwell)
    d |= static_cast<uint64_t>(2) << 61;                           // SWIZZLE_128B
    return d;
}

// Instruction descriptor: bf16 x bf16 -> fp32, both operands K-major, tile M x N.
__host__ __device__ __forceinline__ uint32_t umma_idesc_bf16(uint32_t m, uint32_t n) {
    uint32_t d = 0;
    d |= 1u << 4;          // D format: f32
    d |= 1u << 7;          // A format: bf16
    d |= 1u << 10;         // B format: bf16
    d |= (n >> 3) << 17;   // N / 8
    d |= (m >> 4) << 24;   // M / 16
    return d;
}

// Shared-memory descriptor of an MN-major bf16 operand stored as rows of 64 MN-elements (128 bytes, the row
// index is K) with the 128-byte swizzle: 8-row K groups are 1024 bytes apart (SBO); further 64-element MN
// atoms start `lbo_bytes` apart (LBO).  This is what a TMA box of [rows, 64 columns] of a row-major matrix
// whose *rows* are the reduction dimension gives (V for P V, dO for P^T dO, ...).
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

// bf16 x bf16 -> fp32 instruction descriptor with explicit operand majors (0 = K-major, 1 = MN-major)
__host__ __device__ __forceinline__ uint32_t umma_idesc_bf16_major(uint32_t m, uint32_t n, uint32_t a_mn, uint32_t b_mn) {
    return umma_idesc_bf16(m, n) | (a_mn << 15) | (b_mn << 16);
}

// ---- counter-based random bits -------------------------------------------------------
// Stateless generator: 128 random bits as a pure function of (seed, stream, counter), built from
// two SplitMix64 finalisers (2 x 64-bit multiply-xorshift).  Dropout needs reproducibility — the
// backward and a checkpoint recompute must regenerate the forward's mask from (seed, stream) alone —
// and decent uniformity, not cryptographic quality; this costs ~30 integer instructions per four
// elements, which matters in GEMM epilogues that sit on a kernel's critical path.
struct CounterRng {
    uint64_t key;
    __device__ __forceinline__ explicit CounterRng(uint64_t seed)
        : key(seed * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull) {}
    __device__ __forceinline__ static uint64_t mix(uint64_t z) {
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    // 4 x 32 random bits for `counter`; `stream` separates call sites.
    __device__ __forceinline__ uint4 operator()(uint64_t counter, uint64_t stream) const {
        const uint64_t x = (key ^ (stream * 0xC2B2AE3D27D4EB4Full)) + counter * 0x9E3779B97F4A7C15ull;
        const uint64_t a = mix(x), b = mix(x ^ 0xA5A5A5A5A5A5A5A5ull);
        return make_uint4(static_cast<uint32_t>(a), static_cast<uint32_t>(a >> 32), static_cast<uint32_t>(b),
                          static_cast<uint32_t>(b >> 32));
    }
};

// Dropout keep-decision for element `idx` of a tensor: one generator call serves the
// four elements idx&~3 .. idx|3.  keep iff u32 >= threshold (threshold = p * 2^32).
__device__ __forceinline__ bool dropout_keep(const CounterRng& ph, uint64_t stream, uint64_t idx, uint32_t threshold) {
    uint4 r = ph(idx >> 2, stream);
    uint32_t v = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
    return v >= threshold;
}

// 16-byte vector reduction to global memory (one L2 atomic operation for four floats).
__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace pb2

#endif  // __CUDACC__
