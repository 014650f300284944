// Host-side runtime shared by the kernels: error reporting, launch counter,
// TMA tensor-map encoding through the driver entry point.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace pb2 {

std::atomic<long long> g_launches{0};
static thread_local char t_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return -1;
}

EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(p);
    return fn;
}

// A tensor map is a pure function of (address, shape, stride, box): weights hit the cache every
// time, and activations mostly do too because the caching allocator hands the same blocks back.
// Direct-mapped, per host thread (no locking on the launch path).
namespace {
struct TmapKey {
    const void* ptr;
    uint64_t rows, cols, ld;
    uint32_t box_rows, box_cols;
    bool operator==(const TmapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
               box_cols == o.box_cols;
    }
};
struct TmapSlot {
    TmapKey key;
    CUtensorMap map;
    bool valid;
};
constexpr size_t kTmapSlots = 8192;
thread_local TmapSlot* t_tmap_cache = nullptr;
inline size_t tmap_hash(const TmapKey& k) {
    uint64_t h = reinterpret_cast<uint64_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows * 0xC2B2AE3D27D4EB4Full) ^ (k.cols << 17) ^ (k.ld << 29) ^ (static_cast<uint64_t>(k.box_rows) << 47);
    return static_cast<size_t>((h >> 20) % kTmapSlots);
}
}  // namespace

int make_tmap_bf16_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols) {
    if (!t_tmap_cache) t_tmap_cache = new TmapSlot[kTmapSlots]();
    const TmapKey key{ptr, rows, cols, ld_elems, box_rows, box_cols};
    TmapSlot& slot = t_tmap_cache[tmap_hash(key)];
    if (slot.valid && slot.key == key) {
        *out = slot.map;
        return 0;
    }
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return -1;
    cuuint64_t dims[2] = {cols, rows};                       // innermost first
    cuuint64_t strides[1] = {ld_elems * 2};                  // bytes, dims 1..
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] ld %llu box [%u x %u]", static_cast<int>(r),
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems, box_rows, box_cols);
        return -1;
    }
    slot.key = key;
    slot.map = *out;
    slot.valid = true;
    return 0;
}

int make_tmap_f32_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                     uint32_t box_rows, uint32_t box_cols) {
    // same direct-mapped cache as the bf16 maps; the element size is folded into the key through ld (in bytes) | 1
    if (!t_tmap_cache) t_tmap_cache = new TmapSlot[kTmapSlots]();
    const TmapKey key{ptr, rows, cols, (ld_elems * 4) | 1, box_rows, box_cols};
    TmapSlot& slot = t_tmap_cache[tmap_hash(key)];
    if (slot.valid && slot.key == key) {
        *out = slot.map;
        return 0;
    }
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return -1;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld_elems * 4};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (fp32) failed (%d) for [%llu x %llu] ld %llu box [%u x %u]", static_cast<int>(r),
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems, box_rows, box_cols);
        return -1;
    }
    slot.key = key;
    slot.map = *out;
    slot.valid = true;
    return 0;
}

// cudaFuncSetAttribute(max dynamic smem) once per (kernel, device) instead of once per launch.
// One timing-less event per device is enough for record-then-wait pairs: cudaStreamWaitEvent captures the
// work recorded by the most recent cudaEventRecord at the time of the call, later records do not disturb it.
int record_and_wait(cudaStream_t recorded_on, cudaStream_t waiter) {
    static cudaEvent_t events[64] = {nullptr};
    static std::atomic_flag busy = ATOMIC_FLAG_INIT;
    int dev = 0;
    PB2_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("record_and_wait: device ordinal %d out of range", dev); return -1; }
    while (busy.test_and_set(std::memory_order_acquire)) {}
    int rc = 0;
    if (!events[dev] && check_cuda(cudaEventCreateWithFlags(&events[dev], cudaEventDisableTiming), "cudaEventCreate")) rc = -1;
    if (!rc && check_cuda(cudaEventRecord(events[dev], recorded_on), "cudaEventRecord")) rc = -1;
    if (!rc && check_cuda(cudaStreamWaitEvent(waiter, events[dev], 0), "cudaStreamWaitEvent")) rc = -1;
    busy.clear(std::memory_order_release);
    return rc;
}

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("PB2_NO_PDL");
        on = (e && atoi(e) != 0) ? 0 : 1;
    }
    return on != 0;
}

int ensure_max_smem(const void* func, int bytes) {
    struct Seen { const void* f; int dev; };
    thread_local Seen seen[32];
    thread_local int n_seen = 0;
    int dev = 0;
    if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return -1;
    for (int i = 0; i < n_seen; ++i)
        if (seen[i].f == func && seen[i].dev == dev) return 0;
    if (check_cuda(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes),
                   "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)"))
        return -1;
    if (n_seen < 32) seen[n_seen++] = Seen{func, dev};
    return 0;
}

}  // namespace pb2

extern "C" const char* pb2_last_error(void) { return pb2::t_error; }
extern "C" long long pb2_launch_count(void) { return pb2::g_launches.load(); }
extern "C" int pb2_version(void) { return 1; }

extern "C" int pb2_stream_wait_stream(cudaStream_t waiter, cudaStream_t waited) {
    return pb2::record_and_wait(waited, waiter);
}

// Strided fill on a raw stream (rows of `width` bytes, `pitch` bytes apart): the Python layer uses it where a region has
// to be cleared on exactly the stream a kernel was launched on (a torch op under an ExternalStream wrapper of the legacy
// default stream is NOT ordered with torch's own default stream).
extern "C" int pb2_memset2d(void* ptr, size_t pitch, int value, size_t width, size_t height, cudaStream_t stream) {
    if (!ptr || width == 0 || height == 0) return 0;
    PB2_CUDA(cudaMemset2DAsync(ptr, pitch, value, width, height, stream));
    return 0;
}
