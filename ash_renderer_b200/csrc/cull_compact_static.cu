// cull_compact_static.cu — ashcuda_cull_compact_static: the ordered draw-command compaction of the occlusion cull
// (k_cull_compact, occlusion_cull.cu; occlusion_cull.comp:227-232 + PIN-7) for callers that cannot call cudaMalloc:
// code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation.
//
// NO COUNTERPART IN THE REFERENCE.  It changes no existing code path.
//
// The compaction needs scratch memory: the number of visible objects per CTA of 256 objects, scanned to give each CTA
// its base slot.  ashcuda_occlusion_cull keeps that scratch in the context (cudaMalloc'ed by ashcuda_create); here it
// is a file-scope static __device__ array of a fixed maximum size, so the entry point allocates NOTHING: no
// cudaMalloc / cudaFree / cudaMallocAsync, no stream or event creation, no allocating library routine.  Consequences,
// all documented in include/ashcuda.h:
//   * every GPU that loads the library holds the array (ASH_STATIC_COMPACT_MAX_OBJECTS / 256 x 4 B = 16 KB) until exit;
//   * object_count above ASH_STATIC_COMPACT_MAX_OBJECTS is rejected with ASH_ERR_INVALID and a message;
//   * every call on a device shares the one array and finds the previous call's contents in it.  Nothing has to start at
//     zero: pass 1 overwrites entries [0, ceil(object_count / 256)) before pass 2 reads exactly those, and *visible_count
//     is stored, not accumulated.  Two calls that OVERLAP on one GPU (different streams, or different host threads
//     without ordering) race on the array: serialise them.
//
// In host code the array's name is a host-side stand-in, not a device address.  The device address is obtained once
// per device, with that device current, by cudaGetSymbolAddress (return code checked) and passed to the kernels.
#include <cstdio>
#include <mutex>
#include "ash_internal.cuh"

namespace ash {

constexpr int kStaticThreads = 256;                                         // objects per CTA, as kCullThreads
constexpr uint32_t kStaticMaxCtas = ASH_STATIC_COMPACT_MAX_OBJECTS / kStaticThreads;
static_assert(ASH_STATIC_COMPACT_MAX_OBJECTS % kStaticThreads == 0, "whole CTAs");

static __device__ __align__(256) uint32_t g_static_cta_counts[kStaticMaxCtas];   // the scratch: 4096 x 4 B = 16 KB per GPU

// pass 1: visible objects per CTA (visibility[i] != 0, the shader's `visibility[objectIdx] = 1u`, :227)
__global__ void __launch_bounds__(kStaticThreads)
k_static_count(const uint32_t* __restrict__ visibility, uint32_t object_count, uint32_t* __restrict__ cta_counts) {
    __shared__ uint32_t s_warp[kStaticThreads / 32];
    const uint32_t idx = blockIdx.x * kStaticThreads + threadIdx.x;
    const bool vis = idx < object_count && __ldg(visibility + idx) != 0u;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, vis);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = __popc(ballot);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t n = 0;
#pragma unroll
        for (int w = 0; w < kStaticThreads / 32; ++w) n += s_warp[w];
        cta_counts[blockIdx.x] = n;                                          // every entry this call reads is overwritten here
    }
}

// pass 2: slot = number of visible objects with a smaller index (PIN-7: ascending object index, the serial-execution
// outcome of the shader's atomicAdd, :229); drawCommandsOut[slot] = drawCommandsIn[obj.drawIndex] (:231-232), zeros for a
// drawIndex past the template buffer (robustBufferAccess), exactly as k_cull_compact
__global__ void __launch_bounds__(kStaticThreads)
k_static_compact(const AshCullObjectData* __restrict__ objs, const AshIndirectDrawCommand* __restrict__ draws_in,
                 uint32_t n_draws, const uint32_t* __restrict__ visibility, uint32_t object_count,
                 const uint32_t* __restrict__ cta_counts, AshIndirectDrawCommand* __restrict__ draws_out,
                 uint32_t* __restrict__ visible_count) {
    __shared__ uint32_t s_red[kStaticThreads / 32];
    __shared__ uint32_t s_warp_cnt[kStaticThreads / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t part = 0;                                                       // sum of the counts of all earlier CTAs
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += kStaticThreads) part += cta_counts[b];
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, s);
    if (lane == 0) s_red[warp] = part;
    const uint32_t idx = blockIdx.x * kStaticThreads + threadIdx.x;
    const bool vis = idx < object_count && __ldg(visibility + idx) != 0u;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, vis);
    if (lane == 0) s_warp_cnt[warp] = __popc(ballot);
    __syncthreads();
    uint32_t base = 0, woff = 0;
#pragma unroll
    for (int w = 0; w < kStaticThreads / 32; ++w) {
        base += s_red[w];
        if (w < warp) woff += s_warp_cnt[w];
    }
    if (vis) {
        const uint32_t slot = base + woff + __popc(ballot & ((1u << lane) - 1u));
        const uint32_t di = objs[idx].draw_index;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(draws_in + di);
        uint32_t* dst = reinterpret_cast<uint32_t*>(draws_out + slot);
#pragma unroll
        for (int k = 0; k < 5; ++k) dst[k] = di < n_draws ? __ldg(src + k) : 0u;
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *visible_count = base + cta_counts[blockIdx.x];
}

__global__ void k_static_zero(uint32_t* __restrict__ visible_count) { *visible_count = 0u; }

// device address of the scratch array, resolved once per device (the array's name means nothing to a kernel launch)
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mutex;
static uint32_t* g_static_ptr[kStaticMaxDevices] = {};

static cudaError_t static_scratch(int device, uint32_t** out) {            // `device` is current
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_cta_counts);
        if (e != cudaSuccess) return e;
        g_static_ptr[device] = static_cast<uint32_t*>(p);
    }
    *out = g_static_ptr[device];
    return cudaSuccess;
}

}  // namespace ash

using namespace ash;

static int static_fail(char* err, size_t err_len, int code, const char* what, const char* detail = "") {
    if (err && err_len) std::snprintf(err, err_len, "ashcuda_cull_compact_static: %s%s", what, detail);
    return code;
}

extern "C" int ashcuda_cull_compact_static(int device, void* cuda_stream, const AshCullObjectData* objs,
                                           const AshIndirectDrawCommand* draws_in, uint32_t n_draws,
                                           const uint32_t* visibility, uint32_t object_count,
                                           AshIndirectDrawCommand* draws_out, uint32_t* visible_count,
                                           char* err, size_t err_len) {
    if (err && err_len) err[0] = '\0';
    if (!visible_count) return static_fail(err, err_len, ASH_ERR_INVALID, "visible_count is NULL");
    if (object_count > ASH_STATIC_COMPACT_MAX_OBJECTS) {
        char d[96];
        std::snprintf(d, sizeof d, "%u > ASH_STATIC_COMPACT_MAX_OBJECTS (%u): the static scratch array cannot hold it",
                      object_count, ASH_STATIC_COMPACT_MAX_OBJECTS);
        return static_fail(err, err_len, ASH_ERR_INVALID, "object_count ", d);
    }
    if (object_count && (!objs || !visibility || !draws_out || (n_draws && !draws_in)))
        return static_fail(err, err_len, ASH_ERR_INVALID, "NULL device pointer");
    if (device < 0 || device >= kStaticMaxDevices) return static_fail(err, err_len, ASH_ERR_DEVICE, "device index out of range");
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_fail(err, err_len, ASH_ERR_DEVICE, "no usable CUDA device: ", cudaGetErrorString(e));
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    uint32_t* scratch = nullptr;
    e = static_scratch(device, &scratch);
    if (e == cudaSuccess) {
        if (object_count == 0) {
            k_static_zero<<<1, 1, 0, s>>>(visible_count);
        } else {
            const uint32_t grid = (object_count + kStaticThreads - 1) / kStaticThreads;
            k_static_count<<<grid, kStaticThreads, 0, s>>>(visibility, object_count, scratch);
            k_static_compact<<<grid, kStaticThreads, 0, s>>>(objs, draws_in, n_draws, visibility, object_count, scratch,
                                                             draws_out, visible_count);
        }
        e = cudaGetLastError();
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) return static_fail(err, err_len, ASH_ERR_DEVICE, "CUDA error: ", cudaGetErrorString(e));
    return ASH_OK;
}
