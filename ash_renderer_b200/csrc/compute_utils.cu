// compute_utils.cu — test harness for ash_compute_utils.cuh (shaders/compute_utils.glsl as a CUDA header).
// One kernel that calls every function of the header on caller data, behind a context-free C entry point
// (ashcuda_debug_compute_utils) so that tests/ can compare it with the CPU restatement bit for bit.
#include <cstdint>
#include "../../include/ashcuda.h"
#include "ash_compute_utils.cuh"

namespace ash {

struct CuParams { float f[18]; };       // op 4/5/7: view_proj columns; op 6: inv_proj columns, near, far

__global__ void __launch_bounds__(256)
k_compute_utils(int op, const float* __restrict__ in, uint32_t n, CuParams p, float* __restrict__ out) {
    __shared__ cu::GroupScratch scratch;
    const uint32_t lid = threadIdx.x, gid = blockIdx.x * cu::kGroup + lid;
    M4 m;
#pragma unroll
    for (int c = 0; c < 4; ++c) m.c[c] = {p.f[c * 4 + 0], p.f[c * 4 + 1], p.f[c * 4 + 2], p.f[c * 4 + 3]};
    switch (op) {
        case ASH_CU_MIN: case ASH_CU_MAX: case ASH_CU_SUM: {        // n is a multiple of 256 (the GLSL has no tail)
            const float v = in[gid];
            const float r = op == ASH_CU_MIN ? cu::parallel_min(scratch, lid, v)
                          : op == ASH_CU_MAX ? cu::parallel_max(scratch, lid, v) : cu::parallel_sum(scratch, lid, v);
            if (lid == 0) out[blockIdx.x] = r;
            break;
        }
        case ASH_CU_SCAN: {
            const uint32_t v = reinterpret_cast<const uint32_t*>(in)[gid];
            reinterpret_cast<uint32_t*>(out)[gid] = cu::exclusive_prefix_sum(scratch, lid, v);
            break;
        }
        case ASH_CU_SPHERE_IN_FRUSTUM: case ASH_CU_AABB_IN_FRUSTUM: {
            cu::Plane planes[6];
#pragma unroll
            for (int f = 0; f < 6; ++f) planes[f] = cu::extract_plane(m, f);
            if (gid >= n) break;
            bool vis;
            if (op == ASH_CU_SPHERE_IN_FRUSTUM) vis = cu::sphere_in_frustum(v3(in[gid * 4], in[gid * 4 + 1], in[gid * 4 + 2]), in[gid * 4 + 3], planes);
            else vis = cu::aabb_in_frustum(v3(in[gid * 6], in[gid * 6 + 1], in[gid * 6 + 2]), v3(in[gid * 6 + 3], in[gid * 6 + 4], in[gid * 6 + 5]), planes);
            reinterpret_cast<uint32_t*>(out)[gid] = vis ? 1u : 0u;
            break;
        }
        case ASH_CU_DEPTH: {
            if (gid >= n) break;
            const float u = in[gid * 3], v = in[gid * 3 + 1], d = in[gid * 3 + 2];
            const V3 q = cu::uv_depth_to_view(u, v, d, m);
            out[gid * 4 + 0] = cu::linearize_depth(d, p.f[16], p.f[17]);
            out[gid * 4 + 1] = q.x; out[gid * 4 + 2] = q.y; out[gid * 4 + 3] = q.z;
            break;
        }
        case ASH_CU_PLANES: {
            if (gid < 6) {
                const cu::Plane pl = cu::extract_plane(m, (int)gid);
                out[gid * 4 + 0] = pl.normal.x; out[gid * 4 + 1] = pl.normal.y; out[gid * 4 + 2] = pl.normal.z; out[gid * 4 + 3] = pl.distance;
            }
            break;
        }
    }
}

}  // namespace ash

extern "C" int ashcuda_debug_compute_utils(int device, int op, const void* in, uint32_t n, const float* params18, void* out) {
    using namespace ash;
    if (op < ASH_CU_MIN || op > ASH_CU_PLANES || !out || (op != ASH_CU_PLANES && (!in || !n))) return ASH_ERR_INVALID;
    if (op <= ASH_CU_SCAN && n % cu::kGroup) return ASH_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return ASH_ERR_DEVICE;
    const size_t in_elems[] = {1, 1, 1, 1, 4, 6, 3, 0};
    size_t in_bytes = (size_t)n * in_elems[op] * 4, out_bytes;
    if (op <= ASH_CU_SUM) out_bytes = (size_t)(n / cu::kGroup) * 4;
    else if (op == ASH_CU_DEPTH) out_bytes = (size_t)n * 16;
    else if (op == ASH_CU_PLANES) { out_bytes = 96; n = 6; }
    else out_bytes = (size_t)n * 4;
    CuParams p{};
    if (params18) for (int i = 0; i < 18; ++i) p.f[i] = params18[i];
    float *d_in = nullptr, *d_out = nullptr;
    int rc = ASH_OK;
    if (in_bytes && cudaMalloc(&d_in, in_bytes) != cudaSuccess) return ASH_ERR_OUT_OF_MEMORY;
    if (cudaMalloc(&d_out, out_bytes) != cudaSuccess) { cudaFree(d_in); return ASH_ERR_OUT_OF_MEMORY; }
    if (in_bytes && cudaMemcpy(d_in, in, in_bytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = ASH_ERR_DEVICE;
    if (rc == ASH_OK) {
        k_compute_utils<<<(n + cu::kGroup - 1) / cu::kGroup, cu::kGroup>>>(op, d_in, n, p, d_out);
        if (cudaGetLastError() != cudaSuccess || cudaMemcpy(out, d_out, out_bytes, cudaMemcpyDeviceToHost) != cudaSuccess) rc = ASH_ERR_DEVICE;
    }
    cudaFree(d_in); cudaFree(d_out);
    return rc;
}
