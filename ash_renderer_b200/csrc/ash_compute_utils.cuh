// ash_compute_utils.cuh — shaders/compute_utils.glsl as a CUDA header (third item).
//
// The reference ships this GLSL library (workgroup reductions, a 256-wide exclusive scan, frustum plane
// helpers, depth helpers) but no shader includes it.  Each function below cites the GLSL it restates and
// returns bit-identical results: the float reductions keep the reference's pairing (element i with
// i + stride, stride = 128 .. 1), so parallel_sum rounds exactly like the shared-memory tree; the strides
// that stay inside warp 0 (16 .. 1) run on shuffles instead of shared memory + barriers, the cross-warp
// strides (128, 64, 32) need one shared-memory round each.  Workgroup = 256 threads (one CTA).
#pragma once
#include "ash_math.cuh"

namespace ash {
namespace cu {

constexpr int kGroup = 256;            // compute_utils.glsl:16 `shared float sharedData[256]`

// shared storage a caller provides once per CTA (the GLSL declares it at file scope, :16 and :73)
struct GroupScratch { float f[kGroup]; uint32_t u[kGroup]; };

// The reduction tree of parallelMin / parallelMax / parallelSum (:22-65) for any op(a, b) where the result
// of a step is op(data[i], data[i + stride]).  Valid in thread 0 (like the GLSL); every thread must call.
template <class Op>
__device__ __forceinline__ float group_reduce(GroupScratch& s, uint32_t lid, float value, Op op) {
    s.f[lid] = value;
    __syncthreads();
#pragma unroll
    for (uint32_t stride = 128; stride >= 32; stride >>= 1) {         // cross-warp strides: shared memory
        if (lid < stride) s.f[lid] = op(s.f[lid], s.f[lid + stride]);
        __syncthreads();
    }
    float v = s.f[lid & 31];                                           // warp 0 holds data[0..31]
    if (lid < 32) {
#pragma unroll
        for (uint32_t stride = 16; stride > 0; stride >>= 1) {         // data[i] = op(data[i], data[i + stride]), i < stride
            const float o = __shfl_down_sync(0xFFFFFFFFu, v, stride);
            if (lid < stride) v = op(v, o);
        }
    }
    return v;                                                          // thread 0: data[0]
}
struct OpMin { __device__ float operator()(float a, float b) const { return gmin(a, b); } };   // min(x, y), :29
struct OpMax { __device__ float operator()(float a, float b) const { return gmax(a, b); } };   // :44
struct OpSum { __device__ float operator()(float a, float b) const { return fadd(a, b); } };   // :58
__device__ __forceinline__ float parallel_min(GroupScratch& s, uint32_t lid, float v) { return group_reduce(s, lid, v, OpMin()); }
__device__ __forceinline__ float parallel_max(GroupScratch& s, uint32_t lid, float v) { return group_reduce(s, lid, v, OpMax()); }
__device__ __forceinline__ float parallel_sum(GroupScratch& s, uint32_t lid, float v) { return group_reduce(s, lid, v, OpSum()); }

// exclusivePrefixSum (:80-111).  Integer addition is associative, so any correct scan is bit-identical to
// the Blelloch up-sweep / down-sweep of the GLSL: warp scans by shuffle + one shared round for the 8 warp
// totals (2 barriers instead of 17).
__device__ __forceinline__ uint32_t exclusive_prefix_sum(GroupScratch& s, uint32_t lid, uint32_t value) {
    uint32_t incl = value;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((lid & 31) >= (uint32_t)d) incl += o;
    }
    if ((lid & 31) == 31) s.u[lid >> 5] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (uint32_t w = 0; w < (lid >> 5); ++w) base += s.u[w];
    __syncthreads();                                                   // s.u may be reused by the caller
    return base + incl - value;
}

// ---- frustum culling helpers (:117-191)
struct Plane { V3 normal; float distance; };

__device__ __forceinline__ Plane extract_plane(const M4& m, int face) {             // :126-160
    V4 row;
    const V4 &c0 = m.c[0], &c1 = m.c[1], &c2 = m.c[2], &c3 = m.c[3];
    switch (face) {
        case 0: row = {fadd(c3.x, c0.x), fadd(c3.y, c0.y), fadd(c3.z, c0.z), fadd(c3.w, c0.w)}; break;   // left
        case 1: row = {fsub(c3.x, c0.x), fsub(c3.y, c0.y), fsub(c3.z, c0.z), fsub(c3.w, c0.w)}; break;   // right
        case 2: row = {fadd(c3.x, c1.x), fadd(c3.y, c1.y), fadd(c3.z, c1.z), fadd(c3.w, c1.w)}; break;   // bottom
        case 3: row = {fsub(c3.x, c1.x), fsub(c3.y, c1.y), fsub(c3.z, c1.z), fsub(c3.w, c1.w)}; break;   // top
        case 4: row = {fadd(c3.x, c2.x), fadd(c3.y, c2.y), fadd(c3.z, c2.z), fadd(c3.w, c2.w)}; break;   // near
        case 5: row = {fsub(c3.x, c2.x), fsub(c3.y, c2.y), fsub(c3.z, c2.z), fsub(c3.w, c2.w)}; break;   // far
        default: row = {0.0f, 0.0f, 0.0f, 1.0f};
    }
    const float len = length3(v3(row.x, row.y, row.z));
    return {divs3(v3(row.x, row.y, row.z), len), fdiv(row.w, len)};
}
__device__ __forceinline__ bool sphere_outside_plane(V3 center, float radius, const Plane& p) {   // :164-167
    return fadd(dot3(p.normal, center), p.distance) < -radius;
}
__device__ __forceinline__ bool sphere_in_frustum(V3 center, float radius, const Plane planes[6]) {   // :171-178
    for (int i = 0; i < 6; ++i)
        if (sphere_outside_plane(center, radius, planes[i])) return false;
    return true;
}
// step(edge, x) = x < edge ? 0 : 1 ; mix(a, b, t) = a*(1-t) + b*t with t in {0, 1}
__device__ __forceinline__ float mix_step(float lo, float hi, float n) {
    const float t = n < 0.0f ? 0.0f : 1.0f;
    return lerp1(lo, hi, t);
}
__device__ __forceinline__ bool aabb_outside_plane(V3 mn, V3 mx, const Plane& p) {                  // :181-185
    const V3 pv = {mix_step(mn.x, mx.x, p.normal.x), mix_step(mn.y, mx.y, p.normal.y), mix_step(mn.z, mx.z, p.normal.z)};
    return fadd(dot3(p.normal, pv), p.distance) < 0.0f;
}
__device__ __forceinline__ bool aabb_in_frustum(V3 mn, V3 mx, const Plane planes[6]) {              // :188-195
    for (int i = 0; i < 6; ++i)
        if (aabb_outside_plane(mn, mx, planes[i])) return false;
    return true;
}

// ---- depth helpers (:201-215)
__device__ __forceinline__ float linearize_depth(float depth, float near_, float far_) {            // :205-207
    return fdiv(fmul(near_, far_), fadd(far_, fmul(depth, fsub(near_, far_))));
}
__device__ __forceinline__ float depth_to_view_z(float depth, float near_, float far_) { return -linearize_depth(depth, near_, far_); }
__device__ __forceinline__ V3 uv_depth_to_view(float u, float v, float depth, const M4& inv_proj) {  // :215-219
    const V4 q = mul4(inv_proj, fsub(fmul(u, 2.0f), 1.0f), fsub(fmul(v, 2.0f), 1.0f), depth, 1.0f);
    return {fdiv(q.x, q.w), fdiv(q.y, q.w), fdiv(q.z, q.w)};
}

}  // namespace cu
}  // namespace ash
