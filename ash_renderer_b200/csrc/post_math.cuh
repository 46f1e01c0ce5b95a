// post_math.cuh — device helpers shared by the general (post.cu) and fast (post_fast.cu) post kernels.
#pragma once
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

struct Tex16 { const uint2* p; int W, H; };

__device__ __forceinline__ V4 fetch16(const Tex16& t, int x, int y) {
    return unpack_rgba16f(__ldg(t.p + (size_t)y * t.W + x));
}
__device__ __forceinline__ V4 fetch16_clamped(const Tex16& t, int x, int y) {
    return fetch16(t, min(max(x, 0), t.W - 1), min(max(y, 0), t.H - 1));
}
__device__ __forceinline__ V4 texture16(const Tex16& t, float u, float v) {
    int x0, x1, y0, y1; float fx, fy;
    bilinear_setup(u, t.W, x0, x1, fx);
    bilinear_setup(v, t.H, y0, y1, fy);
    const V4 a = fetch16(t, x0, y0), b = fetch16(t, x1, y0), c = fetch16(t, x0, y1), d = fetch16(t, x1, y1);
    const V4 top = {lerp1(a.x, b.x, fx), lerp1(a.y, b.y, fx), lerp1(a.z, b.z, fx), lerp1(a.w, b.w, fx)};
    const V4 bot = {lerp1(c.x, d.x, fx), lerp1(c.y, d.y, fx), lerp1(c.z, d.z, fx), lerp1(c.w, d.w, fx)};
    return {lerp1(top.x, bot.x, fy), lerp1(top.y, bot.y, fy), lerp1(top.z, bot.z, fy), lerp1(top.w, bot.w, fy)};
}
__device__ __forceinline__ V3 rgb(V4 v) { return {v.x, v.y, v.z}; }

// ---------------------------------------------------------------- TAA (taa_resolve.comp)
__device__ __forceinline__ V3 rgb_to_ycocg(V3 c) {                        // :34-40
    return {fadd(fadd(fmul(0.25f, c.x), fmul(0.5f, c.y)), fmul(0.25f, c.z)),
            fsub(fmul(0.5f, c.x), fmul(0.5f, c.z)),
            fsub(fadd(fmul(-0.25f, c.x), fmul(0.5f, c.y)), fmul(0.25f, c.z))};
}
__device__ __forceinline__ V3 ycocg_to_rgb(V3 c) {                        // :43-52
    return {fsub(fadd(c.x, c.y), c.z), fadd(c.x, c.z), fsub(fsub(c.x, c.y), c.z)};
}
__device__ __forceinline__ V3 taa_tonemap(V3 c) {                         // :55-57
    return {fdiv(c.x, fadd(1.0f, c.x)), fdiv(c.y, fadd(1.0f, c.y)), fdiv(c.z, fadd(1.0f, c.z))};
}
__device__ __forceinline__ V3 taa_untonemap(V3 c) {                       // :59-61
    return {fdiv(c.x, fadd(fsub(1.0f, c.x), 0.0001f)), fdiv(c.y, fadd(fsub(1.0f, c.y), 0.0001f)),
            fdiv(c.z, fadd(fsub(1.0f, c.z), 0.0001f))};
}
__device__ __forceinline__ V3 clip_aabb(V3 mn, V3 mx, V3 hist) {          // :64-77
    const V3 center = scale3(add3(mx, mn), 0.5f);
    const V3 half = adds3(scale3(sub3(mx, mn), 0.5f), 0.001f);
    const V3 delta = sub3(hist, center);
    const V3 v = div3(delta, half);
    const float maxv = gmax(gmax(fabsf(v.x), fabsf(v.y)), fabsf(v.z));
    if (maxv > 1.0f) return add3(center, divs3(delta, maxv));
    return hist;
}

__device__ __forceinline__ V3 soft_threshold(V3 c, float threshold) {
    const float knee = 0.5f;                                                               // :32
    const float br = gmax(c.x, gmax(c.y, c.z));
    float soft = fadd(fsub(br, threshold), knee);
    soft = gclamp(soft, 0.0f, fmul(2.0f, knee));
    soft = fdiv(fmul(soft, soft), fadd(fmul(4.0f, knee), 0.00001f));
    float contrib = gmax(soft, fsub(br, threshold));
    contrib = fdiv(contrib, gmax(br, 0.00001f));
    return scale3(c, contrib);
}

__device__ __forceinline__ float aces(float x) {                                           // :20-27
    return gclamp(fdiv(fmul(x, fadd(fmul(2.51f, x), 0.03f)), fadd(fmul(x, fadd(fmul(2.43f, x), 0.59f)), 0.14f)), 0.0f, 1.0f);
}
// TonemapOperator (features/tonemapping.rs:9-20); PIN-13 in include/ashcuda.h.  x = exposed HDR channel.
__device__ __forceinline__ float hable(float x) {        // ((x (A x + C B) + D E) / (x (A x + B) + D F)) - E / F
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return fsub(fdiv(fadd(fmul(x, fadd(fmul(A, x), fmul(C, B))), fmul(D, E)), fadd(fmul(x, fadd(fmul(A, x), B)), fmul(D, F))), fdiv(E, F));
}
__device__ __forceinline__ float tonemap_curve(float x, uint32_t op) {
    switch (op) {
        case ASH_TONEMAP_REINHARD:   return gclamp(fdiv(x, fadd(1.0f, x)), 0.0f, 1.0f);
        case ASH_TONEMAP_UNCHARTED2: return gclamp(fdiv(hable(fmul(2.0f, x)), hable(11.2f)), 0.0f, 1.0f);
        case ASH_TONEMAP_NONE:       return gclamp(x, 0.0f, 1.0f);
        default:                     return aces(x);
    }
}
__device__ __forceinline__ uint32_t to_unorm8(float l) {
    return (uint32_t)floorf(fadd(fmul(gclamp(l, 0.0f, 1.0f), 255.0f), 0.5f));
}

// ---------------------------------------------------------------- 1:2 tent upsample tap tables (bloom_upsample.frag:16-41, PIN-9)
// Destination pixel x = 2i + p samples the source at i - 1/4 + p/2 + m/2 for tap offsets m = -2..2 (in units of
// half a source texel): window column (origin i-2) c0 and weight g in {1/4, 3/4}; see post_fast.cu.
__host__ __device__ constexpr int up_fd4(int q) { return q >= 0 ? q / 4 : -((-q + 3) / 4); }
__host__ __device__ constexpr int up_c0(int p, int m) { return up_fd4(2 * p - 1 + 2 * m) + 2; }
__host__ __device__ constexpr bool up_g75(int p, int m) { return (2 * p - 1 + 2 * m) - 4 * up_fd4(2 * p - 1 + 2 * m) == 3; }

// ---------------------------------------------------------------- cp.async staging (LDGSTS)
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

}  // namespace ash
