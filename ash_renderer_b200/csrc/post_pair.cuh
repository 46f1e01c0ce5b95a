// post_pair.cuh — the post-chain math of post_math.cuh / post_fast.cu on TWO pixels at once (P2 lanes,
// ash_f32x2.cuh).  Every function is the scalar function's operation sequence, lane by lane.
#pragma once
#include "ash_f32x2.cuh"
#include "post_math.cuh"

namespace ash {

struct V3P { P2 x, y, z; };

__device__ __forceinline__ V3P pk3(V3 a, V3 b) { return {pk(a.x, b.x), pk(a.y, b.y), pk(a.z, b.z)}; }
__device__ __forceinline__ V3 lo3(V3P a) { return {lo(a.x), lo(a.y), lo(a.z)}; }
__device__ __forceinline__ V3 hi3(V3P a) { return {hi(a.x), hi(a.y), hi(a.z)}; }
__device__ __forceinline__ V3P add3(V3P a, V3P b) { return {fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)}; }
__device__ __forceinline__ V3P sub3(V3P a, V3P b) { return {fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)}; }
__device__ __forceinline__ V3P scale3(V3P a, float s) { const P2 k = splat(s); return {fmul(a.x, k), fmul(a.y, k), fmul(a.z, k)}; }
__device__ __forceinline__ V3P scale3(V3P a, P2 k) { return {fmul(a.x, k), fmul(a.y, k), fmul(a.z, k)}; }
__device__ __forceinline__ V3P adds3(V3P a, float s) { const P2 k = splat(s); return {fadd(a.x, k), fadd(a.y, k), fadd(a.z, k)}; }
__device__ __forceinline__ V3P sel3(bool c_lo, bool c_hi, V3P a, V3P b) { return {sel(c_lo, c_hi, a.x, b.x), sel(c_lo, c_hi, a.y, b.y), sel(c_lo, c_hi, a.z, b.z)}; }

// Every product here is a scaling by a power of two, i.e. exact (the operands are tonemapped colours, zero or
// >= 2^-41 in magnitude: no fp32 subnormals), so a sum of two products rounds once whether or not the second
// product is folded into an FMA: 7 instructions instead of 13, same bits as the scalar rgb_to_ycocg.
__device__ __forceinline__ V3P rgb_to_ycocg(V3P c) {                       // taa_resolve.comp:34-40
    const P2 hy = fmul(0.5f, c.y);
    return {ffma(c.z, splat(0.25f), ffma(c.x, splat(0.25f), hy)),          // (0.25 x + 0.5 y) + 0.25 z
            ffma(c.z, splat(-0.5f), fmul(0.5f, c.x)),                      // 0.5 x - 0.5 z
            ffma(c.z, splat(-0.25f), ffma(c.x, splat(-0.25f), hy))};       // (-0.25 x + 0.5 y) - 0.25 z
}
__device__ __forceinline__ V3P ycocg_to_rgb(V3P c) {                       // :43-52
    return {fsub(fadd(c.x, c.y), c.z), fadd(c.x, c.z), fsub(fsub(c.x, c.y), c.z)};
}
__device__ __forceinline__ V3P taa_tonemap_safe(V3P c) {                   // :55-57, c in [0, 65504]
    return {fdiv_safe(c.x, fadd(1.0f, c.x)), fdiv_safe(c.y, fadd(1.0f, c.y)), fdiv_safe(c.z, fadd(1.0f, c.z))};
}

// :59-61.  The divisor (1 - c) + 1e-4 can reach 0, so the scalar kernel uses the guarded IEEE division.
// Here both lanes take the unguarded sequence and a lane is redone with __fdiv_rn when its operands leave
// the range in which that sequence is proven identical (fdiv_safe, ash_math.cuh): |c| <= 16 and
// |divisor| >= 2^-20 on all three channels (c is zero or >= 2^-100 in magnitude: sums and differences of
// quotients of fp16 texels).  A NaN lane yields NaN on either path (stored as the canonical fp16 NaN).
__device__ __forceinline__ V3P taa_untonemap_pair(V3P c) {
    const V3P d = {fadd(fsub(1.0f, c.x), 0.0001f), fadd(fsub(1.0f, c.y), 0.0001f), fadd(fsub(1.0f, c.z), 0.0001f)};
    V3P q = {fdiv_safe(c.x, d.x), fdiv_safe(c.y, d.y), fdiv_safe(c.z, d.z)};
    const float kMinDiv = 9.5367431640625e-07f;   // 2^-20
    const bool ok_lo = fmaxf(fmaxf(fabsf(lo(c.x)), fabsf(lo(c.y))), fabsf(lo(c.z))) <= 16.0f &&
                       fminf(fminf(fabsf(lo(d.x)), fabsf(lo(d.y))), fabsf(lo(d.z))) >= kMinDiv;
    const bool ok_hi = fmaxf(fmaxf(fabsf(hi(c.x)), fabsf(hi(c.y))), fabsf(hi(c.z))) <= 16.0f &&
                       fminf(fminf(fabsf(hi(d.x)), fabsf(hi(d.y))), fabsf(hi(d.z))) >= kMinDiv;
    if (!(ok_lo && ok_hi)) {
        const V3 l = ok_lo ? lo3(q) : V3{fdiv(lo(c.x), lo(d.x)), fdiv(lo(c.y), lo(d.y)), fdiv(lo(c.z), lo(d.z))};
        const V3 h = ok_hi ? hi3(q) : V3{fdiv(hi(c.x), hi(d.x)), fdiv(hi(c.y), hi(d.y)), fdiv(hi(c.z), hi(d.z))};
        q = pk3(l, h);
    }
    return q;
}

// clip_aabb_fast (post_fast.cu) per lane: a lane is clipped iff |delta| > half on some channel
__device__ __forceinline__ V3P clip_aabb_pair(V3P mn, V3P mx, V3P hist) {   // :64-77
    const V3P center = scale3(add3(mx, mn), 0.5f);
    const V3P dm = sub3(mx, mn);                                           // 0.5 * dm is exact: fma == fmul + fadd
    const V3P half = {ffma(dm.x, splat(0.5f), splat(0.001f)), ffma(dm.y, splat(0.5f), splat(0.001f)), ffma(dm.z, splat(0.5f), splat(0.001f))};
    const V3P delta = sub3(hist, center);
    const bool c_lo = fabsf(lo(delta.x)) > lo(half.x) || fabsf(lo(delta.y)) > lo(half.y) || fabsf(lo(delta.z)) > lo(half.z);
    const bool c_hi = fabsf(hi(delta.x)) > hi(half.x) || fabsf(hi(delta.y)) > hi(half.y) || fabsf(hi(delta.z)) > hi(half.z);
    if (c_lo || c_hi) {
        const V3P v = {fdiv_safe(delta.x, half.x), fdiv_safe(delta.y, half.y), fdiv_safe(delta.z, half.z)};
        // an unclipped lane may have maxv == 0 (its quotient is discarded by sel3 below)
        const P2 maxv = pk(fmaxf(fmaxf(fabsf(lo(v.x)), fabsf(lo(v.y))), fabsf(lo(v.z))),
                           fmaxf(fmaxf(fabsf(hi(v.x)), fabsf(hi(v.y))), fabsf(hi(v.z))));
        const V3P clipped = add3(center, V3P{fdiv_safe(delta.x, maxv), fdiv_safe(delta.y, maxv), fdiv_safe(delta.z, maxv)});
        return sel3(c_lo, c_hi, clipped, hist);
    }
    return hist;
}

__device__ __forceinline__ V3P soft_threshold_pair(V3P c, float threshold) {   // bloom_threshold.frag:18-27
    const float knee = 0.5f;
    const P2 br = pk(gmax(lo(c.x), gmax(lo(c.y), lo(c.z))), gmax(hi(c.x), gmax(hi(c.y), hi(c.z))));
    P2 soft = fadd(fsub(br, threshold), knee);
    soft = gclamp(soft, 0.0f, fmul(2.0f, knee));
    soft = fdiv_safe(fmul(soft, soft), fadd(fmul(4.0f, knee), 0.00001f));
    P2 contrib = gmax(soft, fsub(br, threshold));
    contrib = fdiv_safe(contrib, gmax(br, 0.00001f));
    return scale3(c, contrib);
}

}  // namespace ash
