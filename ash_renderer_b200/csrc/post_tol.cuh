// post_tol.cuh — arithmetic of the TOLERANCE precision mode (AshCudaConfig.precision = FAST) shared by its kernels
// (post_taa_tol.cu, post_down_box.cu, post_up_pair.cu): packed fp32 ops that ptxas may contract, MUFU reciprocals, and the
// bloom threshold of a pair of stored texels.
#pragma once
#include "ash_f32x2.cuh"
#include "post_math.cuh"

namespace ash {

namespace tol {
__device__ __forceinline__ P2 add(P2 a, P2 b) { P2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 sub(P2 a, P2 b) { P2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 mul(P2 a, P2 b) { P2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }   // ptxas may contract: welcome here
__device__ __forceinline__ P2 fma(P2 a, P2 b, P2 c) { P2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ P2 rcp(P2 x) { return pk(rcp(lo(x)), rcp(hi(x))); }
__device__ __forceinline__ P2 rcp_abs(P2 x) { return pk(rcp(fabsf(lo(x))), rcp(fabsf(hi(x)))); }
struct V3P { P2 x, y, z; };
__device__ __forceinline__ V3P pk3(V3 a, V3 b) { return {pk(a.x, b.x), pk(a.y, b.y), pk(a.z, b.z)}; }

// c / (1 + c)                                                               taa_resolve.comp:55-57
__device__ __forceinline__ V3P tonemap(V3P c) {
    const P2 one = splat(1.0f);
    return {mul(c.x, rcp(add(c.x, one))), mul(c.y, rcp(add(c.y, one))), mul(c.z, rcp(add(c.z, one)))};
}
// (0.25 r + 0.5 g + 0.25 b, 0.5 r - 0.5 b, -0.25 r + 0.5 g - 0.25 b)        :34-40
__device__ __forceinline__ V3P rgb_to_ycocg(V3P c) {
    const P2 rb = add(c.x, c.z), hg = mul(c.y, splat(0.5f));
    return {fma(rb, splat(0.25f), hg), mul(sub(c.x, c.z), splat(0.5f)), fma(rb, splat(-0.25f), hg)};
}
// (y + co - cg, y + cg, y - co - cg)                                        :43-52
__device__ __forceinline__ V3P ycocg_to_rgb(V3P c) {
    const P2 a = sub(c.x, c.z);
    return {add(a, c.y), add(c.x, c.z), sub(a, c.y)};
}
}  // namespace tol

__device__ __forceinline__ V3 tol_unpack(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

// bloom_threshold.frag:18-27 of two resolved texels AS STORED (fp16), result as stored (fp16, alpha = one): D_0.
// ONE definition: the TAA kernel that stores D_0 and the kernels that rebuild it on the fly from the resolve (the first
// downsample, the last upsample) must produce the same bits.
__device__ __forceinline__ void tol_threshold_texels(uint2 ra, uint2 rb, float threshold, float one, uint2& da, uint2& db) {
    using namespace tol;
    const V3 ta = tol_unpack(ra), tb = tol_unpack(rb);
    const P2 br = pk(fmaxf(fmaxf(ta.x, ta.y), ta.z), fmaxf(fmaxf(tb.x, tb.y), tb.z));
    const P2 over = sub(br, splat(threshold));                               // brightness - threshold
    const P2 sf = add(over, splat(0.5f));                                    // + knee
    const P2 sc = pk(fminf(fmaxf(lo(sf), 0.0f), 1.0f), fminf(fmaxf(hi(sf), 0.0f), 1.0f));
    const P2 soft = mul(mul(sc, sc), splat(1.0f / 2.00001f));                // soft^2 / (4 knee + 1e-5)
    const P2 ct = pk(fmaxf(lo(soft), lo(over)), fmaxf(hi(soft), hi(over)));
    const P2 w = mul(ct, pk(rcp(fmaxf(lo(br), 0.00001f)), rcp(fmaxf(hi(br), 0.00001f))));
    const tol::V3P t3 = tol::pk3(ta, tb);
    const P2 bx = mul(t3.x, w), by = mul(t3.y, w), bz = mul(t3.z, w);
    da = pack_rgba16f(lo(bx), lo(by), lo(bz), one);
    db = pack_rgba16f(hi(bx), hi(by), hi(bz), one);
}

}  // namespace ash
