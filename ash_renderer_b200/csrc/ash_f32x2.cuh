// ash_f32x2.cuh — two IEEE fp32 lanes in one 64-bit register pair (Blackwell FADD2 / FMUL2 / FFMA2).
//
// sm_100a adds packed fp32 arithmetic: add/sub/mul/fma.rn.f32x2 round each lane exactly like the scalar
// .rn instruction (no FTZ), so a kernel that evaluates the pinned IEEE sequence (PIN-1) for TWO pixels in
// lock-step keeps bit-exact parity while issuing one instruction instead of two.  The FMA pipe itself
// is not faster (tools/microbench/f32x2.cu: 36.3 vs 34.7 TFMA/s) but the post kernels are bound by
// instruction ISSUE (ncu: ~80 % issue-active, FMA pipe 45 %), and half of what they issue is fp32
// add / mul / fma.  ptxas folds per-lane negation into the FFMA2 operand modifier, so neg() is free.
//
// Only arithmetic is packed; min / max / compare / select / conversions stay scalar per lane (there is
// no packed form) and go through lo() / hi() / pk(), which ptxas coalesces into register pairing.
#pragma once
#include "ash_math.cuh"

namespace ash {

struct P2 { unsigned long long v; };

__device__ __forceinline__ P2 pk(float lo, float hi) { P2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ P2 splat(float s) { return pk(s, s); }
__device__ __forceinline__ float lo(P2 a) { float l; [[maybe_unused]] float h; asm("mov.b64 {%0,%1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v)); return l; }
__device__ __forceinline__ float hi(P2 a) { [[maybe_unused]] float l; float h; asm("mov.b64 {%0,%1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v)); return h; }

__device__ __forceinline__ P2 fadd(P2 a, P2 b) { P2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 fsub(P2 a, P2 b) { P2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
// ptxas 12.9 contracts mul.rn.f32x2 + add/sub.rn.f32x2 into FFMA2 although both carry .rn (it does not for
// the scalar forms, and neither -fmad=false nor spelling the product fma(a, b, -0.0) stops it), which
// would change the rounding.  The product is therefore written as fma(a, b, Z) with Z = (-0.0, -0.0) read
// from constant memory: a*b + (-0) == RN(a*b) for every a*b (signed zeros, subnormals, inf, NaN
// included), ptxas cannot see the value, and an FFMA2 cannot be fused into the add that follows.
// Same issue cost as FMUL2 (Z sits in a uniform register).
static __constant__ unsigned long long c_negzero2 = 0x8000000080000000ull;
__device__ __forceinline__ P2 fmul(P2 a, P2 b) { P2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c_negzero2)); return r; }
// single rounding: only where the scalar code has an FMA too, or where the product is exact (fma_exact)
__device__ __forceinline__ P2 ffma(P2 a, P2 b, P2 c) { P2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ P2 neg(P2 a) { return pk(-lo(a), -hi(a)); }
// mixed forms (scalar broadcast on the right)
__device__ __forceinline__ P2 fadd(P2 a, float s) { return fadd(a, splat(s)); }
__device__ __forceinline__ P2 fsub(P2 a, float s) { return fsub(a, splat(s)); }
__device__ __forceinline__ P2 fsub(float s, P2 a) { return fsub(splat(s), a); }
__device__ __forceinline__ P2 fmul(P2 a, float s) { return fmul(a, splat(s)); }
__device__ __forceinline__ P2 fadd(float s, P2 a) { return fadd(splat(s), a); }
__device__ __forceinline__ P2 fmul(float s, P2 a) { return fmul(splat(s), a); }

// per-lane scalar helpers with the GLSL semantics of ash_math.cuh
__device__ __forceinline__ P2 gmin(P2 a, P2 b) { return pk(gmin(lo(a), lo(b)), gmin(hi(a), hi(b))); }
__device__ __forceinline__ P2 gmax(P2 a, P2 b) { return pk(gmax(lo(a), lo(b)), gmax(hi(a), hi(b))); }
__device__ __forceinline__ P2 gmax(P2 a, float s) { return pk(gmax(lo(a), s), gmax(hi(a), s)); }
__device__ __forceinline__ P2 gclamp(P2 x, float l, float h) { return pk(gclamp(lo(x), l, h), gclamp(hi(x), l, h)); }
__device__ __forceinline__ P2 sel(bool c_lo, bool c_hi, P2 a, P2 b) { return pk(c_lo ? lo(a) : lo(b), c_hi ? hi(a) : hi(b)); }

// fdiv_safe (ash_math.cuh) on both lanes: 2 MUFU.RCP + 5 FFMA2, the same operation sequence per lane
__device__ __forceinline__ P2 fdiv_safe(P2 a, P2 b) {
    float r0, r1;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(lo(b)));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(hi(b)));
    P2 r = pk(r0, r1);
    const P2 nb = neg(b);
    const P2 t = ffma(nb, r, splat(1.0f));
    r = ffma(r, t, r);
    const P2 q = ffma(a, r, splat(0.0f));
    const P2 e = ffma(nb, q, a);
    return ffma(r, e, q);
}
// fdiv_safe(1, b): the quotient step q = fma(1, r, 0) is r itself (r > 0 or r < 0, never -0), so it is dropped
__device__ __forceinline__ P2 frcp_safe(P2 b) {
    float r0, r1;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(lo(b)));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(hi(b)));
    P2 r = pk(r0, r1);
    const P2 nb = neg(b);
    const P2 t = ffma(nb, r, splat(1.0f));
    r = ffma(r, t, r);
    const P2 e = ffma(nb, r, splat(1.0f));
    return ffma(r, e, r);
}
__device__ __forceinline__ P2 fdiv_safe(float a, P2 b) { return fdiv_safe(splat(a), b); }
__device__ __forceinline__ P2 fdiv_safe(P2 a, float b) { return fdiv_safe(a, splat(b)); }

// scalar fma spelled like the packed one, so that templates read the same for float and P2
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float splat1(float s) { return s; }

}  // namespace ash
