// post_fast.cuh — the per-texel / per-tile device functions of post_fast.cu that more than one kernel uses: the one-shot
// kernels of post_fast.cu and the persistent bloom-tail kernel (post_tail.cu).  See post_fast.cu for why they are
// bit-identical to the general kernels of post.cu.  NC: post_math.cuh, ld_texel.
#pragma once
#include "post_math.cuh"

namespace ash {

__device__ __forceinline__ float fma_exact(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ V3 unpack_rgb(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

// bloom_downsample.frag:16-108, source extent == 2 x destination extent: the stored texel of destination pixel (x, y)
// from its 4x4 source window, fetched with 12 independent loads (8 B, 16 B, 8 B per row) issued back to back.
template <bool NC>
__device__ __forceinline__ uint2 down_fast_texel(const uint2* __restrict__ src, int sw, int sh, int x, int y) {
    // 4x4 window: window (i,j) = source texel (2x-1+i, 2y-1+j), CLAMP_TO_EDGE
    const int xl = max(2 * x - 1, 0), xr = min(2 * x + 2, sw - 1);
    uint2 e0[4], e3[4]; uint4 m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint2* row = src + (size_t)min(max(2 * y - 1 + j, 0), sh - 1) * sw;
        e0[j] = ld_texel<NC>(row + xl);
        m[j] = ld_texel2<NC>(reinterpret_cast<const uint4*>(row + 2 * x));       // texels 2x, 2x+1 (16-byte aligned: sw is even)
        e3[j] = ld_texel<NC>(row + xr);
    }
    V3 hs[4][3];     // hs[j][i] = t[j][i] + t[j][i+1]            (= 2 * horizontal lerp at 1/2)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const V3 t0 = unpack_rgb(e0[j]), t1 = unpack_rgb(make_uint2(m[j].x, m[j].y));
        const V3 t2 = unpack_rgb(make_uint2(m[j].z, m[j].w)), t3 = unpack_rgb(e3[j]);
        hs[j][0] = add3(t0, t1); hs[j][1] = add3(t1, t2); hs[j][2] = add3(t2, t3);
    }
    V3 res = {0.0f, 0.0f, 0.0f};
    float sum = 0.0f;
    // tap(i,j): 4 * bilinear sample whose 2x2 footprint starts at window (i,j); s = tap/4
    auto tap = [&](int i, int j, float wt) {
        const V3 t4 = add3(hs[j][i], hs[j + 1][i]);
        const float luma4 = fadd(fadd(fmul(t4.x, 0.2126f), fmul(t4.y, 0.7152f)), fmul(t4.z, 0.0722f));
        const float w = fdiv_safe(1.0f, fma_exact(luma4, 0.25f, 1.0f));     // 1 / (1 + luma), luma in [0, 65504]   :87-101
        const float k = 0.25f * wt;                                          // power of two
        res = {fma_exact(fmul(t4.x, w), k, res.x), fma_exact(fmul(t4.y, w), k, res.y), fma_exact(fmul(t4.z, w), k, res.z)};
        sum = fma_exact(w, wt, sum);
    };
    tap(1, 1, 0.25f);                                                        // centre
    tap(0, 0, 0.0625f); tap(2, 0, 0.0625f); tap(0, 2, 0.0625f); tap(2, 2, 0.0625f);   // a b c d   :25-28
    tap(0, 1, 0.125f);  tap(2, 1, 0.125f);  tap(1, 0, 0.125f);  tap(1, 2, 0.125f);    // e f g h   :31-34
    // sum in [1.5e-5, 1], res in [0, 65504]: safe domain                                          :107
    return pack_rgba16f(fdiv_safe(res.x, sum), fdiv_safe(res.y, sum), fdiv_safe(res.z, sum), 1.0f);
}


// =====================================================================================================
// 1:2 tent upsample + additive blend (bloom_upsample.frag:16-41, PIN-9), destination == 2 x source
// =====================================================================================================
// 4 * lerp(a, b, g): g = 3/4 -> a + 3b ; g = 1/4 -> 3a + b   (3 * fp16 texel is exact)
template <bool G75>
__device__ __forceinline__ V3 hlerp4(V3 a, V3 b) {
    if (G75) return {fma_exact(b.x, 3.0f, a.x), fma_exact(b.y, 3.0f, a.y), fma_exact(b.z, 3.0f, a.z)};
    return {fma_exact(a.x, 3.0f, b.x), fma_exact(a.y, 3.0f, b.y), fma_exact(a.z, 3.0f, b.z)};
}
// 16 * lerp(H0, H1, g) from h = 4H: g = 3/4 -> h0 + RN(3 h1) ; g = 1/4 -> RN(3 h0) + h1
template <bool G75>
__device__ __forceinline__ V3 vlerp16(V3 h0, V3 h1) {
    if (G75) return {fadd(h0.x, fmul(h1.x, 3.0f)), fadd(h0.y, fmul(h1.y, 3.0f)), fadd(h0.z, fmul(h1.z, 3.0f))};
    return {fadd(fmul(h0.x, 3.0f), h1.x), fadd(fmul(h0.y, 3.0f), h1.y), fadd(fmul(h0.z, 3.0f), h1.z)};
}

struct UpH {            // horizontal lerps (x4) of one destination-column parity
    V3 m2[3], m1[3], z[5], p1[3], p2[3];   // tap m = -2,-1: window rows 1..3 ; m = 0: rows 0..4 ; m = +1,+2: rows 1..3
};

template <int PX>
__device__ __forceinline__ void up_hlerps(const uint2 (*s_t)[20], int tj, int ti, UpH& h) {
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        V3 t[5];
#pragma unroll
        for (int c = 0; c < 4; ++c) t[c + PX] = unpack_rgb(s_t[tj + r][ti + c + PX]);   // window cols PX..PX+3
        h.z[r] = hlerp4<up_g75(PX, 0)>(t[up_c0(PX, 0)], t[up_c0(PX, 0) + 1]);
        if (r >= 1 && r <= 3) {
            h.m2[r - 1] = hlerp4<up_g75(PX, -2)>(t[up_c0(PX, -2)], t[up_c0(PX, -2) + 1]);
            h.m1[r - 1] = hlerp4<up_g75(PX, -1)>(t[up_c0(PX, -1)], t[up_c0(PX, -1) + 1]);
            h.p1[r - 1] = hlerp4<up_g75(PX, 1)>(t[up_c0(PX, 1)], t[up_c0(PX, 1) + 1]);
            h.p2[r - 1] = hlerp4<up_g75(PX, 2)>(t[up_c0(PX, 2)], t[up_c0(PX, 2) + 1]);
        }
    }
}

// 256 * (tent sum / 16), i.e. R16 of the file header; rows via the same (c0, g) tables with p = PY
template <int PY>
__device__ __forceinline__ V3 up_combine(const UpH& h) {
#define ASH_UPV(arr, base, n) vlerp16<up_g75(PY, n)>(arr[up_c0(PY, n) - (base)], arr[up_c0(PY, n) + 1 - (base)])
    const V3 s0 = ASH_UPV(h.m2, 1, 0);      // (-2d, 0)      bloom_upsample.frag:21
    const V3 s1 = ASH_UPV(h.m1, 1, -1);     // (-d, -d)
    const V3 s2 = ASH_UPV(h.z, 0, -2);      // (0, -2d)
    const V3 s3 = ASH_UPV(h.p1, 1, -1);     // (d, -d)
    const V3 s4 = ASH_UPV(h.p2, 1, 0);      // (2d, 0)
    const V3 s5 = ASH_UPV(h.p1, 1, 1);      // (d, d)
    const V3 s6 = ASH_UPV(h.z, 0, 2);       // (0, 2d)
    const V3 s7 = ASH_UPV(h.m1, 1, 1);      // (-d, d)
    const V3 s8 = ASH_UPV(h.z, 0, 0);       // (0, 0)
#undef ASH_UPV
    const V3 q = add3(add3(add3(s1, s3), s5), s7);                           // :33
    const V3 p = add3(add3(add3(s0, s2), s4), s6);                           // :34
    const V3 r = {fadd(fadd(fmul(s8.x, 4.0f), fmul(q.x, 2.0f)), p.x), fadd(fadd(fmul(s8.y, 4.0f), fmul(q.y, 2.0f)), p.y),
                  fadd(fadd(fmul(s8.z, 4.0f), fmul(q.z, 2.0f)), p.z)};       // :32-34
    return r;
}

// ---- tonemapping.frag:20-48 for the 8-bit target only (PIN-11: within 1 LSB of the reference): FMA contraction,
//      MUFU reciprocal / lg2 / ex2.  Not used for any fp16 / fp32 output.
__device__ __forceinline__ float lg2_ftz(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ex2_ftz(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rcp_ftz(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ uint32_t tonemap8(float c, float bloom, const AshPostProcessPushConstants& pc, float inv_gamma) {
    const float x = __fmaf_rn(bloom, pc.bloom_intensity, c) * pc.exposure;              // :38, :41
    const float n = x * __fmaf_rn(2.51f, x, 0.03f);                                      // :20-27
    const float d = __fmaf_rn(x, __fmaf_rn(2.43f, x, 0.59f), 0.14f);
    const float a = __saturatef(n * rcp_ftz(d));
    const float g = ex2_ftz(inv_gamma * lg2_ftz(a));                                     // pow(a, 1/gamma), a in [0,1]   :47
    return (uint32_t)__float2int_rd(__fmaf_rn(__saturatef(g), 255.0f, 0.5f));            // PIN-11
}
__device__ __forceinline__ uint32_t tonemap_pack8(V3 c, V3 bloom, const AshPostProcessPushConstants& pc, float inv_gamma) {
    return tonemap8(c.x, bloom.x, pc, inv_gamma) | (tonemap8(c.y, bloom.y, pc, inv_gamma) << 8) |
           (tonemap8(c.z, bloom.z, pc, inv_gamma) << 16) | 0xFF000000u;
}

// One tile: source texels [i0, i0 + 16) x [j0, j0 + 16) -> 32x32 destination pixels, CTA of 256 threads, s_t = the CTA's
// staged source tile.  Ends after the last store; callers that reuse s_t put a barrier before the next tile.
template <bool TONEMAP, bool STORE_U, bool NC>
__device__ __forceinline__ void up_fast_tile(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, float intensity,
                                             const uint2* __restrict__ hdr, uint32_t* __restrict__ ldr8,
                                             const AshPostProcessPushConstants& pc, int i0, int j0, uint2 (*s_t)[20]) {
    const int ti = threadIdx.x & 15, tj = threadIdx.x >> 4;
    const int i = i0 + ti, j = j0 + tj;
    const bool active = i < sw && j < sh;
    const int dw = 2 * sw;
    // every global load of the thread is issued up front: the source tile and its own 2x2 destination block
    uint2 st[2];
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const int idx = threadIdx.x + it * 256;
        if (idx < 400) {
            const int r = idx / 20, c = idx - r * 20;
            st[it] = ld_texel<NC>(src + (size_t)min(max(j0 - 2 + r, 0), sh - 1) * sw + min(max(i0 - 2 + c, 0), sw - 1));   // CLAMP_TO_EDGE
        }
    }
    uint4 dpre[2], cpre[2];
    if (active) {
#pragma unroll
        for (int py = 0; py < 2; ++py) {
            const size_t o = (size_t)(2 * j + py) * dw + 2 * i;               // 16-byte aligned pair of texels
            dpre[py] = *reinterpret_cast<const uint4*>(dst + o);
            if (TONEMAP) cpre[py] = __ldg(reinterpret_cast<const uint4*>(hdr + o));
        }
    }
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const int idx = threadIdx.x + it * 256;
        if (idx < 400) s_t[idx / 20][idx % 20] = st[it];
    }
    __syncthreads();
    if (!active) return;
    const float k = fmul(intensity, 0.00390625f);                 // (R16 / 256) * intensity == R16 * (intensity / 256)
    const float inv_gamma = TONEMAP ? __fdividef(1.0f, pc.gamma) : 0.0f;

    V3 up[2][2];     // [py][px] upsampled value, before the additive blend
    {
        UpH h;
        up_hlerps<0>(s_t, tj, ti, h);
        const V3 a = up_combine<0>(h), b = up_combine<1>(h);
        up[0][0] = {fmul(a.x, k), fmul(a.y, k), fmul(a.z, k)};
        up[1][0] = {fmul(b.x, k), fmul(b.y, k), fmul(b.z, k)};
    }
    {
        UpH h;
        up_hlerps<1>(s_t, tj, ti, h);
        const V3 a = up_combine<0>(h), b = up_combine<1>(h);
        up[0][1] = {fmul(a.x, k), fmul(a.y, k), fmul(a.z, k)};
        up[1][1] = {fmul(b.x, k), fmul(b.y, k), fmul(b.z, k)};
    }
#pragma unroll
    for (int py = 0; py < 2; ++py) {
        const size_t o = (size_t)(2 * j + py) * dw + 2 * i;                   // 16-byte aligned pair of texels
        const uint4 d = dpre[py];
        const V3 d0 = unpack_rgb(make_uint2(d.x, d.y)), d1 = unpack_rgb(make_uint2(d.z, d.w));
        const uint2 u0 = pack_rgba16f(fadd(up[py][0].x, d0.x), fadd(up[py][0].y, d0.y), fadd(up[py][0].z, d0.z), 1.0f);
        const uint2 u1 = pack_rgba16f(fadd(up[py][1].x, d1.x), fadd(up[py][1].y, d1.y), fadd(up[py][1].z, d1.z), 1.0f);
        if (STORE_U) *reinterpret_cast<uint4*>(dst + o) = make_uint4(u0.x, u0.y, u1.x, u1.y);
        if (TONEMAP) {
            const uint4 c = cpre[py];
            const uint32_t a8 = tonemap_pack8(unpack_rgb(make_uint2(c.x, c.y)), unpack_rgb(u0), pc, inv_gamma);
            const uint32_t b8 = tonemap_pack8(unpack_rgb(make_uint2(c.z, c.w)), unpack_rgb(u1), pc, inv_gamma);
            *reinterpret_cast<uint2*>(ldr8 + o) = make_uint2(a8, b8);
        }
    }
}


}  // namespace ash
