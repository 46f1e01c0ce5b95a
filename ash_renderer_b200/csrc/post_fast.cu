// post_fast.cu — the post chain as it runs at 4K: tiled, fused kernels that are BIT-IDENTICAL to the
// general one-thread-per-pixel kernels of post.cu (and therefore to the reference semantics) but do
// a fraction of the work:
//
//   k_taa_fast           TAA resolve (+ fused bloom threshold): the tonemapped YCoCg neighbourhood is
//                        computed once per texel into shared memory instead of 9x per pixel; each
//                        thread resolves a 1x4 pixel strip and shares the row-wise min/max.
//   k_downsample_fast    2:1 Karis downsample: every tap of the pinned sampler lands exactly midway
//                        between 4 texels, so the 9 taps are 2x2 box sums over one 4x4 window.
//   k_upsample_fast      1:2 tent upsample + additive blend (+ fused ACES tonemap / 8-bit pack for the
//                        last level): weights are exactly 1/4, 3/4; each thread makes a 2x2 block and
//                        shares the horizontal lerps between its two rows.
//
// Why identical: (1) exact_taps.cu proves, per extent, that the sampler selects exactly the texels
// and weights assumed here; (2) scaling by a power of two commutes with IEEE rounding (no fp32
// subnormals can arise from fp16 texels), so factors 1/2, 1/4, 1/16 are deferred; (3) a product that
// is exactly representable (fp16 texel x 3, x * 2^k) may feed an FMA without changing the single
// rounding of the following add.  Every other operation is the pinned IEEE sequence of ash_math.cuh.
// Preconditions: colour texels are finite and carry no negative zeros.
#include "post_math.cuh"

namespace ash {

__device__ __forceinline__ float fma_exact(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ V3 unpack_rgb(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

// =====================================================================================================
// TAA resolve (+ bloom threshold) — taa_resolve.comp:83-167 (+ bloom_threshold.frag:18-35)
// No velocity buffer (PIN-12: velocity == 0 -> jitter-delta branch, uniform over the frame).
// CTA = 32x32 pixels, 256 threads, thread = 1x4 strip.
// =====================================================================================================
// Safe-domain helpers (operands bounded by construction: fdiv_safe, ash_math.cuh)
__device__ __forceinline__ V3 taa_tonemap_safe(V3 c) {      // c in [0, 65504]: 1 + c in [1, 65505]
    return {fdiv_safe(c.x, fadd(1.0f, c.x)), fdiv_safe(c.y, fadd(1.0f, c.y)), fdiv_safe(c.z, fadd(1.0f, c.z))};
}
__device__ __forceinline__ V3 soft_threshold_safe(V3 c, float threshold) {   // bloom_threshold.frag:18-27
    const float knee = 0.5f;
    const float br = gmax(c.x, gmax(c.y, c.z));
    float soft = fadd(fsub(br, threshold), knee);
    soft = gclamp(soft, 0.0f, fmul(2.0f, knee));
    soft = fdiv_safe(fmul(soft, soft), fadd(fmul(4.0f, knee), 0.00001f));      // [0,1] / 2.00001
    float contrib = gmax(soft, fsub(br, threshold));
    contrib = fdiv_safe(contrib, gmax(br, 0.00001f));                            // divisor in [1e-5, 65504]
    return scale3(c, contrib);
}

__device__ __forceinline__ V3 clip_aabb_fast(V3 mn, V3 mx, V3 hist) {       // :64-77
    const V3 center = scale3(add3(mx, mn), 0.5f);
    const V3 half = adds3(scale3(sub3(mx, mn), 0.5f), 0.001f);
    const V3 delta = sub3(hist, center);
    // RN(|d|/h) > 1  <=>  |d| > h  for h >= 0.001 (a float above h is at least h + ulp(h) > h(1 + 2^-24))
    if (fabsf(delta.x) > half.x || fabsf(delta.y) > half.y || fabsf(delta.z) > half.z) {
        // tonemapped YCoCg: |delta| <= 3, half in [0.001, 2], 1 < maxv <= 3000: safe domain
        const V3 v = {fdiv_safe(delta.x, half.x), fdiv_safe(delta.y, half.y), fdiv_safe(delta.z, half.z)};
        const float maxv = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fabsf(v.z));
        return add3(center, v3(fdiv_safe(delta.x, maxv), fdiv_safe(delta.y, maxv), fdiv_safe(delta.z, maxv)));
    }
    return hist;
}

constexpr int kTaaTile = 32, kTaaHalo = kTaaTile + 2, kTaaTexels = kTaaHalo * kTaaHalo;   // 34 x 34 = 1156
constexpr int kTaaLoads = (kTaaTexels + 255) / 256;                                      // 5 per thread

// HIST_SMEM: |jitter delta| <= 0.99 px, so every history tap of a tile lies inside its 1-texel halo.
template <bool THRESH, bool HIST_SMEM>
__global__ void __launch_bounds__(256, 3)
k_taa_fast(Tex16 cur, Tex16 hist, uint2* __restrict__ out, uint2* __restrict__ thr,
           AshTaaPushConstants pc, float threshold) {
    __shared__ float4 s_nb[kTaaHalo][kTaaHalo];      // tonemapped YCoCg of `current`, tile + halo
    __shared__ uint2  s_c[kTaaTile][kTaaTile];       // raw `current`, tile
    __shared__ uint2  s_h[HIST_SMEM ? kTaaHalo : 1][HIST_SMEM ? kTaaHalo : 1];   // raw history, tile + halo
    const int W = cur.W, H = cur.H;
    const int x0 = blockIdx.x * kTaaTile, y0 = blockIdx.y * kTaaTile;
    const bool clamp_on = pc.params[1] > 0.5f;
    {   // stage: all global loads of the CTA are issued here, back to back
        uint2 tc[kTaaLoads], th[kTaaLoads];
#pragma unroll
        for (int it = 0; it < kTaaLoads; ++it) {
            const int idx = threadIdx.x + it * 256;
            if (idx < kTaaTexels) {
                const int r = idx / kTaaHalo, c = idx - r * kTaaHalo;
                const size_t g = (size_t)min(max(y0 - 1 + r, 0), H - 1) * W + min(max(x0 - 1 + c, 0), W - 1);   // CLAMP_TO_EDGE
                tc[it] = __ldg(cur.p + g);
                if (HIST_SMEM) th[it] = __ldg(hist.p + g);
            }
        }
#pragma unroll
        for (int it = 0; it < kTaaLoads; ++it) {
            const int idx = threadIdx.x + it * 256;
            if (idx < kTaaTexels) {
                const int r = idx / kTaaHalo, c = idx - r * kTaaHalo;
                if (HIST_SMEM) s_h[r][c] = th[it];
                if (r >= 1 && r <= kTaaTile && c >= 1 && c <= kTaaTile) s_c[r - 1][c - 1] = tc[it];
                if (clamp_on) {
                    const V3 s = rgb_to_ycocg(taa_tonemap_safe(unpack_rgb(tc[it])));                  // :120
                    s_nb[r][c] = make_float4(s.x, s.y, s.z, 0.0f);
                }
            }
        }
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int x = x0 + tx, yb = y0 + 4 * ty;
    if (x >= W || yb >= H) return;

    const float sw = pc.screen_params[0], sh = pc.screen_params[1];
    const float u = fdiv_safe(fadd((float)x, 0.5f), sw);                                   // :92 (extent in [1, 32768])
    const float vx = fdiv_safe(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);                 // :97-99
    const float vy = fdiv_safe(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
    const float hu = gclamp(fsub(u, vx), 0.0f, 1.0f);                                      // :102
    int hx0, hx1; float fx;
    bilinear_setup(hu, W, hx0, hx1, fx);
    float blend = pc.params[0];                                                            // :145-150
    if (pc.params[2] > 0.5f) {
        const float vl = length2(fmul(vx, sw), fmul(vy, sh));
        blend = fmul(blend, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
    }
    const float ib = fsub(1.0f, blend);
    const float ifx = fsub(1.0f, fx);

    // row-wise min / max of the 3 horizontal neighbours, for the 6 rows this strip touches
    V3 rmn[6], rmx[6];
    if (clamp_on) {
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            const float4 a = s_nb[4 * ty + r][tx], b = s_nb[4 * ty + r][tx + 1], c = s_nb[4 * ty + r][tx + 2];
            rmn[r] = {fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z)};
            rmx[r] = {fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z)};
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int y = yb + k;
        if (y >= H) break;
        const float v = fdiv_safe(fadd((float)y, 0.5f), sh);
        const float hv = gclamp(fsub(v, vy), 0.0f, 1.0f);
        int hy0, hy1; float fy;
        bilinear_setup(hv, H, hy0, hy1, fy);
        const V3 c = unpack_rgb(s_c[4 * ty + k][tx]);                                      // :105 (centre tap == texel)
        V3 h;                                                                              // :108 bilinear history
        {
            V3 a, b, cc, d;
            if (HIST_SMEM) {
                const int cx0 = hx0 - (x0 - 1), cx1 = hx1 - (x0 - 1), ry0 = hy0 - (y0 - 1), ry1 = hy1 - (y0 - 1);
                a = unpack_rgb(s_h[ry0][cx0]); b = unpack_rgb(s_h[ry0][cx1]);
                cc = unpack_rgb(s_h[ry1][cx0]); d = unpack_rgb(s_h[ry1][cx1]);
            } else {
                a = unpack_rgb(__ldg(hist.p + (size_t)hy0 * W + hx0)); b = unpack_rgb(__ldg(hist.p + (size_t)hy0 * W + hx1));
                cc = unpack_rgb(__ldg(hist.p + (size_t)hy1 * W + hx0)); d = unpack_rgb(__ldg(hist.p + (size_t)hy1 * W + hx1));
            }
            const float ify = fsub(1.0f, fy);
            const V3 top = {fadd(fmul(a.x, ifx), fmul(b.x, fx)), fadd(fmul(a.y, ifx), fmul(b.y, fx)), fadd(fmul(a.z, ifx), fmul(b.z, fx))};
            const V3 bot = {fadd(fmul(cc.x, ifx), fmul(d.x, fx)), fadd(fmul(cc.y, ifx), fmul(d.y, fx)), fadd(fmul(cc.z, ifx), fmul(d.z, fx))};
            h = {fadd(fmul(top.x, ify), fmul(bot.x, fy)), fadd(fmul(top.y, ify), fmul(bot.y, fy)), fadd(fmul(top.z, ify), fmul(bot.z, fy))};
        }
        if (clamp_on) {                                                                    // :112-142
            V3 mn = {fminf(fminf(rmn[k].x, rmn[k + 1].x), rmn[k + 2].x), fminf(fminf(rmn[k].y, rmn[k + 1].y), rmn[k + 2].y),
                     fminf(fminf(rmn[k].z, rmn[k + 1].z), rmn[k + 2].z)};
            V3 mx = {fmaxf(fmaxf(rmx[k].x, rmx[k + 1].x), rmx[k + 2].x), fmaxf(fmaxf(rmx[k].y, rmx[k + 1].y), rmx[k + 2].y),
                     fmaxf(fmaxf(rmx[k].z, rmx[k + 1].z), rmx[k + 2].z)};
            const V3 ctr = scale3(add3(mn, mx), 0.5f), ext = scale3(sub3(mx, mn), 0.5f);   // :132-135
            mn = sub3(ctr, scale3(ext, 1.25f));
            mx = add3(ctr, scale3(ext, 1.25f));
            const V3 hy = rgb_to_ycocg(taa_tonemap_safe(h));
            h = taa_untonemap(ycocg_to_rgb(clip_aabb_fast(mn, mx, hy)));                   // :140-141 (divisor may reach 0: guarded div)
        }
        const uint2 o = pack_rgba16f(fadd(fmul(c.x, ib), fmul(h.x, blend)), fadd(fmul(c.y, ib), fmul(h.y, blend)),
                                     fadd(fmul(c.z, ib), fmul(h.z, blend)), 1.0f);         // mix(), :159
        out[(size_t)y * W + x] = o;
        if (THRESH) {                      // bloom threshold of the resolved texel as stored (fp16)
            const V3 b = soft_threshold_safe(unpack_rgb(o), threshold);
            thr[(size_t)y * W + x] = pack_rgba16f(b.x, b.y, b.z, 1.0f);
        }
    }
}

bool taa_in_halo(const AshTaaPushConstants& pc) {
    return fabsf(pc.jitter[0] - pc.prev_jitter[0]) <= 0.99f && fabsf(pc.jitter[1] - pc.prev_jitter[1]) <= 0.99f;
}

cudaError_t launch_taa_fast(const void* cur, const void* hist, const void* vel_or_null, void* out, void* thr_or_null,
                            const AshTaaPushConstants& pc, float threshold, const Launch& L) {
    const int W = (int)pc.screen_params[0], H = (int)pc.screen_params[1];
    const Tex16 c{static_cast<const uint2*>(cur), W, H}, h{static_cast<const uint2*>(hist), W, H};
    const dim3 grid((W + 31) / 32, (H + 31) / 32);
    const bool in_halo = taa_in_halo(pc);
    uint2* o = static_cast<uint2*>(out);
    uint2* t = static_cast<uint2*>(thr_or_null);
    // the usual case (sub-pixel jitter delta): two pixels per thread on the packed fp32 pipe, post_pair.cu
    // (with a velocity buffer the history taps are gathered per pixel, so the halo condition does not apply)
    if (vel_or_null || in_halo) return launch_taa_pair(cur, hist, vel_or_null, out, thr_or_null, pc, threshold, L);
    Timed timed(L, thr_or_null ? ASH_KERNEL_TAA_THRESHOLD : ASH_KERNEL_TAA);
    if (t) k_taa_fast<true, false><<<grid, 256, 0, L.s>>>(c, h, o, t, pc, threshold);
    else   k_taa_fast<false, false><<<grid, 256, 0, L.s>>>(c, h, o, t, pc, threshold);
    return cudaGetLastError();
}

// =====================================================================================================
// Bloom downsample, source extent == 2 x destination extent (bloom_downsample.frag:16-108)
// One thread = one destination pixel = one 4x4 source window, fetched with 12 independent loads
// (8 B, 16 B, 8 B per row) issued back to back; neighbouring threads share texels through L1.
// =====================================================================================================
__global__ void __launch_bounds__(256)
k_downsample_fast(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int dw, int dh) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    // 4x4 window: window (i,j) = source texel (2x-1+i, 2y-1+j), CLAMP_TO_EDGE
    const int xl = max(2 * x - 1, 0), xr = min(2 * x + 2, sw - 1);
    uint2 e0[4], e3[4]; uint4 m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint2* row = src + (size_t)min(max(2 * y - 1 + j, 0), sh - 1) * sw;
        e0[j] = __ldg(row + xl);
        m[j] = __ldg(reinterpret_cast<const uint4*>(row + 2 * x));       // texels 2x, 2x+1 (16-byte aligned: sw is even)
        e3[j] = __ldg(row + xr);
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
    dst[(size_t)y * dw + x] = pack_rgba16f(fdiv_safe(res.x, sum), fdiv_safe(res.y, sum), fdiv_safe(res.z, sum), 1.0f);
}

cudaError_t launch_bloom_downsample_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                         uint32_t dh, const Launch& L) {
#ifndef ASH_DOWN_PAIR
#define ASH_DOWN_PAIR 1
#endif
#ifndef ASH_DOWN_BOX
#define ASH_DOWN_BOX 1
#endif
    if (ASH_DOWN_BOX && dw >= 64u && dw % 32u == 0u) return launch_bloom_downsample_box(src, sw, sh, dst, dw, dh, L);           // every box once, post_down_box.cu
    if (ASH_DOWN_PAIR && (dw & 1u) == 0u) return launch_bloom_downsample_pair(src, sw, sh, dst, dw, dh, L);    // two pixels per thread, post_pair.cu
    const dim3 grid((dw + 31) / 32, (dh + 7) / 8);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_DOWNSAMPLE, dw, dh));
    k_downsample_fast<<<grid, 256, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh,
                                             static_cast<uint2*>(dst), (int)dw, (int)dh);
    return cudaGetLastError();
}

// =====================================================================================================
// Bloom upsample + additive blend, destination extent == 2 x source extent
// (bloom_upsample.frag:16-41, PIN-9), optionally fused with tonemapping.frag:29-48 + PIN-11.
// CTA = 16x16 source texels = 32x32 destination pixels; thread = 2x2 destination block.
//
// Destination pixel x = 2i + p samples the source at i - 1/4 + p/2 + m/2 for tap offsets
// m = -2..2 (in units of d = half a source texel): window column (origin i-2) c0 and weight g:
//   c0 = floor((2p - 1 + 2m) / 4) + 2,   g = ((2p - 1 + 2m) mod 4) / 4  in {1/4, 3/4}.
// =====================================================================================================
// (up_fd4 / up_c0 / up_g75: post_math.cuh)

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

#ifndef ASH_UPS_MINB
#define ASH_UPS_MINB 3
#endif
template <bool TONEMAP, bool STORE_U>
__global__ void __launch_bounds__(256, ASH_UPS_MINB)
k_upsample_fast(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, float intensity,
                const uint2* __restrict__ hdr, uint32_t* __restrict__ ldr8, AshPostProcessPushConstants pc) {
    __shared__ uint2 s_t[20][20];
    const int i0 = blockIdx.x * 16, j0 = blockIdx.y * 16;
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
            st[it] = __ldg(src + (size_t)min(max(j0 - 2 + r, 0), sh - 1) * sw + min(max(i0 - 2 + c, 0), sw - 1));   // CLAMP_TO_EDGE
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

// (The persistent, cp.async double-buffered form of this kernel — k_upsample_pipe, 90 -> 68 us for the fused last level
//  at 4K — was superseded by k_upsample_pair, post_up_pair.cu: the same pipeline with two source texels per thread on
//  the packed fp32 pipe, 60 us.  Levels too small to fill the persistent grid twice run the one-shot kernel above.)

#ifndef ASH_UPS_PAIR          // two source texels per thread on the packed fp32 pipe (post_up_pair.cu)
#define ASH_UPS_PAIR 1
#endif
cudaError_t launch_bloom_upsample_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity,
                                       const Launch& L) {
    const dim3 grid((sw + 15) / 16, (sh + 15) / 16);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_UPSAMPLE, 2u * sw, 2u * sh));
    if (ASH_UPS_PAIR && upsample_pair_applies(sw, sh)) return launch_upsample_pair(src, sw, sh, dst, intensity, L.fast, L.s);
    k_upsample_fast<false, true><<<grid, 256, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh,
                                                        static_cast<uint2*>(dst), intensity, nullptr, nullptr,
                                                        AshPostProcessPushConstants{});
    return cudaGetLastError();
}

cudaError_t launch_upsample_tonemap_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity,
                                         const void* hdr, void* ldr8, const AshPostProcessPushConstants& pc,
                                         bool store_u, const Launch& L) {
    const dim3 grid((sw + 15) / 16, (sh + 15) / 16);
    Timed timed(L, ASH_KERNEL_UPSAMPLE_TONEMAP);
    if (ASH_UPS_PAIR && upsample_pair_applies(sw, sh))
        return launch_upsample_tonemap_pair(src, sw, sh, dst, intensity, hdr, ldr8, pc, store_u, L.fast, L.s);
    if (store_u)
        k_upsample_fast<true, true><<<grid, 256, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh, static_cast<uint2*>(dst),
                                                           intensity, static_cast<const uint2*>(hdr), static_cast<uint32_t*>(ldr8), pc);
    else
        k_upsample_fast<true, false><<<grid, 256, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh, static_cast<uint2*>(dst),
                                                            intensity, static_cast<const uint2*>(hdr), static_cast<uint32_t*>(ldr8), pc);
    return cudaGetLastError();
}

// =====================================================================================================
// Tonemap alone (tonemapping.frag:29-48), centre taps validated: one thread = 2 pixels (16 B in, 8 B out)
// =====================================================================================================
__global__ void __launch_bounds__(256)
k_tonemap_fast(const uint4* __restrict__ hdr, const uint4* __restrict__ bloom, uint2* __restrict__ ldr8, size_t pairs,
               AshPostProcessPushConstants pc) {
    const float inv_gamma = __fdividef(1.0f, pc.gamma);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < pairs; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 c = __ldg(hdr + i);
        uint4 b = make_uint4(0, 0, 0, 0);
        if (bloom) b = __ldg(bloom + i);
        ldr8[i] = make_uint2(tonemap_pack8(unpack_rgb(make_uint2(c.x, c.y)), unpack_rgb(make_uint2(b.x, b.y)), pc, inv_gamma),
                             tonemap_pack8(unpack_rgb(make_uint2(c.z, c.w)), unpack_rgb(make_uint2(b.z, b.w)), pc, inv_gamma));
    }
}

cudaError_t launch_tonemap_fast(const void* hdr, const void* bloom_or_null, void* ldr8, uint32_t W, uint32_t H,
                                const AshPostProcessPushConstants& pc, const Launch& L) {
    const size_t pairs = (size_t)W * H / 2;
    const unsigned grid = (unsigned)min((pairs + 255) / 256, (size_t)148 * 16);
    Timed timed(L, ASH_KERNEL_TONEMAP);
    k_tonemap_fast<<<grid, 256, 0, L.s>>>(static_cast<const uint4*>(hdr), static_cast<const uint4*>(bloom_or_null),
                                          static_cast<uint2*>(ldr8), pairs, pc);
    return cudaGetLastError();
}

}  // namespace ash
