// post.cu — TAA resolve, bloom threshold / downsample / upsample, tonemap: the GENERAL kernels, one thread per pixel and one
// kernel per reference shader dispatch, valid for any extent and any sampler position.  They are the fallback: at the
// extents the benchmark runs, ashcuda.cu dispatches to post_fast.cu / post_pair.cu / post_taa_tol.cu / post_up_pair.cu /
// post_down_box.cu whenever exact_taps.cu validates their sampler shortcuts, and says so once on stderr when a level of
// 2^20 texels or more falls back here.
//
// All colour targets are RGBA16F (8 B / texel, renderer/hdr_framebuffer.rs:34, taa_resolve.comp:19),
// sampled LINEAR + CLAMP_TO_EDGE with 8-bit sub-texel weights (PIN-6).  A tap that lands on a
// texel centre of a same-sized image (error << 1/512 texel for extents <= 4096) therefore equals
// that texel exactly; CENTER=true kernels fetch it directly, CENTER=false kernels run the general
// sampler for every tap (any extent).  Arithmetic is the pinned-order IEEE sequence of ash_math.cuh,
// so outputs are bit-identical to the reference semantics, not merely within tolerance.
#include "post_math.cuh"

namespace ash {

template <bool CENTER, bool HAS_VEL>
__global__ void __launch_bounds__(256)
k_taa_resolve(Tex16 cur, Tex16 hist, const __half2* __restrict__ vel, uint2* __restrict__ out,
              AshTaaPushConstants pc) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    const int W = cur.W, H = cur.H;
    if (x >= W || y >= H) return;
    const float sw = pc.screen_params[0], sh = pc.screen_params[1];
    const float u = fdiv(fadd((float)x, 0.5f), sw), v = fdiv(fadd((float)y, 0.5f), sh);     // :92
    float vx = 0.0f, vy = 0.0f;
    if (HAS_VEL) {
        if (CENTER) {
            const float2 q = __half22float2(__ldg(vel + (size_t)y * W + x));
            vx = q.x; vy = q.y;
        } else {
            int x0, x1, y0, y1; float fx, fy;
            bilinear_setup(u, W, x0, x1, fx); bilinear_setup(v, H, y0, y1, fy);
            const float2 a = __half22float2(__ldg(vel + (size_t)y0 * W + x0)), b = __half22float2(__ldg(vel + (size_t)y0 * W + x1));
            const float2 c = __half22float2(__ldg(vel + (size_t)y1 * W + x0)), d = __half22float2(__ldg(vel + (size_t)y1 * W + x1));
            vx = lerp1(lerp1(a.x, b.x, fx), lerp1(c.x, d.x, fx), fy);
            vy = lerp1(lerp1(a.y, b.y, fx), lerp1(c.y, d.y, fx), fy);
        }
    }
    if (length2(vx, vy) < 0.0001f) {                                                       // :97-99
        vx = fdiv(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);
        vy = fdiv(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
    }
    const float hu = gclamp(fsub(u, vx), 0.0f, 1.0f), hv = gclamp(fsub(v, vy), 0.0f, 1.0f); // :102
    const V3 c = CENTER ? rgb(fetch16(cur, x, y)) : rgb(texture16(cur, u, v));             // :105
    V3 h = rgb(texture16(hist, hu, hv));                                                   // :108
    if (pc.params[1] > 0.5f) {                                                             // :112-142
        V3 mn, mx;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                V3 s;
                if (CENTER) s = rgb(fetch16_clamped(cur, x + dx, y + dy));
                else s = rgb(texture16(cur, fadd(u, fmul((float)dx, pc.screen_params[2])),
                                       fadd(v, fmul((float)dy, pc.screen_params[3]))));
                s = rgb_to_ycocg(taa_tonemap(s));
                if (dy == -1 && dx == -1) { mn = s; mx = s; } else { mn = min3(mn, s); mx = max3(mx, s); }
            }
        const V3 ctr = scale3(add3(mn, mx), 0.5f), ext = scale3(sub3(mx, mn), 0.5f);       // :132-135
        mn = sub3(ctr, scale3(ext, 1.25f));
        mx = add3(ctr, scale3(ext, 1.25f));
        const V3 hy = rgb_to_ycocg(taa_tonemap(h));
        h = taa_untonemap(ycocg_to_rgb(clip_aabb(mn, mx, hy)));                            // :140-141
    }
    float blend = pc.params[0];                                                            // :145-150
    if (pc.params[2] > 0.5f) {
        const float vl = length2(fmul(vx, sw), fmul(vy, sh));
        blend = fmul(blend, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
    }
    if (hu < 0.0f || hu > 1.0f || hv < 0.0f || hv > 1.0f) blend = 0.0f;                    // :153-156
    const float ib = fsub(1.0f, blend);                                                    // mix(), :159
    out[(size_t)y * W + x] = pack_rgba16f(fadd(fmul(c.x, ib), fmul(h.x, blend)), fadd(fmul(c.y, ib), fmul(h.y, blend)),
                                          fadd(fmul(c.z, ib), fmul(h.z, blend)), 1.0f);
}

cudaError_t launch_taa(const void* cur, const void* hist, const void* vel, void* out,
                       const AshTaaPushConstants& pc, bool center, const Launch& L) {
    const int W = (int)pc.screen_params[0], H = (int)pc.screen_params[1];
    const Tex16 c{static_cast<const uint2*>(cur), W, H}, h{static_cast<const uint2*>(hist), W, H};
    const dim3 grid((W + 31) / 32, (H + 7) / 8);
    Timed timed(L, ASH_KERNEL_TAA);
    const __half2* v = static_cast<const __half2*>(vel);
    uint2* o = static_cast<uint2*>(out);
    if (center) { if (v) k_taa_resolve<true, true><<<grid, 256, 0, L.s>>>(c, h, v, o, pc); else k_taa_resolve<true, false><<<grid, 256, 0, L.s>>>(c, h, v, o, pc); }
    else        { if (v) k_taa_resolve<false, true><<<grid, 256, 0, L.s>>>(c, h, v, o, pc); else k_taa_resolve<false, false><<<grid, 256, 0, L.s>>>(c, h, v, o, pc); }
    return cudaGetLastError();
}

// ---------------------------------------------------------------- bloom threshold (bloom_threshold.frag:18-35)
template <bool CENTER>
__global__ void __launch_bounds__(256)
k_bloom_threshold(Tex16 hdr, uint2* __restrict__ out, float threshold) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= hdr.W || y >= hdr.H) return;
    V3 c;
    if (CENTER) c = rgb(fetch16(hdr, x, y));
    else c = rgb(texture16(hdr, fdiv(fadd((float)x, 0.5f), (float)hdr.W), fdiv(fadd((float)y, 0.5f), (float)hdr.H)));
    const V3 b = soft_threshold(c, threshold);
    out[(size_t)y * hdr.W + x] = pack_rgba16f(b.x, b.y, b.z, 1.0f);
}

cudaError_t launch_bloom_threshold(const void* hdr, void* out, uint32_t W, uint32_t H, float threshold,
                                   bool center, const Launch& L) {
    const Tex16 t{static_cast<const uint2*>(hdr), (int)W, (int)H};
    const dim3 grid((W + 31) / 32, (H + 7) / 8);
    Timed timed(L, ASH_KERNEL_BLOOM_THRESHOLD);
    if (center) k_bloom_threshold<true><<<grid, 256, 0, L.s>>>(t, static_cast<uint2*>(out), threshold);
    else k_bloom_threshold<false><<<grid, 256, 0, L.s>>>(t, static_cast<uint2*>(out), threshold);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- bloom downsample (bloom_downsample.frag:16-108)
__global__ void __launch_bounds__(256)
k_bloom_downsample(Tex16 src, uint2* __restrict__ dst, int dw, int dh) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    const float dx = fdiv(1.0f, (float)src.W), dy = fdiv(1.0f, (float)src.H);              // texelSize, :11
    const float u = fdiv(fadd((float)x, 0.5f), (float)dw), v = fdiv(fadd((float)y, 0.5f), (float)dh);
    // tap order centre,a,b,c,d,e,f,g,h (:22-34), weights (:87-101)
    const float ou[9] = {0.0f, -dx, dx, -dx, dx, -dx, dx, 0.0f, 0.0f};
    const float ov[9] = {0.0f, -dy, -dy, dy, dy, 0.0f, 0.0f, -dy, dy};
    const float wt[9] = {0.25f, 0.0625f, 0.0625f, 0.0625f, 0.0625f, 0.125f, 0.125f, 0.125f, 0.125f};
    V3 res = {0.0f, 0.0f, 0.0f};
    float sum = 0.0f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const V3 s = rgb(texture16(src, fadd(u, ou[k]), fadd(v, ov[k])));
        const float w = fdiv(1.0f, fadd(1.0f, dot3(s, v3(0.2126f, 0.7152f, 0.0722f))));
        res = add3(res, scale3(scale3(s, w), wt[k]));
        sum = fadd(sum, fmul(w, wt[k]));
    }
    dst[(size_t)y * dw + x] = pack_rgba16f(fdiv(res.x, sum), fdiv(res.y, sum), fdiv(res.z, sum), 1.0f);   // :107
}

cudaError_t launch_bloom_downsample(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                    uint32_t dh, const Launch& L) {
    const Tex16 t{static_cast<const uint2*>(src), (int)sw, (int)sh};
    const dim3 grid((dw + 31) / 32, (dh + 7) / 8);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_DOWNSAMPLE, dw, dh));
    k_bloom_downsample<<<grid, 256, 0, L.s>>>(t, static_cast<uint2*>(dst), (int)dw, (int)dh);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- bloom upsample + additive blend (bloom_upsample.frag:16-41, PIN-9)
__global__ void __launch_bounds__(256)
k_bloom_upsample_add(Tex16 src, uint2* __restrict__ dst, int dw, int dh, float intensity) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    const float dx = fmul(fdiv(1.0f, (float)src.W), 0.5f), dy = fmul(fdiv(1.0f, (float)src.H), 0.5f);   // :18
    const float u = fdiv(fadd((float)x, 0.5f), (float)dw), v = fdiv(fadd((float)y, 0.5f), (float)dh);
    const V3 s0 = rgb(texture16(src, fadd(u, fmul(-dx, 2.0f)), fadd(v, 0.0f)));            // :21-29
    const V3 s1 = rgb(texture16(src, fadd(u, -dx), fadd(v, -dy)));
    const V3 s2 = rgb(texture16(src, fadd(u, 0.0f), fadd(v, fmul(-dy, 2.0f))));
    const V3 s3 = rgb(texture16(src, fadd(u, dx), fadd(v, -dy)));
    const V3 s4 = rgb(texture16(src, fadd(u, fmul(dx, 2.0f)), fadd(v, 0.0f)));
    const V3 s5 = rgb(texture16(src, fadd(u, dx), fadd(v, dy)));
    const V3 s6 = rgb(texture16(src, fadd(u, 0.0f), fadd(v, fmul(dy, 2.0f))));
    const V3 s7 = rgb(texture16(src, fadd(u, -dx), fadd(v, dy)));
    const V3 s8 = rgb(texture16(src, u, v));
    V3 r = scale3(s8, 4.0f);                                                               // :32-38
    r = add3(r, scale3(add3(add3(add3(s1, s3), s5), s7), 2.0f));
    r = add3(r, add3(add3(add3(s0, s2), s4), s6));
    r = divs3(r, 16.0f);
    r = scale3(r, intensity);
    const V4 d = unpack_rgba16f(dst[(size_t)y * dw + x]);
    dst[(size_t)y * dw + x] = pack_rgba16f(fadd(r.x, d.x), fadd(r.y, d.y), fadd(r.z, d.z), 1.0f);
}

cudaError_t launch_bloom_upsample_add(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                      uint32_t dh, float intensity, const Launch& L) {
    const Tex16 t{static_cast<const uint2*>(src), (int)sw, (int)sh};
    const dim3 grid((dw + 31) / 32, (dh + 7) / 8);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_UPSAMPLE, dw, dh));
    k_bloom_upsample_add<<<grid, 256, 0, L.s>>>(t, static_cast<uint2*>(dst), (int)dw, (int)dh, intensity);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- tonemap (tonemapping.frag:20-49, PIN-11)
template <bool CENTER>
__global__ void __launch_bounds__(256)
k_tonemap(Tex16 hdr, Tex16 bloom, uint32_t* __restrict__ ldr8, AshPostProcessPushConstants pc, uint32_t op) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= hdr.W || y >= hdr.H) return;
    V3 c, b = {0.0f, 0.0f, 0.0f};
    if (CENTER) {
        c = rgb(fetch16(hdr, x, y));
        if (bloom.p) b = rgb(fetch16(bloom, x, y));
    } else {
        const float u = fdiv(fadd((float)x, 0.5f), (float)hdr.W), v = fdiv(fadd((float)y, 0.5f), (float)hdr.H);
        c = rgb(texture16(hdr, u, v));
        if (bloom.p) b = rgb(texture16(bloom, u, v));
    }
    c = add3(c, scale3(b, pc.bloom_intensity));                                            // :38
    c = scale3(c, pc.exposure);                                                            // :41
    const float ig = fdiv(1.0f, pc.gamma);
    const uint32_t r8 = to_unorm8(powf(tonemap_curve(c.x, op), ig));                       // :44-47 (aces, or PIN-13)
    const uint32_t g8 = to_unorm8(powf(tonemap_curve(c.y, op), ig));
    const uint32_t b8 = to_unorm8(powf(tonemap_curve(c.z, op), ig));
    ldr8[(size_t)y * hdr.W + x] = r8 | (g8 << 8) | (b8 << 16) | 0xFF000000u;
}

cudaError_t launch_tonemap(const void* hdr, const void* bloom_or_null, void* ldr8, uint32_t W, uint32_t H,
                           const AshPostProcessPushConstants& pc, bool center, uint32_t op, const Launch& L) {
    const Tex16 h{static_cast<const uint2*>(hdr), (int)W, (int)H}, b{static_cast<const uint2*>(bloom_or_null), (int)W, (int)H};
    const dim3 grid((W + 31) / 32, (H + 7) / 8);
    Timed timed(L, ASH_KERNEL_TONEMAP);
    if (center) k_tonemap<true><<<grid, 256, 0, L.s>>>(h, b, static_cast<uint32_t*>(ldr8), pc, op);
    else k_tonemap<false><<<grid, 256, 0, L.s>>>(h, b, static_cast<uint32_t*>(ldr8), pc, op);
    return cudaGetLastError();
}

}  // namespace ash
