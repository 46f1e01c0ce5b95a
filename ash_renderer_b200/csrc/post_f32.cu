// post_f32.cu — the post chain in the fp32_intermediates storage mode of SURVEY.md PIN-6 (AshCudaConfig.fp32_intermediates):
// the TAA output / history and every bloom target are RGBA32F (16 B / texel) instead of RGBA16F, so nothing is rounded
// to fp16 between stages and BASELINE.json's "max-abs 1e-5 in fp32" criterion can be exercised as stated.  The HDR
// colour INPUT keeps the reference's format (RGBA16F, hdr_framebuffer.rs:34): it is an input, not an intermediate.
//
// One kernel per reference shader dispatch, one thread per destination pixel, every tap through the general LINEAR
// sampler (any extent), the pinned IEEE operation sequence of ash_math.cuh — the same arithmetic as post.cu and as the
// oracle's `_fmt` functions, on textures whose texel type is a run-time flag.  This is the parity mode, not the fast one:
// twice the bytes of the RGBA16F chain and none of the fused kernels; the benchmark never runs it.
#include "post_math.cuh"

namespace ash {

struct TexF { const void* p; int W, H, f32; };          // RGBA16F (uint2 texels) or RGBA32F (float4 texels)

__device__ __forceinline__ V4 fetchf(const TexF& t, int x, int y) {
    const size_t i = (size_t)y * t.W + x;
    if (t.f32) { const float4 q = __ldg(static_cast<const float4*>(t.p) + i); return {q.x, q.y, q.z, q.w}; }
    return unpack_rgba16f(__ldg(static_cast<const uint2*>(t.p) + i));
}
__device__ __forceinline__ V4 texturef(const TexF& t, float u, float v) {      // PIN-6: horizontal then vertical, a*(1-f) + b*f
    int x0, x1, y0, y1; float fx, fy;
    bilinear_setup(u, t.W, x0, x1, fx);
    bilinear_setup(v, t.H, y0, y1, fy);
    const V4 a = fetchf(t, x0, y0), b = fetchf(t, x1, y0), c = fetchf(t, x0, y1), d = fetchf(t, x1, y1);
    const V4 top = {lerp1(a.x, b.x, fx), lerp1(a.y, b.y, fx), lerp1(a.z, b.z, fx), lerp1(a.w, b.w, fx)};
    const V4 bot = {lerp1(c.x, d.x, fx), lerp1(c.y, d.y, fx), lerp1(c.z, d.z, fx), lerp1(c.w, d.w, fx)};
    return {lerp1(top.x, bot.x, fy), lerp1(top.y, bot.y, fy), lerp1(top.z, bot.z, fy), lerp1(top.w, bot.w, fy)};
}
__device__ __forceinline__ void storef(void* img, int f32, size_t i, float r, float g, float b, float a) {
    if (f32) static_cast<float4*>(img)[i] = make_float4(r, g, b, a);
    else static_cast<uint2*>(img)[i] = pack_rgba16f(r, g, b, a);
}

// ---------------------------------------------------------------- taa_resolve.comp:83-167
__global__ void __launch_bounds__(256)
k_taa_resolve_f(TexF cur, TexF hist, const __half2* __restrict__ vel, void* __restrict__ out, int out_f32,
                AshTaaPushConstants pc) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    const int W = cur.W, H = cur.H;
    if (x >= W || y >= H) return;
    const float sw = pc.screen_params[0], sh = pc.screen_params[1];
    const float u = fdiv(fadd((float)x, 0.5f), sw), v = fdiv(fadd((float)y, 0.5f), sh);     // :92
    float vx = 0.0f, vy = 0.0f;
    if (vel) {
        int x0, x1, y0, y1; float fx, fy;
        bilinear_setup(u, W, x0, x1, fx); bilinear_setup(v, H, y0, y1, fy);
        const float2 a = __half22float2(__ldg(vel + (size_t)y0 * W + x0)), b = __half22float2(__ldg(vel + (size_t)y0 * W + x1));
        const float2 c = __half22float2(__ldg(vel + (size_t)y1 * W + x0)), d = __half22float2(__ldg(vel + (size_t)y1 * W + x1));
        vx = lerp1(lerp1(a.x, b.x, fx), lerp1(c.x, d.x, fx), fy);
        vy = lerp1(lerp1(a.y, b.y, fx), lerp1(c.y, d.y, fx), fy);
    }
    if (length2(vx, vy) < 0.0001f) {                                                       // :97-99
        vx = fdiv(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);
        vy = fdiv(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
    }
    const float hu = gclamp(fsub(u, vx), 0.0f, 1.0f), hv = gclamp(fsub(v, vy), 0.0f, 1.0f); // :102
    const V3 c = rgb(texturef(cur, u, v));                                                 // :105
    V3 h = rgb(texturef(hist, hu, hv));                                                    // :108
    if (pc.params[1] > 0.5f) {                                                             // :112-142
        V3 mn, mx;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                V3 s = rgb(texturef(cur, fadd(u, fmul((float)dx, pc.screen_params[2])), fadd(v, fmul((float)dy, pc.screen_params[3]))));
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
    storef(out, out_f32, (size_t)y * W + x, fadd(fmul(c.x, ib), fmul(h.x, blend)), fadd(fmul(c.y, ib), fmul(h.y, blend)),
           fadd(fmul(c.z, ib), fmul(h.z, blend)), 1.0f);
}

// ---------------------------------------------------------------- bloom_threshold.frag:18-35
__global__ void __launch_bounds__(256)
k_bloom_threshold_f(TexF hdr, void* __restrict__ out, int out_f32, float threshold) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= hdr.W || y >= hdr.H) return;
    const V3 c = rgb(texturef(hdr, fdiv(fadd((float)x, 0.5f), (float)hdr.W), fdiv(fadd((float)y, 0.5f), (float)hdr.H)));
    const V3 b = soft_threshold(c, threshold);
    storef(out, out_f32, (size_t)y * hdr.W + x, b.x, b.y, b.z, 1.0f);
}

// ---------------------------------------------------------------- bloom_downsample.frag:16-108
__global__ void __launch_bounds__(256)
k_bloom_downsample_f(TexF src, void* __restrict__ dst, int dw, int dh) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    const float dx = fdiv(1.0f, (float)src.W), dy = fdiv(1.0f, (float)src.H);              // texelSize, :11
    const float u = fdiv(fadd((float)x, 0.5f), (float)dw), v = fdiv(fadd((float)y, 0.5f), (float)dh);
    const float ou[9] = {0.0f, -dx, dx, -dx, dx, -dx, dx, 0.0f, 0.0f};                     // centre,a,b,c,d,e,f,g,h (:22-34)
    const float ov[9] = {0.0f, -dy, -dy, dy, dy, 0.0f, 0.0f, -dy, dy};
    const float wt[9] = {0.25f, 0.0625f, 0.0625f, 0.0625f, 0.0625f, 0.125f, 0.125f, 0.125f, 0.125f};
    V3 res = {0.0f, 0.0f, 0.0f};
    float sum = 0.0f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const V3 s = rgb(texturef(src, fadd(u, ou[k]), fadd(v, ov[k])));
        const float w = fdiv(1.0f, fadd(1.0f, dot3(s, v3(0.2126f, 0.7152f, 0.0722f))));    // :87-101
        res = add3(res, scale3(scale3(s, w), wt[k]));
        sum = fadd(sum, fmul(w, wt[k]));
    }
    storef(dst, src.f32, (size_t)y * dw + x, fdiv(res.x, sum), fdiv(res.y, sum), fdiv(res.z, sum), 1.0f);   // :107
}

// ---------------------------------------------------------------- bloom_upsample.frag:16-41 + additive blend (PIN-9)
__global__ void __launch_bounds__(256)
k_bloom_upsample_add_f(TexF src, void* __restrict__ dst, int dw, int dh, float intensity) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    const float dx = fmul(fdiv(1.0f, (float)src.W), 0.5f), dy = fmul(fdiv(1.0f, (float)src.H), 0.5f);   // :18
    const float u = fdiv(fadd((float)x, 0.5f), (float)dw), v = fdiv(fadd((float)y, 0.5f), (float)dh);
    const V3 s0 = rgb(texturef(src, fadd(u, fmul(-dx, 2.0f)), fadd(v, 0.0f)));             // :21-29
    const V3 s1 = rgb(texturef(src, fadd(u, -dx), fadd(v, -dy)));
    const V3 s2 = rgb(texturef(src, fadd(u, 0.0f), fadd(v, fmul(-dy, 2.0f))));
    const V3 s3 = rgb(texturef(src, fadd(u, dx), fadd(v, -dy)));
    const V3 s4 = rgb(texturef(src, fadd(u, fmul(dx, 2.0f)), fadd(v, 0.0f)));
    const V3 s5 = rgb(texturef(src, fadd(u, dx), fadd(v, dy)));
    const V3 s6 = rgb(texturef(src, fadd(u, 0.0f), fadd(v, fmul(dy, 2.0f))));
    const V3 s7 = rgb(texturef(src, fadd(u, -dx), fadd(v, dy)));
    const V3 s8 = rgb(texturef(src, u, v));
    V3 r = scale3(s8, 4.0f);                                                               // :32-38
    r = add3(r, scale3(add3(add3(add3(s1, s3), s5), s7), 2.0f));
    r = add3(r, add3(add3(add3(s0, s2), s4), s6));
    r = divs3(r, 16.0f);
    r = scale3(r, intensity);
    const TexF d{dst, dw, dh, src.f32};
    const V4 o = fetchf(d, x, y);
    storef(dst, src.f32, (size_t)y * dw + x, fadd(r.x, o.x), fadd(r.y, o.y), fadd(r.z, o.z), 1.0f);
}

// ---------------------------------------------------------------- tonemapping.frag:20-49 (+ PIN-13 operators), PIN-11
__global__ void __launch_bounds__(256)
k_tonemap_f(TexF hdr, TexF bloom, uint32_t* __restrict__ ldr8, AshPostProcessPushConstants pc, uint32_t op) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= hdr.W || y >= hdr.H) return;
    const float u = fdiv(fadd((float)x, 0.5f), (float)hdr.W), v = fdiv(fadd((float)y, 0.5f), (float)hdr.H);
    V3 c = rgb(texturef(hdr, u, v)), b = {0.0f, 0.0f, 0.0f};
    if (bloom.p) b = rgb(texturef(bloom, u, v));
    c = add3(c, scale3(b, pc.bloom_intensity));                                            // :38
    c = scale3(c, pc.exposure);                                                            // :41
    const float ig = fdiv(1.0f, pc.gamma);
    const uint32_t r8 = to_unorm8(powf(tonemap_curve(c.x, op), ig));                       // :44-47
    const uint32_t g8 = to_unorm8(powf(tonemap_curve(c.y, op), ig));
    const uint32_t b8 = to_unorm8(powf(tonemap_curve(c.z, op), ig));
    ldr8[(size_t)y * hdr.W + x] = r8 | (g8 << 8) | (b8 << 16) | 0xFF000000u;
}

// ---------------------------------------------------------------- launchers (fmt flags: 1 = that image is RGBA32F)
static dim3 grid_of(uint32_t w, uint32_t h) { return dim3((w + 31) / 32, (h + 7) / 8); }

cudaError_t launch_taa_f(const void* cur, int cur_f32, const void* hist, int hist_f32, const void* vel, void* out, int out_f32,
                         const AshTaaPushConstants& pc, const Launch& L) {
    const int W = (int)pc.screen_params[0], H = (int)pc.screen_params[1];
    Timed timed(L, ASH_KERNEL_TAA);
    k_taa_resolve_f<<<grid_of(W, H), 256, 0, L.s>>>(TexF{cur, W, H, cur_f32}, TexF{hist, W, H, hist_f32},
                                                   static_cast<const __half2*>(vel), out, out_f32, pc);
    return cudaGetLastError();
}
cudaError_t launch_bloom_threshold_f(const void* hdr, int hdr_f32, void* out, int out_f32, uint32_t W, uint32_t H, float threshold,
                                     const Launch& L) {
    Timed timed(L, ASH_KERNEL_BLOOM_THRESHOLD);
    k_bloom_threshold_f<<<grid_of(W, H), 256, 0, L.s>>>(TexF{hdr, (int)W, (int)H, hdr_f32}, out, out_f32, threshold);
    return cudaGetLastError();
}
cudaError_t launch_bloom_downsample_f(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw, uint32_t dh, int f32, const Launch& L) {
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_DOWNSAMPLE, dw, dh));
    k_bloom_downsample_f<<<grid_of(dw, dh), 256, 0, L.s>>>(TexF{src, (int)sw, (int)sh, f32}, dst, (int)dw, (int)dh);
    return cudaGetLastError();
}
cudaError_t launch_bloom_upsample_add_f(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw, uint32_t dh, float intensity,
                                        int f32, const Launch& L) {
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_UPSAMPLE, dw, dh));
    k_bloom_upsample_add_f<<<grid_of(dw, dh), 256, 0, L.s>>>(TexF{src, (int)sw, (int)sh, f32}, dst, (int)dw, (int)dh, intensity);
    return cudaGetLastError();
}
cudaError_t launch_tonemap_f(const void* hdr, int hdr_f32, const void* bloom_or_null, int bloom_f32, void* ldr8, uint32_t W, uint32_t H,
                             const AshPostProcessPushConstants& pc, uint32_t op, const Launch& L) {
    Timed timed(L, ASH_KERNEL_TONEMAP);
    k_tonemap_f<<<grid_of(W, H), 256, 0, L.s>>>(TexF{hdr, (int)W, (int)H, hdr_f32}, TexF{bloom_or_null, (int)W, (int)H, bloom_f32},
                                               static_cast<uint32_t*>(ldr8), pc, op);
    return cudaGetLastError();
}

}  // namespace ash
