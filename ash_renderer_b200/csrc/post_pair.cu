// post_pair.cu — post kernels that evaluate two pixels per thread in lock-step on the packed fp32 pipe
// (FADD2 / FMUL2 / FFMA2, ash_f32x2.cuh).  Bit-identical to post_fast.cu / post.cu: the same pinned
// IEEE operation sequence per lane; only the instruction count changes.
#include "post_pair.cuh"

namespace ash {

__device__ __forceinline__ V3 unpack_rgb_(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

// =====================================================================================================
// TAA resolve (+ bloom threshold) — taa_resolve.comp:83-167 (+ bloom_threshold.frag:18-35), no velocity
// buffer (PIN-12), |jitter delta| <= 0.99 px so that every history tap lies in the tile's 1-texel halo.
// CTA = 32x32 pixels, 256 threads; thread = 1x4 strip = two (row k, row k+1) pairs.
// The history sampling coordinates are separable: 32 column entries and 32 row entries per tile are
// computed once (by 64 threads, while the global loads are in flight) instead of once per pixel.
// =====================================================================================================
constexpr int kT = 32, kHalo = kT + 2, kTexels = kHalo * kHalo;      // 34 x 34 = 1156
constexpr int kLoads = (kTexels + 255) / 256;                        // 5 per thread

struct AxisTap { int i0, i1; float f; float pad; };                  // halo-local texel indices + 8-bit weight

template <bool THRESH>
__global__ void __launch_bounds__(256, 3)
k_taa_pair(Tex16 cur, Tex16 hist, uint2* __restrict__ out, uint2* __restrict__ thr,
           AshTaaPushConstants pc, float threshold) {
    __shared__ float4 s_nb[kHalo][kHalo];      // tonemapped YCoCg of `current`, tile + halo
    __shared__ uint2  s_c[kT][kT];             // raw `current`, tile
    __shared__ uint2  s_h[kHalo][kHalo];       // raw history, tile + halo
    __shared__ AxisTap s_col[kT], s_row[kT];
    __shared__ float  s_blend;
    const int W = cur.W, H = cur.H;
    const int x0 = blockIdx.x * kT, y0 = blockIdx.y * kT;
    const bool clamp_on = pc.params[1] > 0.5f;
    {   // stage: all global loads of the CTA are issued here, back to back
        uint2 tc[kLoads], th[kLoads];
#pragma unroll
        for (int it = 0; it < kLoads; ++it) {
            const int idx = threadIdx.x + it * 256;
            if (idx < kTexels) {
                const int r = idx / kHalo, c = idx - r * kHalo;
                const size_t g = (size_t)min(max(y0 - 1 + r, 0), H - 1) * W + min(max(x0 - 1 + c, 0), W - 1);   // CLAMP_TO_EDGE
                tc[it] = __ldg(cur.p + g);
                th[it] = __ldg(hist.p + g);
            }
        }
        if (threadIdx.x < 2 * kT) {     // coordinate tables, :92-102 + PIN-6, while the loads fly
            const bool is_row = threadIdx.x >= kT;
            const int k = threadIdx.x & (kT - 1);
            const float ext = is_row ? pc.screen_params[1] : pc.screen_params[0];
            const int n = is_row ? H : W, p0 = is_row ? y0 : x0;
            const float jd = is_row ? fsub(pc.jitter[1], pc.prev_jitter[1]) : fsub(pc.jitter[0], pc.prev_jitter[0]);
            const float u = fdiv_safe(fadd((float)(p0 + k), 0.5f), ext);            // :92 (extent in [1, 32768])
            const float vel = fdiv_safe(jd, ext);                                    // :97-99
            const float hu = gclamp(fsub(u, vel), 0.0f, 1.0f);                       // :102
            AxisTap t;
            bilinear_setup(hu, n, t.i0, t.i1, t.f);
            t.i0 = min(max(t.i0 - (p0 - 1), 0), kHalo - 1);                          // halo-local (clamp only guards rows/cols past the extent)
            t.i1 = min(max(t.i1 - (p0 - 1), 0), kHalo - 1);
            t.pad = 0.0f;
            if (is_row) s_row[k] = t; else s_col[k] = t;
            if (threadIdx.x == 0) {                                                  // :145-150
                float blend = pc.params[0];
                if (pc.params[2] > 0.5f) {
                    const float sw = pc.screen_params[0], sh = pc.screen_params[1];
                    const float vx = fdiv_safe(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);
                    const float vy = fdiv_safe(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
                    const float vl = length2(fmul(vx, sw), fmul(vy, sh));
                    blend = fmul(blend, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
                }
                s_blend = blend;
            }
        }
        // tonemapped YCoCg of the staged texels, two at a time
        auto store_raw = [&](int it) {
            const int idx = threadIdx.x + it * 256;
            const int r = idx / kHalo, c = idx - r * kHalo;
            s_h[r][c] = th[it];
            if (r >= 1 && r <= kT && c >= 1 && c <= kT) s_c[r - 1][c - 1] = tc[it];
        };
#pragma unroll
        for (int it = 0; it + 1 < kLoads; it += 2) {        // both indices < 1024 + 255 < kTexels?  (it+1)*256+255 = 1023 for it = 2: always in range
            store_raw(it); store_raw(it + 1);
            if (clamp_on) {
                const V3P s = rgb_to_ycocg(taa_tonemap_safe(pk3(unpack_rgb_(tc[it]), unpack_rgb_(tc[it + 1]))));   // :120
                const int i0 = threadIdx.x + it * 256, i1 = i0 + 256;
                s_nb[i0 / kHalo][i0 % kHalo] = make_float4(lo(s.x), lo(s.y), lo(s.z), 0.0f);
                s_nb[i1 / kHalo][i1 % kHalo] = make_float4(hi(s.x), hi(s.y), hi(s.z), 0.0f);
            }
        }
        if (kLoads & 1) {
            const int it = kLoads - 1, idx = threadIdx.x + it * 256;
            if (idx < kTexels) {
                store_raw(it);
                if (clamp_on) {
                    const V3 c3 = unpack_rgb_(tc[it]);
                    const V3 tm = {fdiv_safe(c3.x, fadd(1.0f, c3.x)), fdiv_safe(c3.y, fadd(1.0f, c3.y)), fdiv_safe(c3.z, fadd(1.0f, c3.z))};
                    const V3 s = rgb_to_ycocg(tm);
                    s_nb[idx / kHalo][idx % kHalo] = make_float4(s.x, s.y, s.z, 0.0f);
                }
            }
        }
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int x = x0 + tx, yb = y0 + 4 * ty;
    if (x >= W || yb >= H) return;

    const AxisTap col = s_col[tx];
    const float blend = s_blend, ib = fsub(1.0f, blend);
    const float fx = col.f, ifx = fsub(1.0f, fx);

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
    for (int k = 0; k < 4; k += 2) {                 // lanes: lo = row yb + k, hi = row yb + k + 1
        const int y = yb + k;
        if (y >= H) break;
        const AxisTap ra = s_row[4 * ty + k], rb = s_row[4 * ty + k + 1];
        const V3P c = pk3(unpack_rgb_(s_c[4 * ty + k][tx]), unpack_rgb_(s_c[4 * ty + k + 1][tx]));     // :105 (centre tap == texel)
        V3P h;                                                                                         // :108 bilinear history
        {
            const V3P a  = pk3(unpack_rgb_(s_h[ra.i0][col.i0]), unpack_rgb_(s_h[rb.i0][col.i0]));
            const V3P b  = pk3(unpack_rgb_(s_h[ra.i0][col.i1]), unpack_rgb_(s_h[rb.i0][col.i1]));
            const V3P cc = pk3(unpack_rgb_(s_h[ra.i1][col.i0]), unpack_rgb_(s_h[rb.i1][col.i0]));
            const V3P d  = pk3(unpack_rgb_(s_h[ra.i1][col.i1]), unpack_rgb_(s_h[rb.i1][col.i1]));
            const P2 fy = pk(ra.f, rb.f), ify = fsub(1.0f, fy);
            const V3P top = add3(scale3(a, ifx), scale3(b, fx));
            const V3P bot = add3(scale3(cc, ifx), scale3(d, fx));
            h = add3(scale3(top, ify), scale3(bot, fy));
        }
        if (clamp_on) {                                                                                // :112-142
            auto mn3 = [&](int r) { return V3{fminf(fminf(rmn[r].x, rmn[r + 1].x), rmn[r + 2].x), fminf(fminf(rmn[r].y, rmn[r + 1].y), rmn[r + 2].y),
                                              fminf(fminf(rmn[r].z, rmn[r + 1].z), rmn[r + 2].z)}; };
            auto mx3 = [&](int r) { return V3{fmaxf(fmaxf(rmx[r].x, rmx[r + 1].x), rmx[r + 2].x), fmaxf(fmaxf(rmx[r].y, rmx[r + 1].y), rmx[r + 2].y),
                                              fmaxf(fmaxf(rmx[r].z, rmx[r + 1].z), rmx[r + 2].z)}; };
            V3P mn = pk3(mn3(k), mn3(k + 1)), mx = pk3(mx3(k), mx3(k + 1));
            const V3P ctr = scale3(add3(mn, mx), 0.5f), ext = scale3(sub3(mx, mn), 0.5f);             // :132-135
            mn = sub3(ctr, scale3(ext, 1.25f));
            mx = add3(ctr, scale3(ext, 1.25f));
            const V3P hy = rgb_to_ycocg(taa_tonemap_safe(h));
            h = taa_untonemap_pair(ycocg_to_rgb(clip_aabb_pair(mn, mx, hy)));                          // :140-141
        }
        const V3P o = add3(scale3(c, ib), scale3(h, blend));                                           // mix(), :159
        const uint2 oa = pack_rgba16f(lo(o.x), lo(o.y), lo(o.z), 1.0f), ob = pack_rgba16f(hi(o.x), hi(o.y), hi(o.z), 1.0f);
        const bool has_b = y + 1 < H;
        out[(size_t)y * W + x] = oa;
        if (has_b) out[(size_t)(y + 1) * W + x] = ob;
        if (THRESH) {                      // bloom threshold of the resolved texel as stored (fp16)
            const V3P b = soft_threshold_pair(pk3(unpack_rgb_(oa), unpack_rgb_(ob)), threshold);
            thr[(size_t)y * W + x] = pack_rgba16f(lo(b.x), lo(b.y), lo(b.z), 1.0f);
            if (has_b) thr[(size_t)(y + 1) * W + x] = pack_rgba16f(hi(b.x), hi(b.y), hi(b.z), 1.0f);
        }
    }
}

cudaError_t launch_taa_pair(const void* cur, const void* hist, void* out, void* thr_or_null,
                            const AshTaaPushConstants& pc, float threshold, const Launch& L) {
    const int W = (int)pc.screen_params[0], H = (int)pc.screen_params[1];
    const Tex16 c{static_cast<const uint2*>(cur), W, H}, h{static_cast<const uint2*>(hist), W, H};
    const dim3 grid((W + kT - 1) / kT, (H + kT - 1) / kT);
    uint2* o = static_cast<uint2*>(out);
    uint2* t = static_cast<uint2*>(thr_or_null);
    Timed timed(L, thr_or_null ? ASH_KERNEL_TAA_THRESHOLD : ASH_KERNEL_TAA);
    if (t) k_taa_pair<true><<<grid, 256, 0, L.s>>>(c, h, o, t, pc, threshold);
    else   k_taa_pair<false><<<grid, 256, 0, L.s>>>(c, h, o, t, pc, threshold);
    return cudaGetLastError();
}

}  // namespace ash
