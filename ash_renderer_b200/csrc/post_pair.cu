// post_pair.cu — post kernels that evaluate two pixels per thread in lock-step on the packed fp32 pipe
// (FADD2 / FMUL2 / FFMA2, ash_f32x2.cuh).  Bit-identical to post_fast.cu / post.cu: the same pinned
// IEEE operation sequence per lane; only the instruction count changes.
#include <cstring>
#include "post_pair.cuh"

namespace ash {

__device__ __forceinline__ V3 unpack_rgb_(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

// =====================================================================================================
// TAA resolve (+ bloom threshold) — taa_resolve.comp:83-167 (+ bloom_threshold.frag:18-35), no velocity
// buffer (PIN-12), |jitter delta| <= 0.99 px so that every history tap lies in the tile's 1-texel halo.
// Tile = 32x32 pixels, 256 threads; thread = 1x4 strip = two (row k, row k+1) pairs.
// The history sampling coordinates are separable: 32 column entries and 32 row entries per tile are
// computed once (by 64 threads) instead of once per pixel.
// Staging map (no divisions): step it < 4: warp w stages halo row w + 8*it, lane = halo column (the 32x32
// block at the halo origin); step 4: threads 0..63 stage rows 32..33, threads 64..131 columns 32..33.
// =====================================================================================================
#ifndef ASH_TAA_TH
#define ASH_TAA_TH 16                                                  // tile height: 32 (256 threads, 114.8 us at 4K), 16 (128 threads, 109.4 us) or 8 (64 threads, 114.0 us)
#endif
constexpr int kT = 32, kHalo = kT + 2;                                 // tile width; 34 staged columns
constexpr int kTH = ASH_TAA_TH, kHaloH = kTH + 2, kTaaThreads = 8 * kTH, kTaaWarps = kTaaThreads / 32;

struct AxisTap { int i0, i1; float f; float pad; };                  // halo-local texel BYTE offsets (column: 8 i, row: 8 kHalo i) + 8-bit weight
struct TaaTables { AxisTap col[kT], row[kTH]; float blend; };
struct TaaRaw { uint2 cur[kHaloH][kHalo], hist[kHaloH][kHalo]; };       // RGBA16F texels, tile + halo

// halo entries beyond the 32-column x kTH-row block: 64 of rows kTH, kTH+1 and 2 * kHaloH of columns 32, 33, one per
// thread (staging them as packed pairs on half the threads was measured slower: 106.8 vs 105.0 us, it lengthens the
// two warps that also build the coordinate tables)
constexpr int kExtraEntries = 64 + 2 * kHaloH, kExtraSlots = (kExtraEntries + kTaaThreads - 1) / kTaaThreads;
struct StageMap {
    int lane, warp, xr[kExtraSlots], rr[kExtraSlots]; bool extra[kExtraSlots];
    __device__ __forceinline__ StageMap() {
        lane = threadIdx.x & 31; warp = threadIdx.x >> 5;
#pragma unroll
        for (int n = 0; n < kExtraSlots; ++n) {
            const int e = (int)threadIdx.x + n * kTaaThreads;
            xr[n] = e < 64 ? (e & 31) : 32 + (e & 1);
            rr[n] = e < 64 ? kTH + (e >> 5) : (e - 64) >> 1;
            extra[n] = e < kExtraEntries;
        }
    }
};

// coordinate tables of one tile, :92-102 + PIN-6 (threads 0..63; thread 0 also the blend factor, :145-150)
__device__ __forceinline__ void taa_tables(TaaTables& tb, const AshTaaPushConstants& pc, int W, int H, int x0, int y0) {
    if (threadIdx.x >= kT + kTH) return;
    const bool is_row = threadIdx.x >= kT;
    const int k = is_row ? (int)threadIdx.x - kT : (int)threadIdx.x;
    const float ext = is_row ? pc.screen_params[1] : pc.screen_params[0];
    const int n = is_row ? H : W, p0 = is_row ? y0 : x0;
    const float jd = is_row ? fsub(pc.jitter[1], pc.prev_jitter[1]) : fsub(pc.jitter[0], pc.prev_jitter[0]);
    const float u = fdiv_safe(fadd((float)(p0 + k), 0.5f), ext);            // :92 (extent in [1, 32768])
    const float vel = fdiv_safe(jd, ext);                                    // :97-99
    const float hu = gclamp(fsub(u, vel), 0.0f, 1.0f);                       // :102
    AxisTap t;
    bilinear_setup(hu, n, t.i0, t.i1, t.f);
    const int pitch = is_row ? 8 * kHalo : 8;                                // byte offsets: one add per tap address in the strip
    t.i0 = min(max(t.i0 - (p0 - 1), 0), (is_row ? kHaloH : kHalo) - 1) * pitch;                  // halo-local (the clamp only guards rows / columns past the extent)
    t.i1 = min(max(t.i1 - (p0 - 1), 0), (is_row ? kHaloH : kHalo) - 1) * pitch;
    t.pad = 0.0f;
    if (is_row) tb.row[k] = t; else tb.col[k] = t;
    if (threadIdx.x == 0) {
        float blend = pc.params[0];
        if (pc.params[2] > 0.5f) {
            const float sw = pc.screen_params[0], sh = pc.screen_params[1];
            const float vx = fdiv_safe(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);
            const float vy = fdiv_safe(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
            const float vl = length2(fmul(vx, sw), fmul(vy, sh));
            blend = fmul(blend, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
        }
        tb.blend = blend;
    }
}

// tonemapped YCoCg (:120) of this thread's staged `current` texels -> s_nb, two at a time
__device__ __forceinline__ void taa_stage_nb(float4 (*s_nb)[kHalo], const uint2 (&tc)[4 + kExtraSlots], const StageMap& m) {
#pragma unroll
    for (int it = 0; it < 4; it += 2) {
        const int r0 = m.warp + kTaaWarps * it, r1 = r0 + kTaaWarps;
        const V3P s = rgb_to_ycocg(taa_tonemap_safe(pk3(unpack_rgb_(tc[it]), unpack_rgb_(tc[it + 1]))));
        s_nb[r0][m.lane] = make_float4(lo(s.x), lo(s.y), lo(s.z), 0.0f);
        s_nb[r1][m.lane] = make_float4(hi(s.x), hi(s.y), hi(s.z), 0.0f);
    }
#pragma unroll
    for (int n = 0; n < kExtraSlots; ++n)
        if (m.extra[n]) {
            const V3 c3 = unpack_rgb_(tc[4 + n]);
            const V3 tm = {fdiv_safe(c3.x, fadd(1.0f, c3.x)), fdiv_safe(c3.y, fadd(1.0f, c3.y)), fdiv_safe(c3.z, fadd(1.0f, c3.z))};
            const V3 s = rgb_to_ycocg(tm);
            s_nb[m.rr[n]][m.xr[n]] = make_float4(s.x, s.y, s.z, 0.0f);
        }
}

// resolve this thread's 1x4 strip of the tile at (x0, y0).
// VEL = false: no velocity buffer, the history taps come from the staged halo through the per-tile tables.
// VEL = true : per-pixel velocity (RG16F, centre tap == texel): velocity test (:97-99), history UV (:102), 8-bit
//              bilinear setup and velocity-weighted blend (:145-150) per lane; the four history texels of each
//              lane are gathered from global memory (L2), the rest of the resolve is the same packed code.
template <bool THRESH, bool VEL>
__device__ __forceinline__ void taa_strip(const float4 (*s_nb)[kHalo], const TaaRaw& raw, const TaaTables& tb, bool clamp_on,
                                          int W, int H, int x0, int y0, uint2* __restrict__ out, uint2* __restrict__ thr,
                                          float threshold, const Tex16& hist, const __half2* __restrict__ vel,
                                          const AshTaaPushConstants& pc) {
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int x = x0 + tx, yb = y0 + 4 * ty;
    if (x >= W || yb >= H) return;
    const AxisTap col = VEL ? AxisTap{0, 0, 0.0f, 0.0f} : tb.col[tx];
    const float blend = VEL ? 0.0f : tb.blend, ib = fsub(1.0f, blend);
    const float fx = col.f, ifx = fsub(1.0f, fx);
    // VEL: per-thread constants of the per-pixel coordinate math
    const float sw = pc.screen_params[0], sh = pc.screen_params[1];
    const float u = VEL ? fdiv_safe(fadd((float)x, 0.5f), sw) : 0.0f;                                 // :92 (extent in [1, 32768])
    const float jvx = VEL ? fdiv(fsub(pc.jitter[0], pc.prev_jitter[0]), sw) : 0.0f;                   // :97-99
    const float jvy = VEL ? fdiv(fsub(pc.jitter[1], pc.prev_jitter[1]), sh) : 0.0f;

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
        const AxisTap ra = VEL ? AxisTap{0, 0, 0.0f, 0.0f} : tb.row[4 * ty + k], rb = VEL ? ra : tb.row[4 * ty + k + 1];
        const V3P c = pk3(unpack_rgb_(raw.cur[4 * ty + k + 1][tx + 1]), unpack_rgb_(raw.cur[4 * ty + k + 2][tx + 1]));   // :105 (centre tap == texel)
        V3P h;                                                                                         // :108 bilinear history
        P2 blend2 = splat(blend), ib2 = splat(ib);
        if (!VEL) {
            const char* hb = reinterpret_cast<const char*>(&raw.hist[0][0]);
            auto tap = [&](int ro, int co) { return unpack_rgb_(*reinterpret_cast<const uint2*>(hb + ro + co)); };
            const V3P a  = pk3(tap(ra.i0, col.i0), tap(rb.i0, col.i0));
            const V3P b  = pk3(tap(ra.i0, col.i1), tap(rb.i0, col.i1));
            const V3P cc = pk3(tap(ra.i1, col.i0), tap(rb.i1, col.i0));
            const V3P d  = pk3(tap(ra.i1, col.i1), tap(rb.i1, col.i1));
            const P2 fy = pk(ra.f, rb.f), ify = fsub(1.0f, fy);
            // texel (11 significant bits) x weight k/256 (<= 9 bits) is exact in fp32 (no subnormals: >= 2^-32), so
            // the horizontal lerps round once with or without the FMA; the vertical ones keep both roundings
            const P2 fx2 = splat(fx);
            const V3P pa = scale3(a, ifx), pc2 = scale3(cc, ifx);
            const V3P top = {ffma(b.x, fx2, pa.x), ffma(b.y, fx2, pa.y), ffma(b.z, fx2, pa.z)};
            const V3P bot = {ffma(d.x, fx2, pc2.x), ffma(d.y, fx2, pc2.y), ffma(d.z, fx2, pc2.z)};
            h = add3(scale3(top, ify), scale3(bot, fy));
        } else {
            const int yl[2] = {y, min(y + 1, H - 1)};            // lane hi of a last odd row recomputes lane lo's row (not stored)
            int hx0[2], hx1[2], hy0[2], hy1[2]; float hfx[2], hfy[2], bl[2];
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                const float2 q = __half22float2(__ldg(vel + (size_t)yl[l] * W + x));
                float vx = q.x, vy = q.y;
                if (length2(vx, vy) < 0.0001f) { vx = jvx; vy = jvy; }                                 // :97-99
                const float v = fdiv_safe(fadd((float)yl[l], 0.5f), sh);                               // :92
                const float hu = gclamp(fsub(u, vx), 0.0f, 1.0f), hv = gclamp(fsub(v, vy), 0.0f, 1.0f);   // :102
                bilinear_setup(hu, W, hx0[l], hx1[l], hfx[l]);
                bilinear_setup(hv, H, hy0[l], hy1[l], hfy[l]);
                float b = pc.params[0];                                                                // :145-150
                if (pc.params[2] > 0.5f) {
                    const float vl = length2(fmul(vx, sw), fmul(vy, sh));
                    b = fmul(b, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
                }
                if (hu < 0.0f || hu > 1.0f || hv < 0.0f || hv > 1.0f) b = 0.0f;                        // :153-156 (dead after the clamp)
                bl[l] = b;
            }
            auto tap = [&](int l, int yy, int xx) { return unpack_rgb_(__ldg(hist.p + (size_t)yy * W + xx)); };
            const V3P a  = pk3(tap(0, hy0[0], hx0[0]), tap(1, hy0[1], hx0[1]));
            const V3P b  = pk3(tap(0, hy0[0], hx1[0]), tap(1, hy0[1], hx1[1]));
            const V3P cc = pk3(tap(0, hy1[0], hx0[0]), tap(1, hy1[1], hx0[1]));
            const V3P d  = pk3(tap(0, hy1[0], hx1[0]), tap(1, hy1[1], hx1[1]));
            const P2 fx2 = pk(hfx[0], hfx[1]), ifx2 = fsub(1.0f, fx2), fy = pk(hfy[0], hfy[1]), ify = fsub(1.0f, fy);
            const V3P pa = scale3(a, ifx2), pc2 = scale3(cc, ifx2);                                   // exact products, as above
            const V3P top = {ffma(b.x, fx2, pa.x), ffma(b.y, fx2, pa.y), ffma(b.z, fx2, pa.z)};
            const V3P bot = {ffma(d.x, fx2, pc2.x), ffma(d.y, fx2, pc2.y), ffma(d.z, fx2, pc2.z)};
            h = add3(scale3(top, ify), scale3(bot, fy));
            blend2 = pk(bl[0], bl[1]);
            ib2 = fsub(1.0f, blend2);
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
        const V3P o = add3(scale3(c, ib2), scale3(h, blend2));                                         // mix(), :159
        const uint2 oa = pack_rgba16f(lo(o.x), lo(o.y), lo(o.z), 1.0f), ob = pack_rgba16f(hi(o.x), hi(o.y), hi(o.z), 1.0f);
        const bool has_b = y + 1 < H;
        const uint32_t o0 = (uint32_t)y * (uint32_t)W + (uint32_t)x, o1 = o0 + (uint32_t)W;        // < 2^32 texels: 32-bit offsets
        out[o0] = oa;
        if (has_b) out[o1] = ob;
        if (THRESH) {                      // bloom threshold of the resolved texel as stored (fp16)
            const V3P b = soft_threshold_pair(pk3(unpack_rgb_(oa), unpack_rgb_(ob)), threshold);
            thr[o0] = pack_rgba16f(lo(b.x), lo(b.y), lo(b.z), 1.0f);
            if (has_b) thr[o1] = pack_rgba16f(hi(b.x), hi(b.y), hi(b.z), 1.0f);
        }
    }
}

// A persistent cp.async double-buffered variant of this kernel (as k_upsample_pair, post_up_pair.cu) was
// measured 20 % SLOWER (0.140 vs 0.116 ms at 4K): three barriers per tile and the extra shared-memory
// pass outweigh the hidden load latency; the one-shot kernel below keeps its loads in registers.
#ifndef ASH_TAA_MINB
#define ASH_TAA_MINB 3
#endif
template <bool THRESH, bool VEL>
#ifndef ASH_TAA_CTAS
#define ASH_TAA_CTAS (ASH_TAA_MINB * 256 / (8 * ASH_TAA_TH))
#endif
__global__ void __launch_bounds__(kTaaThreads, ASH_TAA_CTAS)
k_taa_pair(Tex16 cur, Tex16 hist, const __half2* __restrict__ vel, uint2* __restrict__ out, uint2* __restrict__ thr,
           AshTaaPushConstants pc, float threshold) {
    __shared__ float4 s_nb[kHaloH][kHalo];     // tonemapped YCoCg of `current`, tile + halo
    __shared__ TaaRaw s_raw;
    __shared__ TaaTables s_tb;
    const int W = cur.W, H = cur.H;
    const int x0 = blockIdx.x * kT, y0 = blockIdx.y * kTH;
    const bool clamp_on = pc.params[1] > 0.5f;
    const StageMap m;
    {   // stage: all global loads of the CTA are issued here, back to back
        uint2 tc[4 + kExtraSlots], th[4 + kExtraSlots];
        const int gx = min(max(x0 - 1 + m.lane, 0), W - 1);                                            // CLAMP_TO_EDGE
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const uint32_t g = (uint32_t)min(max(y0 - 1 + m.warp + kTaaWarps * it, 0), H - 1) * (uint32_t)W + (uint32_t)gx;
            tc[it] = __ldg(cur.p + g);
            if (!VEL) th[it] = __ldg(hist.p + g);
        }
#pragma unroll
        for (int n = 0; n < kExtraSlots; ++n) {
            tc[4 + n] = th[4 + n] = make_uint2(0u, 0u);
            if (m.extra[n]) {
                const uint32_t g = (uint32_t)min(max(y0 - 1 + m.rr[n], 0), H - 1) * (uint32_t)W + (uint32_t)min(max(x0 - 1 + m.xr[n], 0), W - 1);
                tc[4 + n] = __ldg(cur.p + g);
                if (!VEL) th[4 + n] = __ldg(hist.p + g);
            }
        }
        if (!VEL) taa_tables(s_tb, pc, W, H, x0, y0);      // while the loads fly
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            s_raw.cur[m.warp + kTaaWarps * it][m.lane] = tc[it];
            if (!VEL) s_raw.hist[m.warp + kTaaWarps * it][m.lane] = th[it];
        }
#pragma unroll
        for (int n = 0; n < kExtraSlots; ++n)
            if (m.extra[n]) { s_raw.cur[m.rr[n]][m.xr[n]] = tc[4 + n]; if (!VEL) s_raw.hist[m.rr[n]][m.xr[n]] = th[4 + n]; }
        if (clamp_on) taa_stage_nb(s_nb, tc, m);
    }
    __syncthreads();
    taa_strip<THRESH, VEL>(s_nb, s_raw, s_tb, clamp_on, W, H, x0, y0, out, thr, threshold, hist, vel, pc);
}

// graph replay: argument 5 of k_taa_pair is the per-view AshTaaPushConstants
void taa_pair_patch(bool thresh, bool vel, const AshTaaPushConstants& pc, PatchArg* out) {
    out->func = thresh ? (vel ? reinterpret_cast<const void*>(&k_taa_pair<true, true>) : reinterpret_cast<const void*>(&k_taa_pair<true, false>))
                       : (vel ? reinterpret_cast<const void*>(&k_taa_pair<false, true>) : reinterpret_cast<const void*>(&k_taa_pair<false, false>));
    out->arg_index = 5; out->n_args = 7;
    std::memcpy(out->value, &pc, sizeof(pc));
}

cudaError_t launch_taa_pair(const void* cur, const void* hist, const void* vel_or_null, void* out, void* thr_or_null,
                            const AshTaaPushConstants& pc, float threshold, const Launch& L) {
    const int W = (int)pc.screen_params[0], H = (int)pc.screen_params[1];
    const Tex16 c{static_cast<const uint2*>(cur), W, H}, h{static_cast<const uint2*>(hist), W, H};
    const dim3 grid((W + kT - 1) / kT, (H + kTH - 1) / kTH);
    uint2* o = static_cast<uint2*>(out);
    uint2* t = static_cast<uint2*>(thr_or_null);
    Timed timed(L, thr_or_null ? ASH_KERNEL_TAA_THRESHOLD : ASH_KERNEL_TAA);
    const __half2* v = static_cast<const __half2*>(vel_or_null);
    if (v) { if (t) k_taa_pair<true, true><<<grid, kTaaThreads, 0, L.s>>>(c, h, v, o, t, pc, threshold);
             else   k_taa_pair<false, true><<<grid, kTaaThreads, 0, L.s>>>(c, h, v, o, t, pc, threshold); }
    else   { if (t) k_taa_pair<true, false><<<grid, kTaaThreads, 0, L.s>>>(c, h, v, o, t, pc, threshold);
             else   k_taa_pair<false, false><<<grid, kTaaThreads, 0, L.s>>>(c, h, v, o, t, pc, threshold); }
    return cudaGetLastError();
}

// =====================================================================================================
// Bloom downsample, source extent == 2 x destination extent, destination width even
// (bloom_downsample.frag:16-108; the tap algebra is k_downsample_fast's, post_fast.cu).
// One thread = two horizontally adjacent destination pixels (lanes) = one 6x4 source window:
// lane lo uses window columns 0..3, lane hi columns 2..5, fetched with 16 independent loads
// (8 B, 16 B, 16 B, 8 B per row).
// =====================================================================================================
#ifndef ASH_DOWN_MINB
#define ASH_DOWN_MINB 3
#endif
__global__ void __launch_bounds__(256, ASH_DOWN_MINB)
k_downsample_pair(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int dw, int dh) {
    const int X = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    const int x = 2 * X;
    if (x >= dw || y >= dh) return;
    // window column i = source texel 4X - 1 + i, CLAMP_TO_EDGE (only columns 0 and 5 can leave the image)
    const int xl = max(4 * X - 1, 0), xr = min(4 * X + 4, sw - 1);
    uint2 e0[4], e5[4]; uint4 ma[4], mb[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint2* row = src + (size_t)min(max(2 * y - 1 + j, 0), sh - 1) * sw;
        e0[j] = __ldg(row + xl);
        ma[j] = __ldg(reinterpret_cast<const uint4*>(row + 4 * X));          // texels 4X, 4X+1 (32-byte aligned)
        mb[j] = __ldg(reinterpret_cast<const uint4*>(row + 4 * X + 2));      // texels 4X+2, 4X+3
        e5[j] = __ldg(row + xr);
    }
    V3P hs[4][3];    // hs[j][i] = t[j][i] + t[j][i+1] of each lane's own window
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const V3 t0 = unpack_rgb_(e0[j]), t1 = unpack_rgb_(make_uint2(ma[j].x, ma[j].y)), t2 = unpack_rgb_(make_uint2(ma[j].z, ma[j].w));
        const V3 t3 = unpack_rgb_(make_uint2(mb[j].x, mb[j].y)), t4 = unpack_rgb_(make_uint2(mb[j].z, mb[j].w)), t5 = unpack_rgb_(e5[j]);
        const V3P a0 = pk3(t0, t2), a1 = pk3(t1, t3), a2 = pk3(t2, t4), a3 = pk3(t3, t5);
        hs[j][0] = add3(a0, a1); hs[j][1] = add3(a1, a2); hs[j][2] = add3(a2, a3);
    }
    V3P res = {splat(0.0f), splat(0.0f), splat(0.0f)};
    P2 sum = splat(0.0f);
    auto tap = [&](int i, int j, float wt) {
        const V3P t4 = add3(hs[j][i], hs[j + 1][i]);
        const P2 luma4 = fadd(fadd(fmul(t4.x, 0.2126f), fmul(t4.y, 0.7152f)), fmul(t4.z, 0.0722f));
        const P2 w = frcp_safe(ffma(luma4, splat(0.25f), splat(1.0f)));         // 1 / (1 + luma), luma in [0, 65504]   :87-101
        const P2 k = splat(0.25f * wt);                                          // power of two
        res = {ffma(fmul(t4.x, w), k, res.x), ffma(fmul(t4.y, w), k, res.y), ffma(fmul(t4.z, w), k, res.z)};
        sum = ffma(w, splat(wt), sum);
    };
    tap(1, 1, 0.25f);                                                        // centre
    tap(0, 0, 0.0625f); tap(2, 0, 0.0625f); tap(0, 2, 0.0625f); tap(2, 2, 0.0625f);   // a b c d   :25-28
    tap(0, 1, 0.125f);  tap(2, 1, 0.125f);  tap(1, 0, 0.125f);  tap(1, 2, 0.125f);    // e f g h   :31-34
    // sum in [1.5e-5, 1], res in [0, 65504]: safe domain                                          :107
    const V3P o = {fdiv_safe(res.x, sum), fdiv_safe(res.y, sum), fdiv_safe(res.z, sum)};
    const uint2 oa = pack_rgba16f(lo(o.x), lo(o.y), lo(o.z), 1.0f), ob = pack_rgba16f(hi(o.x), hi(o.y), hi(o.z), 1.0f);
    *reinterpret_cast<uint4*>(dst + (size_t)y * dw + x) = make_uint4(oa.x, oa.y, ob.x, ob.y);      // dw even: 16-byte aligned
}

cudaError_t launch_bloom_downsample_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                         uint32_t dh, const Launch& L) {
    const dim3 grid((dw / 2 + 31) / 32, (dh + 7) / 8);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_DOWNSAMPLE, dw, dh));
    k_downsample_pair<<<grid, 256, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh,
                                             static_cast<uint2*>(dst), (int)dw, (int)dh);
    return cudaGetLastError();
}

}  // namespace ash
