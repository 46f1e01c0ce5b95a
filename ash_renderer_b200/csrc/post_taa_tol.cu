// post_taa_tol.cu — TAA resolve (+ bloom threshold) in the TOLERANCE precision mode (AshCudaConfig.precision = 1).
//
// Same semantics as k_taa_pair (post_pair.cu) and taa_resolve.comp:83-167 / bloom_threshold.frag:18-35, but the
// colour arithmetic spends the tolerance BASELINE.json's north_star grants the post-processed images instead of
// reproducing the oracle's IEEE sequence bit for bit (the exact mode, precision = 0, stays the default of the
// library and the tests): results are within 1 fp16 ulp of the oracle on every texel (tests/test_parity_gpu.py,
// tolerance-mode cases).  What is relaxed:
//   * a / b        -> a * MUFU.RCP(b)                 (rcp.approx: 1 ulp of fp32, 2 instructions instead of 7)
//   * x*k + y      -> one FMA, everywhere             (the exact mode contracts only provably exact products)
//   * the clip box  center = ctr, half = 1.25 ext + 0.001  (taa_resolve.comp:132-135 then :65-66, algebraically equal)
//   * clipAABB      center + delta * min(1, min_i(half_i / |delta_i|)), branch-free  (= :68-76)
//   * 1/(4 knee + 1e-5) of the soft threshold is a compile-time constant.
// What is NOT relaxed: which texels are sampled and with which weights.  The history coordinates (:92-102, PIN-6 8-bit
// sub-texel weights) are evaluated per tile column / row exactly as in the exact kernels, the bloom threshold reads
// the resolved texel as stored (fp16), and a thread only ever combines the texels the shader would.
//
// Work split (one CTA = 32 x 16 pixels, 128 threads, thread = 1 x 4 strip of its own column):
//   * a thread loads ITS OWN four `current` / history texels (+ one halo texel for 100 of the 128 threads); the
//     `current` texels stay in registers as fp32 and serve both the neighbourhood staging and the resolve's centre
//     tap, so `current` is never written to shared memory in raw form and is converted from fp16 once;
//   * tonemapped YCoCg of every texel of tile + halo: once per texel into shared memory (float4);
//   * history: raw fp16 tile + halo in shared memory, four taps per pixel through the per-tile coordinate tables.
#include <algorithm>
#include <cstring>
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

#ifndef ASH_TOL_S
#define ASH_TOL_S 4                                                // rows per thread (strip height): 4 or 2
#endif
constexpr int kTolS = ASH_TOL_S, kTolP = kTolS / 2;               // P = lane distance of a pixel pair: lanes are rows k and k + P
constexpr int kTolW = 32, kTolThreads = 128, kTolH = (kTolThreads / 32) * kTolS;   // tile; thread = column tx, rows S ty .. S ty + S - 1
constexpr int kTolHaloW = kTolW + 2, kTolHaloH = kTolH + 2;
constexpr int kTolExtras = 2 * kTolHaloW + 2 * kTolH;             // halo ring (100 texels at S = 4), one per thread
static_assert(kTolExtras <= kTolThreads, "one halo texel per thread");

struct TolTap { int i0, i1; float f, pad; };                     // halo-local BYTE offsets (column: 8 i, row: 8 * 34 i), 8-bit weight
struct TolTables { TolTap col[kTolW], row[kTolH]; float blend, pad[3]; };

// coordinate tables of one tile: taa_resolve.comp:92-102 + PIN-6, the very expressions of the exact kernels
// (post_pair.cu: taa_tables) — sampling positions are not subject to the tolerance
__device__ __forceinline__ void tol_tables(TolTables& tb, const AshTaaPushConstants& pc, int W, int H, int x0, int y0) {
    if (threadIdx.x >= kTolW + kTolH) return;
    const bool is_row = threadIdx.x >= kTolW;
    const int k = is_row ? (int)threadIdx.x - kTolW : (int)threadIdx.x;
    const float ext = is_row ? pc.screen_params[1] : pc.screen_params[0];
    const int n = is_row ? H : W, p0 = is_row ? y0 : x0;
    const float jd = is_row ? fsub(pc.jitter[1], pc.prev_jitter[1]) : fsub(pc.jitter[0], pc.prev_jitter[0]);
    const float u = fdiv_safe(fadd((float)(p0 + k), 0.5f), ext);            // :92
    const float vel = fdiv_safe(jd, ext);                                    // :97-99
    const float hu = gclamp(fsub(u, vel), 0.0f, 1.0f);                       // :102
    TolTap t;
    bilinear_setup(hu, n, t.i0, t.i1, t.f);
    const int pitch = is_row ? 8 * kTolHaloW : 8, last = (is_row ? kTolHaloH : kTolHaloW) - 1;
    t.i0 = min(max(t.i0 - (p0 - 1), 0), last) * pitch;
    t.i1 = min(max(t.i1 - (p0 - 1), 0), last) * pitch;
    t.pad = 0.0f;
    if (is_row) tb.row[k] = t; else tb.col[k] = t;
    if (threadIdx.x == 0) {                                                  // :145-150
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

__device__ __forceinline__ V3 tol_unpack(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }

struct TolRingTable { unsigned short v[kTolThreads]; };
constexpr TolRingTable tol_make_ring() {
    TolRingTable t{};
    for (int e = 0; e < kTolThreads; ++e) {
        const int ex = e < 2 * kTolHaloW ? (e < kTolHaloW ? e : e - kTolHaloW) : (e < 2 * kTolHaloW + kTolH ? 0 : kTolHaloW - 1);
        const int ey = e < 2 * kTolHaloW ? (e < kTolHaloW ? 0 : kTolHaloH - 1) : 1 + ((e - 2 * kTolHaloW) & (kTolH - 1));
        t.v[e] = (unsigned short)(ex | (ey << 8));
    }
    return t;
}
__constant__ TolRingTable c_tol_ring = tol_make_ring();

// (y, co, cg) of taa_resolve.comp:34-40 from r + b, g / 2, r - b, written straight into the four registers of one
// 128-bit shared-memory store (the .w lane is never read)
__device__ __forceinline__ float4 tol_nb_record(float rb, float hg, float df) {
    float4 v;
    v.x = fmaf(rb, 0.25f, hg); v.y = 0.5f * df; v.z = fmaf(rb, -0.25f, hg); v.w = v.x;
    return v;
}

// ---- stage: own texels + halo ring -> s_h (raw history), s_nb (tonemapped YCoCg of `current`); returns the thread's own
// `current` texels as fp32
__device__ __forceinline__ void tol_stage(const Tex16& cur, const Tex16& hist, float4 (*s_nb)[kTolHaloW], uint2 (*s_h)[kTolHaloW],
                                          int x0, int y0, int tx, int ty, bool clamp_on, V3 (&c)[kTolS]) {
    using namespace tol;
    const int W = cur.W, H = cur.H;
    uint2 tc[kTolS], th[kTolS], ec = make_uint2(0u, 0u), eh = make_uint2(0u, 0u);
    const uint32_t gx = (uint32_t)min(x0 + tx, W - 1);                                                 // CLAMP_TO_EDGE
#pragma unroll
    for (int k = 0; k < kTolS; ++k) {
        const uint32_t g = (uint32_t)min(y0 + kTolS * ty + k, H - 1) * (uint32_t)W + gx;
        tc[k] = __ldg(cur.p + g);
        th[k] = __ldg(hist.p + g);
    }
    // halo ring: top row (34), bottom row (34), left column (16), right column (16): thread e takes ring texel e
    const uint32_t ring = c_tol_ring.v[threadIdx.x];
    const int ex = (int)(ring & 0xFFu), ey = (int)(ring >> 8);
    const bool extra = threadIdx.x < kTolExtras;
    if (extra) {
        const uint32_t g = (uint32_t)min(max(y0 - 1 + ey, 0), H - 1) * (uint32_t)W + (uint32_t)min(max(x0 - 1 + ex, 0), W - 1);
        ec = __ldg(cur.p + g);
        eh = __ldg(hist.p + g);
    }
#pragma unroll
    for (int k = 0; k < kTolS; ++k) {
        s_h[kTolS * ty + 1 + k][tx + 1] = th[k];
        c[k] = tol_unpack(tc[k]);
    }
    if (extra) s_h[ey][ex] = eh;
    if (clamp_on) {                                      // :116-121, once per texel
#pragma unroll
        for (int k = 0; k < kTolP; ++k) {                // lanes: rows k and k + P (the pairing of the resolve)
            const V3P t = tonemap(pk3(c[k], c[k + kTolP]));
            const P2 rb = add(t.x, t.z), hg = mul(t.y, splat(0.5f)), df = sub(t.x, t.z);
            s_nb[kTolS * ty + 1 + k][tx + 1] = tol_nb_record(lo(rb), lo(hg), lo(df));
            s_nb[kTolS * ty + 1 + kTolP + k][tx + 1] = tol_nb_record(hi(rb), hi(hg), hi(df));
        }
        if (extra) {
            const V3 q = tol_unpack(ec);
            const V3 t = {q.x * rcp(1.0f + q.x), q.y * rcp(1.0f + q.y), q.z * rcp(1.0f + q.z)};
            s_nb[ey][ex] = tol_nb_record(t.x + t.z, 0.5f * t.y, t.x - t.z);
        }
    }
}

// ---- min / max of the 3x3 neighbourhood of the strip's four pixels (:124-129): row-wise over the three horizontal
// neighbours for the six halo rows the strip touches, then column-wise
__device__ __forceinline__ void tol_neighbourhood(const float4 (*s_nb)[kTolHaloW], int tx, int ty, V3 (&mn)[kTolS], V3 (&mx)[kTolS]) {
    V3 rmn[kTolS + 2], rmx[kTolS + 2];
#pragma unroll
    for (int r = 0; r < kTolS + 2; ++r) {
        const float4 a = s_nb[kTolS * ty + r][tx], b = s_nb[kTolS * ty + r][tx + 1], d = s_nb[kTolS * ty + r][tx + 2];
        rmn[r] = {fminf(fminf(a.x, b.x), d.x), fminf(fminf(a.y, b.y), d.y), fminf(fminf(a.z, b.z), d.z)};
        rmx[r] = {fmaxf(fmaxf(a.x, b.x), d.x), fmaxf(fmaxf(a.y, b.y), d.y), fmaxf(fmaxf(a.z, b.z), d.z)};
    }
#pragma unroll
    for (int k = 0; k < kTolS; ++k) {
        mn[k] = {fminf(fminf(rmn[k].x, rmn[k + 1].x), rmn[k + 2].x), fminf(fminf(rmn[k].y, rmn[k + 1].y), rmn[k + 2].y),
                 fminf(fminf(rmn[k].z, rmn[k + 1].z), rmn[k + 2].z)};
        mx[k] = {fmaxf(fmaxf(rmx[k].x, rmx[k + 1].x), rmx[k + 2].x), fmaxf(fmaxf(rmx[k].y, rmx[k + 1].y), rmx[k + 2].y),
                 fmaxf(fmaxf(rmx[k].z, rmx[k + 1].z), rmx[k + 2].z)};
    }
}

// ---- clip the history of two pixels to their neighbourhood boxes, blend, store, threshold (:112-162 + bloom_threshold.frag)
template <bool THRESH>
__device__ __forceinline__ void tol_finish_pair(tol::V3P h, tol::V3P cp, bool clamp_on, tol::V3P mn, tol::V3P mx, P2 b2, P2 ib2,
                                                uint32_t o0, uint32_t o1, bool has1, uint2* __restrict__ out,
                                                uint2* __restrict__ thr, float threshold, float one) {
    using namespace tol;
    if (clamp_on) {
        const P2 hf = splat(0.5f), e125 = splat(0.625f), eps = splat(0.001f);
        const V3P ctr = {mul(add(mn.x, mx.x), hf), mul(add(mn.y, mx.y), hf), mul(add(mn.z, mx.z), hf)};
        const V3P half = {fma(sub(mx.x, mn.x), e125, eps), fma(sub(mx.y, mn.y), e125, eps), fma(sub(mx.z, mn.z), e125, eps)};
        const V3P hy = rgb_to_ycocg(tonemap(h));
        const V3P dl = {sub(hy.x, ctr.x), sub(hy.y, ctr.y), sub(hy.z, ctr.z)};
        // s = min(1, min_i half_i / |delta_i|): 1 inside the box (history kept), 1 / maxV outside (:68-76)
        const P2 sx = mul(half.x, rcp_abs(dl.x)), sy = mul(half.y, rcp_abs(dl.y)), sz = mul(half.z, rcp_abs(dl.z));
        const P2 s = pk(fminf(fminf(fminf(lo(sx), lo(sy)), lo(sz)), 1.0f), fminf(fminf(fminf(hi(sx), hi(sy)), hi(sz)), 1.0f));
        const V3P q = ycocg_to_rgb(V3P{fma(dl.x, s, ctr.x), fma(dl.y, s, ctr.y), fma(dl.z, s, ctr.z)});
        const P2 k1 = splat(1.0001f);                                                                // (1 - c) + 1e-4, :59-61
        h = {mul(q.x, rcp(sub(k1, q.x))), mul(q.y, rcp(sub(k1, q.y))), mul(q.z, rcp(sub(k1, q.z)))};
    }
    const V3P o = {fma(h.x, b2, mul(cp.x, ib2)), fma(h.y, b2, mul(cp.y, ib2)), fma(h.z, b2, mul(cp.z, ib2))};       // mix(), :159
    const uint2 oa = pack_rgba16f(lo(o.x), lo(o.y), lo(o.z), one), ob = pack_rgba16f(hi(o.x), hi(o.y), hi(o.z), one);
    out[o0] = oa;
    if (has1) out[o1] = ob;
    if (THRESH) {                                        // bloom_threshold.frag:18-27 of the resolved texel AS STORED (fp16)
        const V3 ta = tol_unpack(oa), tb = tol_unpack(ob);
        const P2 br = pk(fmaxf(fmaxf(ta.x, ta.y), ta.z), fmaxf(fmaxf(tb.x, tb.y), tb.z));
        const P2 over = sub(br, splat(threshold));                               // brightness - threshold
        const P2 sf = add(over, splat(0.5f));                                    // + knee
        const P2 sc = pk(fminf(fmaxf(lo(sf), 0.0f), 1.0f), fminf(fmaxf(hi(sf), 0.0f), 1.0f));
        const P2 soft = mul(mul(sc, sc), splat(1.0f / 2.00001f));                // soft^2 / (4 knee + 1e-5)
        const P2 ct = pk(fmaxf(lo(soft), lo(over)), fmaxf(hi(soft), hi(over)));
        const P2 w = mul(ct, pk(rcp(fmaxf(lo(br), 0.00001f)), rcp(fmaxf(hi(br), 0.00001f))));
        const V3P t3 = pk3(ta, tb);
        const P2 bx = mul(t3.x, w), by = mul(t3.y, w), bz = mul(t3.z, w);
        thr[o0] = pack_rgba16f(lo(bx), lo(by), lo(bz), one);
        if (has1) thr[o1] = pack_rgba16f(hi(bx), hi(by), hi(bz), one);
    }
}

#ifndef ASH_TOL_CTAS
#define ASH_TOL_CTAS 7                                   // measured at 4K: 5: 80.6 us, 6: 75.5, 7: 67.8, 8: 67.7, 9: 73.5, 10: 77.9 (S = 4)
#endif
// REGULAR = false: history taps through per-tile coordinate tables (any jitter delta inside the halo).
// REGULAR = true : the host has verified (exact_taps.cu: axis_taa_history_regular, the same fp32 expressions) that every
//   column samples texels clamp(x + dx), clamp(x + dx + 1) with ONE weight fx and every row clamp(y + dy), clamp(y + dy + 1)
//   with ONE weight fy — the usual case, the jitter delta being uniform over the frame.  Then the five history rows a
//   strip touches are lerped horizontally once and shared by vertically adjacent pixels, no table is built, and every
//   shared-memory address is a compile-time offset from one base.
template <bool THRESH, bool REGULAR>
__global__ void __launch_bounds__(kTolThreads, ASH_TOL_CTAS)
k_taa_tol(Tex16 cur, Tex16 hist, uint2* __restrict__ out, uint2* __restrict__ thr, TaaTolArgs a) {
    using namespace tol;
    __shared__ float4 s_nb[kTolHaloH][kTolHaloW];        // tonemapped YCoCg of `current`, tile + halo
    __shared__ uint2  s_h[kTolHaloH][kTolHaloW];         // raw history, tile + halo
    __shared__ TolTables s_tb;
    const int W = cur.W, H = cur.H;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int x0 = blockIdx.x * kTolW, y0 = blockIdx.y * kTolH;
    const int x = x0 + tx, yb = y0 + kTolS * ty;
    const bool clamp_on = a.pc.params[1] > 0.5f;

    V3 c[kTolS];                                           // this thread's own `current` texels, fp32, kept for the resolve
    if (!REGULAR) tol_tables(s_tb, a.pc, W, H, x0, y0);
    tol_stage(cur, hist, s_nb, s_h, x0, y0, tx, ty, clamp_on, c);
    __syncthreads();
    if (x >= W || yb >= H) return;

    V3 mn[kTolS], mx[kTolS];
    if (clamp_on) tol_neighbourhood(s_nb, tx, ty, mn, mx);
    float one;                                           // alpha 1.0 kept in ONE register for the eight F2FP packs (an immediate would be re-materialised per use)
    asm volatile("mov.f32 %0, 0f3F800000;" : "=f"(one));
    const float blend = REGULAR ? a.blend : s_tb.blend;
    const P2 b2 = splat(blend), ib2 = splat(1.0f - blend);
    const uint32_t o00 = (uint32_t)yb * (uint32_t)W + (uint32_t)x;

    if (REGULAR) {
        // rows j = 0..4 of the history window, lerped horizontally as (hl0,hl2), (hl1,hl3), (hl2,hl4)
        const uint2* hb = &s_h[kTolS * ty + 1 + a.dy][tx + 1 + a.dx];
        V3 ta[kTolS + 1], tb[kTolS + 1];
#pragma unroll
        for (int j = 0; j < kTolS + 1; ++j) { ta[j] = tol_unpack(hb[j * kTolHaloW]); tb[j] = tol_unpack(hb[j * kTolHaloW + 1]); }
        const P2 fx2 = splat(a.fx), ifx2 = splat(1.0f - a.fx), fy2 = splat(a.fy), ify2 = splat(1.0f - a.fy);
        V3P hl[kTolP + 1];
#pragma unroll
        for (int j = 0; j < kTolP + 1; ++j) {
            const V3P ta2 = pk3(ta[j], ta[j + kTolP]), tb2 = pk3(tb[j], tb[j + kTolP]);
            hl[j] = {fma(tb2.x, fx2, mul(ta2.x, ifx2)), fma(tb2.y, fx2, mul(ta2.y, ifx2)), fma(tb2.z, fx2, mul(ta2.z, ifx2))};
        }
#pragma unroll
        for (int k = 0; k < kTolP; ++k) {                // lanes: lo = row yb + k, hi = row yb + k + P
            if (yb + k >= H) break;
            const V3P h = {fma(hl[k + 1].x, fy2, mul(hl[k].x, ify2)), fma(hl[k + 1].y, fy2, mul(hl[k].y, ify2)),
                           fma(hl[k + 1].z, fy2, mul(hl[k].z, ify2))};
            tol_finish_pair<THRESH>(h, pk3(c[k], c[k + kTolP]), clamp_on, pk3(mn[k], mn[k + kTolP]), pk3(mx[k], mx[k + kTolP]), b2, ib2,
                                    o00 + (uint32_t)k * (uint32_t)W, o00 + (uint32_t)(k + kTolP) * (uint32_t)W, yb + k + kTolP < H, out, thr, a.threshold, one);
        }
    } else {
        const TolTap col = s_tb.col[tx];
        const P2 fx2 = splat(col.f), ifx2 = splat(1.0f - col.f);
        const char* hb = reinterpret_cast<const char*>(&s_h[0][0]);
#pragma unroll
        for (int k = 0; k < kTolP; ++k) {
            if (yb + k >= H) break;
            const TolTap ra = s_tb.row[kTolS * ty + k], rb = s_tb.row[kTolS * ty + k + kTolP];
            auto tap = [&](int ro, int co) { return tol_unpack(*reinterpret_cast<const uint2*>(hb + ro + co)); };
            const V3P t00 = pk3(tap(ra.i0, col.i0), tap(rb.i0, col.i0)), t01 = pk3(tap(ra.i0, col.i1), tap(rb.i0, col.i1));
            const V3P t10 = pk3(tap(ra.i1, col.i0), tap(rb.i1, col.i0)), t11 = pk3(tap(ra.i1, col.i1), tap(rb.i1, col.i1));
            const P2 fy = pk(ra.f, rb.f), ify = sub(splat(1.0f), fy);
            const V3P top = {fma(t01.x, fx2, mul(t00.x, ifx2)), fma(t01.y, fx2, mul(t00.y, ifx2)), fma(t01.z, fx2, mul(t00.z, ifx2))};
            const V3P bot = {fma(t11.x, fx2, mul(t10.x, ifx2)), fma(t11.y, fx2, mul(t10.y, ifx2)), fma(t11.z, fx2, mul(t10.z, ifx2))};
            const V3P h = {fma(bot.x, fy, mul(top.x, ify)), fma(bot.y, fy, mul(top.y, ify)), fma(bot.z, fy, mul(top.z, ify))};
            tol_finish_pair<THRESH>(h, pk3(c[k], c[k + kTolP]), clamp_on, pk3(mn[k], mn[k + kTolP]), pk3(mx[k], mx[k + kTolP]), b2, ib2,
                                    o00 + (uint32_t)k * (uint32_t)W, o00 + (uint32_t)(k + kTolP) * (uint32_t)W, yb + k + kTolP < H, out, thr, a.threshold, one);
        }
    }
}

// A persistent TMA-fed form (k_taa_tma, commit after eb64929: one thread issues two cp.async.bulk.tensor copies per tile
// that complete on an mbarrier, tile i + 1 in flight while tile i is resolved; bit-identical outputs) was measured SLOWER
// as well: 71.8 us under ncu against 64.8.  It does remove the DRAM wait (long scoreboard 2.31 -> 0.91 stalled warps per
// issue) but executes 14 % more instructions (per-tile coordinates, mbarrier polls, own texels re-read from shared
// memory) and stalls 0.9 on its two barriers per tile; with both inputs double-buffered (30.8 KB per CTA) 74.4 us and
// -5.6 % views/s, the shared memory being taken from the other views' kernels.  What bounds this kernel is the length of
// each warp's dependent chain at 7 warps per scheduler, not the memory system.  (TMA detail learnt on the way: the box
// origin must be 16-byte aligned in the innermost dimension — tools/probe/tma_probe.cu — hence 36-texel boxes from x0 - 2.)
// One MUFU.RCP per colour TRIPLE (r = 1/(xyz), 1/x = r yz, ...: 13.8 -> 6.3 reciprocals per pixel, SFU 43 % -> 19 % busy, six
// packed multiplies per pair in exchange; commit eb64929) was measured SLOWER: 70.8 vs 68.5 us.  The kernel waits on
// dependent chains, not on the SFU: the longer chain in front of every reciprocal costs more than the issue slots freed.
// A persistent, cp.async double-buffered variant of this kernel (tile t + 1's raw texels travelling into a second
// shared-memory stage while tile t is resolved; commit 'k_taa_tol_pipe') was measured SLOWER: 81.9 vs 75.7 us at 4K.
// It hides the DRAM wait at the head of each CTA (27 % of the stall samples here) but pays 12 % more instructions
// (own texels re-read from shared memory, per-tile address arithmetic in every thread) and a second barrier per tile.
static const void* tol_kernel(bool thresh, bool regular) {
    return thresh ? (regular ? reinterpret_cast<const void*>(&k_taa_tol<true, true>) : reinterpret_cast<const void*>(&k_taa_tol<true, false>))
                  : (regular ? reinterpret_cast<const void*>(&k_taa_tol<false, true>) : reinterpret_cast<const void*>(&k_taa_tol<false, false>));
}

// graph replay: argument 4 of k_taa_tol is the per-view TaaTolArgs
void taa_tol_patch(bool thresh, const TaaTolArgs& a, PatchArg* out) {
    out->func = tol_kernel(thresh, a.regular != 0);
    out->arg_index = 4; out->n_args = 5;
    static_assert(sizeof(TaaTolArgs) <= sizeof(out->value), "PatchArg too small");
    std::memcpy(out->value, &a, sizeof(a));
}

cudaError_t launch_taa_tol(const void* cur, const void* hist, void* out, void* thr_or_null, const TaaTolArgs& a, const Launch& L) {
    const int W = (int)a.pc.screen_params[0], H = (int)a.pc.screen_params[1];
    Tex16 c{static_cast<const uint2*>(cur), W, H}, h{static_cast<const uint2*>(hist), W, H};
    uint2* o = static_cast<uint2*>(out);
    uint2* t = static_cast<uint2*>(thr_or_null);
    TaaTolArgs args = a;
    void* params[5] = {&c, &h, &o, &t, &args};
    Timed timed(L, thr_or_null ? ASH_KERNEL_TAA_THRESHOLD : ASH_KERNEL_TAA);
    const dim3 grid((W + kTolW - 1) / kTolW, (H + kTolH - 1) / kTolH);
    return cudaLaunchKernel(tol_kernel(t != nullptr, a.regular != 0), grid, dim3(kTolThreads), params, 0, L.s);
}

}  // namespace ash
