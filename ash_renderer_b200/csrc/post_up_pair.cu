// post_up_pair.cu — the 1:2 tent upsample + additive blend (+ ACES tonemap / 8-bit pack) of post_fast.cu
// evaluated for TWO source texels per thread in lock-step on the packed fp32 pipe (ash_f32x2.cuh).
//
// The scalar form of this pipeline (k_upsample_pipe, removed) issued ~150 (upsample) / ~200 (with tonemap) instructions per destination
// pixel, 55 % of them scalar FADD / FMUL / FFMA, and was bound by instruction issue (ncu: 75 % issue-active).
// Here a thread owns source texels (i, j) and (i, j + 8) of a 16x16 tile: both lanes run exactly the scalar
// kernel's operation sequence (same taps, same order, same roundings), every add / mul / fma is one FADD2 /
// FFMA2 for the pair, and the shared-memory addresses of the second lane are immediate offsets of the first.
// Bit-identical to k_upsample_fast (post_fast.cu; bloom_upsample.frag:16-41, PIN-9; tonemapping.frag:20-48,
// PIN-11 for the 8-bit target).
//
// CTA = 128 threads = one 16x16 source tile = 32x32 destination pixels; persistent loop over tiles with the
// next tile's source / destination / HDR blocks in flight (cp.async) while the current one is computed.
// The source tile is converted to fp32 once per tile into the lane-paired layout the window reads use (UpSrcP).
#include "post_pair.cuh"

namespace ash {

__device__ __forceinline__ V3 up_unpack_(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }
__device__ __forceinline__ V3P up_unpack2_(uint2 a, uint2 b) { return pk3(up_unpack_(a), up_unpack_(b)); }

constexpr int kUpLane = 8;     // source-row distance between the two lanes of a thread

// 4 * lerp(a, b, g): g = 3/4 -> a + 3b ; g = 1/4 -> 3a + b   (3 * fp16 texel is exact: one rounding)
template <bool G75>
__device__ __forceinline__ V3P hlerp4p(V3P a, V3P b) {
    const P2 three = splat(3.0f);
    if (G75) return {ffma(b.x, three, a.x), ffma(b.y, three, a.y), ffma(b.z, three, a.z)};
    return {ffma(a.x, three, b.x), ffma(a.y, three, b.y), ffma(a.z, three, b.z)};
}
// 16 * lerp(H0, H1, g) from h = 4H: g = 3/4 -> h0 + RN(3 h1) ; g = 1/4 -> RN(3 h0) + h1   (two roundings)
template <bool G75>
__device__ __forceinline__ V3P vlerp16p(V3P h0, V3P h1) {
    if (G75) return {fadd(h0.x, fmul(h1.x, 3.0f)), fadd(h0.y, fmul(h1.y, 3.0f)), fadd(h0.z, fmul(h1.z, 3.0f))};
    return {fadd(fmul(h0.x, 3.0f), h1.x), fadd(fmul(h0.y, 3.0f), h1.y), fadd(fmul(h0.z, 3.0f), h1.z)};
}

struct UpHP {           // horizontal lerps (x4) of one destination-column parity, both lanes
    V3P m2[3], m1[3], z[5], p1[3], p2[3];   // tap m = -2,-1: window rows 1..3 ; m = 0: rows 0..4 ; m = +1,+2: rows 1..3
};

// The source tile as the lanes consume it: entry (p, c), p = 0..11, holds texel (p, c) of the staged 20x20 tile
// in the lo halves and texel (p + 8, c) in the hi halves, already converted to fp32 — one conversion per texel
// per tile instead of one per thread that has the texel in its 5x5 window (HADD2.F32 costs two FMA-pipe
// cycles, as much as an FFMA2), and the 64-bit lanes arrive in aligned register pairs straight from LDS.
struct UpSrcP {
    ulonglong2 xy[12][20];                   // {(x_p, x_{p+8}), (y_p, y_{p+8})}
    unsigned long long z[12][20];            // (z_p, z_{p+8})
};
__device__ __forceinline__ V3P up_window_(const UpSrcP& S, int p, int c) {
    const ulonglong2 a = S.xy[p][c];
    return {P2{a.x}, P2{a.y}, P2{S.z[p][c]}};
}

template <int PX>
__device__ __forceinline__ void up_hlerps_p(const UpSrcP& S, int tj, int ti, UpHP& h) {
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        V3P t[5];
#pragma unroll
        for (int c = 0; c < 4; ++c) t[c + PX] = up_window_(S, tj + r, ti + c + PX);      // window cols PX..PX+3
        h.z[r] = hlerp4p<up_g75(PX, 0)>(t[up_c0(PX, 0)], t[up_c0(PX, 0) + 1]);
        if (r >= 1 && r <= 3) {
            h.m2[r - 1] = hlerp4p<up_g75(PX, -2)>(t[up_c0(PX, -2)], t[up_c0(PX, -2) + 1]);
            h.m1[r - 1] = hlerp4p<up_g75(PX, -1)>(t[up_c0(PX, -1)], t[up_c0(PX, -1) + 1]);
            h.p1[r - 1] = hlerp4p<up_g75(PX, 1)>(t[up_c0(PX, 1)], t[up_c0(PX, 1) + 1]);
            h.p2[r - 1] = hlerp4p<up_g75(PX, 2)>(t[up_c0(PX, 2)], t[up_c0(PX, 2) + 1]);
        }
    }
}

// 256 * (tent sum / 16); rows via the same (c0, g) tables with p = PY   (up_combine, post_fast.cu)
template <int PY>
__device__ __forceinline__ V3P up_combine_p(const UpHP& h) {
#define ASH_UPV(arr, base, n) vlerp16p<up_g75(PY, n)>(arr[up_c0(PY, n) - (base)], arr[up_c0(PY, n) + 1 - (base)])
    const V3P s0 = ASH_UPV(h.m2, 1, 0);      // (-2d, 0)      bloom_upsample.frag:21
    const V3P s1 = ASH_UPV(h.m1, 1, -1);     // (-d, -d)
    const V3P s2 = ASH_UPV(h.z, 0, -2);      // (0, -2d)
    const V3P s3 = ASH_UPV(h.p1, 1, -1);     // (d, -d)
    const V3P s4 = ASH_UPV(h.p2, 1, 0);      // (2d, 0)
    const V3P s5 = ASH_UPV(h.p1, 1, 1);      // (d, d)
    const V3P s6 = ASH_UPV(h.z, 0, 2);       // (0, 2d)
    const V3P s7 = ASH_UPV(h.m1, 1, 1);      // (-d, d)
    const V3P s8 = ASH_UPV(h.z, 0, 0);       // (0, 0)
#undef ASH_UPV
    const V3P q = add3(add3(add3(s1, s3), s5), s7);                           // :33
    const V3P p = add3(add3(add3(s0, s2), s4), s6);                           // :34
    // (4 s8 + 2 q) + p: both products are exact (powers of two), so fma(s8, 4, 2q) rounds once, exactly like
    // the scalar fadd(fmul(s8, 4), fmul(q, 2))                                  :32-34
    const P2 four = splat(4.0f);
    return {fadd(ffma(s8.x, four, fmul(q.x, 2.0f)), p.x), fadd(ffma(s8.y, four, fmul(q.y, 2.0f)), p.y),
            fadd(ffma(s8.z, four, fmul(q.z, 2.0f)), p.z)};
}

// ---- tonemapping.frag:20-48 for the 8-bit target (PIN-11: within 1 LSB of the reference): the operation
//      sequence of tonemap8 (post_fast.cu) per lane; the polynomial part packed, MUFU / saturate / convert scalar
__device__ __forceinline__ float up_lg2_(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float up_ex2_(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float up_rcp_(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// Per channel, both lanes: x, ACES numerator / denominator, n * rcp(d), inv_gamma * lg2, 255 g + 2^23 are packed; the
// three MUFUs and the final clamp are per lane.  Relative to tonemap8 (post_fast.cu): the clamp of the ACES quotient is
// dropped (a > 1 gives g > 1, a < 0 or NaN gives NaN, and the clamp of g maps those to 1 and 0 exactly as clamping a
// first would, for gamma > 0), and the 8-bit value is read from the low mantissa byte of RN(255 g + 2^23), i.e. round
// to nearest even instead of floor(255 g + 0.5): they differ only on exact ties, inside the 1-LSB tolerance of PIN-11.
__device__ __forceinline__ P2 tonemap8_pair(P2 c, P2 bloom, const AshPostProcessPushConstants& pc, float inv_gamma, float magic) {
    const P2 x = fmul(ffma(bloom, splat(pc.bloom_intensity), c), pc.exposure);           // :38, :41
    const P2 n = fmul(x, ffma(splat(2.51f), x, splat(0.03f)));                           // :20-27
    const P2 d = ffma(x, ffma(splat(2.43f), x, splat(0.59f)), splat(0.14f));
    const P2 a = fmul(n, pk(up_rcp_(lo(d)), up_rcp_(hi(d))));
    const P2 e = fmul(pk(up_lg2_(lo(a)), up_lg2_(hi(a))), inv_gamma);                    // pow(a, 1/gamma)   :47
    const P2 g = pk(__saturatef(up_ex2_(lo(e))), __saturatef(up_ex2_(hi(e))));
    return ffma(g, splat(255.0f), splat(magic));                                         // low byte = 8-bit value (PIN-11)
}
__device__ __forceinline__ void tonemap_pack8_pair(V3P c, V3P bloom, const AshPostProcessPushConstants& pc, float inv_gamma,
                                                   uint32_t& o_lo, uint32_t& o_hi) {
    // 2^23 + 255 g has the 8-bit value in byte 0 and zeros in byte 1; for blue 2^23 + 0xFF00 + 255 g puts alpha = 0xFF there
    const P2 r = tonemap8_pair(c.x, bloom.x, pc, inv_gamma, 8388608.0f), g = tonemap8_pair(c.y, bloom.y, pc, inv_gamma, 8388608.0f),
             b = tonemap8_pair(c.z, bloom.z, pc, inv_gamma, 8388608.0f + 65280.0f);
    // bytes: r0, g0 <- byte 0 of r and g; b0, 0xFF <- bytes 0, 1 of b
    o_lo = __byte_perm(__byte_perm(__float_as_uint(lo(r)), __float_as_uint(lo(g)), 0x0040), __float_as_uint(lo(b)), 0x5410);
    o_hi = __byte_perm(__byte_perm(__float_as_uint(hi(r)), __float_as_uint(hi(g)), 0x0040), __float_as_uint(hi(b)), 0x5410);
}

// ---- TOLERANCE mode (AshCudaConfig.precision = FAST): the nine bilinear taps of bloom_upsample.frag:21-38 are linear in
// the source texels, so for each destination parity (py, px) the whole tent is ONE 4x4 convolution of the source window
// with constant weights (x 256, parity (0,0); the other parities mirror rows / columns):
//      0  1  3  0
//      1 28 46  9        13 non-zero weights: 13 FMAs per channel instead of the 9 x (2 horizontal + 1 vertical) lerps
//      3 46 74 17        + 8 adds of the exact mode.  Same taps, same weights (the exact-tap check of the level still
//      0  9 17  2        gates the kernel); only the order of the fp32 roundings differs: <= 1 fp16 ulp on the result.
// (= [h0 (x) (4 v0 + v-2 + v2) + 2 (h-1 + h1) (x) (v-1 + v1) + (h-2 + h2) (x) v0] / 16 with the 1-D tap vectors of up_c0 / up_g75.)
__host__ __device__ constexpr int up_tol_w00(int r, int c) {
    constexpr int W[4][4] = {{0, 1, 3, 0}, {1, 28, 46, 9}, {3, 46, 74, 17}, {0, 9, 17, 2}};
    return W[r][c];
}
__host__ __device__ constexpr int up_tol_w(int py, int px, int r, int c) { return up_tol_w00(py ? 3 - r : r, px ? 3 - c : c); }
__host__ __device__ constexpr int up_tol_slot(int w) {      // index of a weight value in the 8-entry register table
    return w == 1 ? 0 : w == 2 ? 1 : w == 3 ? 2 : w == 9 ? 3 : w == 17 ? 4 : w == 28 ? 5 : w == 46 ? 6 : 7;
}
__host__ __device__ constexpr float up_tol_value(int slot) { return slot == 0 ? 1.0f : slot == 1 ? 2.0f : slot == 2 ? 3.0f : slot == 3 ? 9.0f : slot == 4 ? 17.0f : slot == 5 ? 28.0f : slot == 6 ? 46.0f : 74.0f; }

// all four parities of one source texel, both lanes: up[py][px] = intensity * conv4x4(window rows py.., cols px..)
__device__ __forceinline__ void up_tol_conv(const UpSrcP& S, int tj, int ti, const P2 (&wv)[8], V3P (&up)[2][2]) {
    const P2 z = splat(0.0f);
#pragma unroll
    for (int py = 0; py < 2; ++py)
#pragma unroll
        for (int px = 0; px < 2; ++px) up[py][px] = {z, z, z};
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        V3P t[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) t[c] = up_window_(S, tj + r, ti + c);
#pragma unroll
        for (int py = 0; py < 2; ++py) {
            const int rr = r - py;
            if (rr < 0 || rr > 3) continue;
#pragma unroll
            for (int px = 0; px < 2; ++px)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int w = up_tol_w(py, px, rr, c);
                    if (w == 0) continue;
                    const P2 k = wv[up_tol_slot(w)];
                    V3P& a = up[py][px];
                    a = {ffma(t[c + px].x, k, a.x), ffma(t[c + px].y, k, a.y), ffma(t[c + px].z, k, a.z)};
                }
        }
    }
}

template <bool TONEMAP>
struct UpTileP {
    uint2 src[20][20];                       // 16x16 source texels + 2-texel halo
    uint4 dst[32][16];                       // 32x32 destination texels (the additive-blend operand)
    uint4 hdr[TONEMAP ? 32 : 1][16];         // 32x32 HDR texels (tonemap input)
};

#ifndef ASH_UPS_PAIR_MINB
#define ASH_UPS_PAIR_MINB 4
#endif
template <bool TONEMAP, bool STORE_U, bool TOL>
__global__ void __launch_bounds__(128, ASH_UPS_PAIR_MINB)
k_upsample_pair(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, float intensity,
                const uint2* __restrict__ hdr, uint32_t* __restrict__ ldr8, AshPostProcessPushConstants pc,
                int tiles_x, int ntiles) {
    __shared__ __align__(16) UpTileP<TONEMAP> T[2];
    __shared__ __align__(16) UpSrcP S;
    const int ti = threadIdx.x & 15, tj = threadIdx.x >> 4;          // tj in 0..7: lanes = source rows tj, tj + 8
    const int dw = 2 * sw, dh = 2 * sh;

    auto issue = [&](int tcx, int tcy, int b) {                      // tile column / row
        const int i0 = tcx * 16, j0 = tcy * 16;
#pragma unroll
        for (int k = 0; k < 4; ++k) {                                // 400 source texels, CLAMP_TO_EDGE
            const int idx = (int)threadIdx.x + 128 * k;
            if (k < 3 || idx < 400) {
                const int r = idx / 20, c = idx - r * 20;
                cp_async8(&T[b].src[r][c], src + ((uint32_t)min(max(j0 - 2 + r, 0), sh - 1) * (uint32_t)sw + (uint32_t)min(max(i0 - 2 + c, 0), sw - 1)));   // < 2^32 texels: 32-bit offsets
            }
        }
        if (i0 + ti < sw) {                                          // 32 rows x 16 blocks of 16 bytes (two texels)
            const uint32_t col = 2u * (uint32_t)(i0 + ti);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int br = tj + 8 * k;
                if (2 * j0 + br < dh) {
                    cp_async16(&T[b].dst[br][ti], dst + ((uint32_t)(2 * j0 + br) * (uint32_t)dw + col));
                    if (TONEMAP) cp_async16(&T[b].hdr[br][ti], hdr + ((uint32_t)(2 * j0 + br) * (uint32_t)dw + col));
                }
            }
        }
    };

    const P2 k2 = splat(fmul(intensity, 0.00390625f));            // (R16 / 256) * intensity == R16 * (intensity / 256)
    P2 wv[8];                                                     // TOL: the eight distinct tent weights, x intensity / 256
#pragma unroll
    for (int k = 0; k < 8; ++k) wv[k] = splat(TOL ? up_tol_value(k) * 0.00390625f * intensity : 0.0f);
    const float inv_gamma = TONEMAP ? __fdividef(1.0f, pc.gamma) : 0.0f;
    // tile coordinates advance by gridDim.x tiles per round without a division in the loop
    const int step_y = (int)gridDim.x / tiles_x, step_x = (int)gridDim.x - step_y * tiles_x;
    int tile = blockIdx.x, b = 0;
    int cy = tile / tiles_x, cx = tile - cy * tiles_x;
    if (tile < ntiles) issue(cx, cy, 0);
    cp_async_commit();
    for (; tile < ntiles; tile += gridDim.x, b ^= 1) {
        const int next = tile + gridDim.x;
        int nx = cx + step_x, ny = cy + step_y;
        if (nx >= tiles_x) { nx -= tiles_x; ++ny; }
        if (next < ntiles) issue(nx, ny, b ^ 1);
        cp_async_commit();
        cp_async_wait<1>();                                       // this thread's copies of `tile` have landed
        __syncthreads();                                          // ... and everybody else's
#pragma unroll
        for (int k = 0; k < 2; ++k) {                             // fp16 -> paired fp32, 240 entries
            const int e = (int)threadIdx.x + 128 * k;
            if (k == 0 || e < 240) {
                const int p = e / 20, c = e - p * 20;
                const V3P t = up_unpack2_(T[b].src[p][c], T[b].src[p + kUpLane][c]);
                S.xy[p][c] = make_ulonglong2(t.x.v, t.y.v);
                S.z[p][c] = t.z.v;
            }
        }
        __syncthreads();
        const int i = cx * 16 + ti, j = cy * 16 + tj;
        cx = nx; cy = ny;
        if (i < sw && j < sh) {
            const bool hi_ok = j + kUpLane < sh;                  // lane hi past the last row: computed on the clamped tile, not stored
            V3P up[2][2];    // [py][px] upsampled value of both lanes, before the additive blend
            if (TOL) {
                up_tol_conv(S, tj, ti, wv, up);
            } else {
                {
                    UpHP h;
                    up_hlerps_p<0>(S, tj, ti, h);
                    up[0][0] = scale3(up_combine_p<0>(h), k2);
                    up[1][0] = scale3(up_combine_p<1>(h), k2);
                }
                {
                    UpHP h;
                    up_hlerps_p<1>(S, tj, ti, h);
                    up[0][1] = scale3(up_combine_p<0>(h), k2);
                    up[1][1] = scale3(up_combine_p<1>(h), k2);
                }
            }
#pragma unroll
            for (int py = 0; py < 2; ++py) {
                const int rl = 2 * tj + py, rh = rl + 2 * kUpLane;                     // tile-local destination rows of the lanes
                const uint32_t ol = (uint32_t)(2 * j + py) * (uint32_t)dw + 2u * (uint32_t)i;   // 16-byte aligned pair of texels
                const uint32_t oh = ol + (uint32_t)(2 * kUpLane) * (uint32_t)dw;
                const uint4 dl = T[b].dst[rl][ti], dhh = T[b].dst[rh][ti];
                const V3P u0 = add3(up[py][0], up_unpack2_(make_uint2(dl.x, dl.y), make_uint2(dhh.x, dhh.y)));
                const V3P u1 = add3(up[py][1], up_unpack2_(make_uint2(dl.z, dl.w), make_uint2(dhh.z, dhh.w)));
                const uint2 u0l = pack_rgba16f(lo(u0.x), lo(u0.y), lo(u0.z), 1.0f), u0h = pack_rgba16f(hi(u0.x), hi(u0.y), hi(u0.z), 1.0f);
                const uint2 u1l = pack_rgba16f(lo(u1.x), lo(u1.y), lo(u1.z), 1.0f), u1h = pack_rgba16f(hi(u1.x), hi(u1.y), hi(u1.z), 1.0f);
                if (STORE_U) {
                    *reinterpret_cast<uint4*>(dst + ol) = make_uint4(u0l.x, u0l.y, u1l.x, u1l.y);
                    if (hi_ok) *reinterpret_cast<uint4*>(dst + oh) = make_uint4(u0h.x, u0h.y, u1h.x, u1h.y);
                }
                if (TONEMAP) {
                    const uint4 cl = T[b].hdr[rl][ti], ch = T[b].hdr[rh][ti];
                    uint32_t a_lo, a_hi, b_lo, b_hi;
                    // the bloom operand is U_0 as stored (fp16); TOL feeds the value before that rounding (half an fp16 ulp: far
                    // inside the 1-LSB bar of the 8-bit result)
                    tonemap_pack8_pair(up_unpack2_(make_uint2(cl.x, cl.y), make_uint2(ch.x, ch.y)), TOL ? u0 : up_unpack2_(u0l, u0h), pc, inv_gamma, a_lo, a_hi);
                    tonemap_pack8_pair(up_unpack2_(make_uint2(cl.z, cl.w), make_uint2(ch.z, ch.w)), TOL ? u1 : up_unpack2_(u1l, u1h), pc, inv_gamma, b_lo, b_hi);
                    *reinterpret_cast<uint2*>(ldr8 + ol) = make_uint2(a_lo, b_lo);
                    if (hi_ok) *reinterpret_cast<uint2*>(ldr8 + oh) = make_uint2(a_hi, b_hi);
                }
            }
        }
        __syncthreads();                                          // buffer b is refilled by the next iteration's issue, S by its conversion
    }
    cp_async_wait<0>();
}

template <bool TONEMAP, bool STORE_U, bool TOL>
static cudaError_t launch_pair_(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity,
                                const void* hdr, void* ldr8, const AshPostProcessPushConstants& pc, cudaStream_t s) {
    const int tiles_x = (int)((sw + 15) / 16), ntiles = tiles_x * (int)((sh + 15) / 16);
    // resident CTAs per SM: launch bound, capped by 2 x sizeof(UpTileP) of shared memory
    const int per_sm = TONEMAP ? (ASH_UPS_PAIR_MINB < 5 ? ASH_UPS_PAIR_MINB : 5) : ASH_UPS_PAIR_MINB;
    const int grid = min(ntiles, 148 * per_sm);
    k_upsample_pair<TONEMAP, STORE_U, TOL><<<grid, 128, 0, s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh, static_cast<uint2*>(dst),
                                                            intensity, static_cast<const uint2*>(hdr), static_cast<uint32_t*>(ldr8),
                                                            pc, tiles_x, ntiles);
    return cudaGetLastError();
}

// enough tiles to keep every resident CTA busy for at least two rounds of the pipeline
bool upsample_pair_applies(uint32_t sw, uint32_t sh) {
    return ((sw + 15) / 16) * ((sh + 15) / 16) >= 148u * ASH_UPS_PAIR_MINB * 2u;
}

cudaError_t launch_upsample_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity, bool tol, cudaStream_t s) {
    return tol ? launch_pair_<false, true, true>(src, sw, sh, dst, intensity, nullptr, nullptr, AshPostProcessPushConstants{}, s)
               : launch_pair_<false, true, false>(src, sw, sh, dst, intensity, nullptr, nullptr, AshPostProcessPushConstants{}, s);
}

cudaError_t launch_upsample_tonemap_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity, const void* hdr,
                                         void* ldr8, const AshPostProcessPushConstants& pc, bool store_u, bool tol, cudaStream_t s) {
    if (tol) return store_u ? launch_pair_<true, true, true>(src, sw, sh, dst, intensity, hdr, ldr8, pc, s)
                            : launch_pair_<true, false, true>(src, sw, sh, dst, intensity, hdr, ldr8, pc, s);
    return store_u ? launch_pair_<true, true, false>(src, sw, sh, dst, intensity, hdr, ldr8, pc, s)
                   : launch_pair_<true, false, false>(src, sw, sh, dst, intensity, hdr, ldr8, pc, s);
}

}  // namespace ash
