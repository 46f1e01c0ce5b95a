// post_down_box.cu — 2:1 Karis bloom downsample (bloom_downsample.frag:16-108) with every 2x2 box evaluated ONCE.
//
// With the source exactly twice the destination (exact_taps.cu: axis_down_taps_exact) tap (i, j), i, j in 0..2,
// of destination pixel (X, Y) is the 2x2 box of source texels whose top-left corner is (2X - 1 + i, 2Y - 1 + j),
// weighted by 1 / (1 + luma(box)).  Boxes at odd positions are shared by neighbouring destination pixels: a tile
// of 32x8 pixels touches 65x17 = 1105 distinct boxes, 4.3 per pixel, where k_downsample_pair (post_pair.cu)
// evaluates 9 per pixel — and the box sum, its luma, the correctly rounded reciprocal and the three weighted
// colour products are 2/3 of that kernel's arithmetic.  The value of a box does not depend on the pixel that
// uses it (same operands, same operation sequence), and each pixel still accumulates its nine taps in the
// shader's order, so the result is bit-identical.
//
// CTA = 256 threads = TWO horizontally adjacent tiles of 32x8 destination pixels; lane lo of every packed
// register (ash_f32x2.cuh) belongs to the left tile, lane hi to the right one, at the same tile-local
// coordinates, so the shared-memory box records are (left, right) pairs and the code is symmetric.
//   phase 1: thread (tx, ty) evaluates the 2x2 block of boxes {2tx, 2tx+1} x {2ty, 2ty+1} (tile-local box
//            coordinates) from its 3x3 source texels; 81 threads add one box of column 64 / row 16;
//   phase 2: thread (tx, ty) accumulates the nine taps of its pixel (four from its own registers, five from
//            shared memory), divides, stores.
#include "post_pair.cuh"

namespace ash {

__device__ __forceinline__ V3 box_unpack_(uint2 t) { V4 v = unpack_rgba16f(t); return {v.x, v.y, v.z}; }
__device__ __forceinline__ V3P box_unpack2_(uint2 a, uint2 b) { return pk3(box_unpack_(a), box_unpack_(b)); }

#ifndef ASH_DOWN_BOX_TH
#define ASH_DOWN_BOX_TH 8                       // destination rows per tile: 8 (256 threads) or 4 (128 threads: measured equal, 4772 vs 4788 views/s)
#endif
constexpr int kBH = ASH_DOWN_BOX_TH, kBoxRows = 2 * kBH + 1, kBoxThreads = 32 * kBH;

struct BoxTile {                              // component k of a box record: 0..2 = colour x weight, 3 = weight; (left, right) pairs
    unsigned long long e[4][kBoxRows][33];          // even box columns 0, 2, .., 64   (component-major: 8-byte accesses of a
    unsigned long long o[4][kBoxRows][32];          // odd box columns 1, 3, .., 63     warp are contiguous, no bank conflicts)
};
struct BoxRec { P2 x, y, z, w; };

// weighted record of one box from its 4x sum: luma, 1 / (1 + luma), colour x weight   (:87-101)
// TOL (AshCudaConfig.precision = FAST): the dot product as an FMA chain, the reciprocal as MUFU.RCP (1 ulp of fp32)
__device__ __forceinline__ float box_rcp_(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
template <bool TOL>
__device__ __forceinline__ BoxRec box_record_(V3P t4) {
    if (TOL) {
        const P2 d = ffma(t4.x, splat(0.25f * 0.2126f), ffma(t4.y, splat(0.25f * 0.7152f), ffma(t4.z, splat(0.25f * 0.0722f), splat(1.0f))));
        const P2 w = pk(box_rcp_(lo(d)), box_rcp_(hi(d)));
        return {fmul(t4.x, w), fmul(t4.y, w), fmul(t4.z, w), w};
    }
    const P2 luma4 = fadd(fadd(fmul(t4.x, 0.2126f), fmul(t4.y, 0.7152f)), fmul(t4.z, 0.0722f));
    const P2 w = frcp_safe(ffma(luma4, splat(0.25f), splat(1.0f)));        // luma in [0, 65504]: safe domain
    return {fmul(t4.x, w), fmul(t4.y, w), fmul(t4.z, w), w};
}
__device__ __forceinline__ void box_store_(BoxTile& B, bool odd, int r, int c, const BoxRec& v) {
    if (odd) { B.o[0][r][c] = v.x.v; B.o[1][r][c] = v.y.v; B.o[2][r][c] = v.z.v; B.o[3][r][c] = v.w.v; }
    else     { B.e[0][r][c] = v.x.v; B.e[1][r][c] = v.y.v; B.e[2][r][c] = v.z.v; B.e[3][r][c] = v.w.v; }
}
__device__ __forceinline__ BoxRec box_load_(const BoxTile& B, bool odd, int r, int c) {
    if (odd) return {P2{B.o[0][r][c]}, P2{B.o[1][r][c]}, P2{B.o[2][r][c]}, P2{B.o[3][r][c]}};
    return {P2{B.e[0][r][c]}, P2{B.e[1][r][c]}, P2{B.e[2][r][c]}, P2{B.e[3][r][c]}};
}

#ifndef ASH_DOWN_BOX_MINB
#define ASH_DOWN_BOX_MINB 4
#endif
template <bool TOL>
__global__ void __launch_bounds__(kBoxThreads, ASH_DOWN_BOX_MINB * 256 / kBoxThreads)
k_downsample_box(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int dw, int dh) {
    __shared__ __align__(16) BoxTile B;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int X0 = blockIdx.x * 64, Y0 = blockIdx.y * kBH;
    const int XA = X0 + tx, XB = XA + 32, Y = Y0 + ty;

    // ---- all global loads up front: 3x3 texels per lane (columns 2X-1 .. 2X+1, rows 2Y-1 .. 2Y+1, CLAMP_TO_EDGE)
    uint2 ra[3][3], rb[3][3];
    // (dw % 32 == 0: a lane past the right edge is a whole tile past it, nothing valid consumes its boxes, so its
    //  16-byte load only has to be in bounds, not clamped texel by texel)
    auto load3 = [&](int X, int row, uint2 (&t)[3]) {
        const uint32_t ro = (uint32_t)min(max(row, 0), sh - 1) * (uint32_t)sw;        // < 2^32 texels: 32-bit offsets
        t[0] = __ldg(src + (ro + (uint32_t)min(max(2 * X - 1, 0), sw - 1)));
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(src + (ro + (uint32_t)min(2 * X, sw - 2))));   // texels 2X, 2X+1 (16-byte aligned)
        t[1] = make_uint2(m.x, m.y); t[2] = make_uint2(m.z, m.w);
    };
#pragma unroll
    for (int r = 0; r < 3; ++r) { load3(XA, 2 * Y - 1 + r, ra[r]); load3(XB, 2 * Y - 1 + r, rb[r]); }
    // the extra boxes: the last row (columns 0..64) and column 64 (the rows above it)
    const bool extra = threadIdx.x < 65 + 2 * kBH;
    const int ec = threadIdx.x < 65 ? (int)threadIdx.x : 64, er = threadIdx.x < 65 ? 2 * kBH : (int)threadIdx.x - 65;
    uint2 xa[2][2], xb[2][2];
    if (extra) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint32_t ro = (uint32_t)min(max(2 * Y0 - 1 + er + j, 0), sh - 1) * (uint32_t)sw;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                xa[j][i] = __ldg(src + (ro + (uint32_t)min(max(2 * X0 - 1 + ec + i, 0), sw - 1)));
                xb[j][i] = __ldg(src + (ro + (uint32_t)min(max(2 * X0 + 63 + ec + i, 0), sw - 1)));
            }
        }
    }

    // ---- phase 1: boxes.  4x box sum = (t[j][i] + t[j][i+1]) + (t[j+1][i] + t[j+1][i+1]), as k_downsample_pair
    BoxRec own[2][2];                        // [bj][bi]: this thread's boxes are also taps (i, j) = (bi, bj) of its own pixel
    {
        V3P hs[3][2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const V3P t0 = box_unpack2_(ra[r][0], rb[r][0]), t1 = box_unpack2_(ra[r][1], rb[r][1]), t2 = box_unpack2_(ra[r][2], rb[r][2]);
            hs[r][0] = add3(t0, t1); hs[r][1] = add3(t1, t2);
        }
#pragma unroll
        for (int bj = 0; bj < 2; ++bj)
#pragma unroll
            for (int bi = 0; bi < 2; ++bi) {
                own[bj][bi] = box_record_<TOL>(add3(hs[bj][bi], hs[bj + 1][bi]));
                box_store_(B, bi == 1, 2 * ty + bj, tx, own[bj][bi]);
            }
        if (extra) {
            const V3P h0 = add3(box_unpack2_(xa[0][0], xb[0][0]), box_unpack2_(xa[0][1], xb[0][1]));
            const V3P h1 = add3(box_unpack2_(xa[1][0], xb[1][0]), box_unpack2_(xa[1][1], xb[1][1]));
            box_store_(B, (ec & 1) != 0, er, ec >> 1, box_record_<TOL>(add3(h0, h1)));
        }
    }
    __syncthreads();

    // ---- phase 2: nine taps per pixel in the shader's order, :25-34, :103-107
    if (Y >= dh) return;                     // (XA < dw always: dw % 32 == 0)
    V3P res = {splat(0.0f), splat(0.0f), splat(0.0f)};
    P2 sum = splat(0.0f);
    auto tap = [&](int i, int j, float wt) {
        const BoxRec v = (i < 2 && j < 2) ? own[j][i] : box_load_(B, i == 1, 2 * ty + j, tx + (i >> 1));
        const P2 k = splat(0.25f * wt);                                          // power of two
        res = {ffma(v.x, k, res.x), ffma(v.y, k, res.y), ffma(v.z, k, res.z)};
        sum = ffma(v.w, splat(wt), sum);
    };
    tap(1, 1, 0.25f);                                                        // centre
    tap(0, 0, 0.0625f); tap(2, 0, 0.0625f); tap(0, 2, 0.0625f); tap(2, 2, 0.0625f);   // a b c d   :25-28
    tap(0, 1, 0.125f);  tap(2, 1, 0.125f);  tap(1, 0, 0.125f);  tap(1, 2, 0.125f);    // e f g h   :31-34
    // sum in [1.5e-5, 1], res in [0, 65504]: safe domain                                          :107
    V3P o;
    if (TOL) {
        const P2 r = pk(box_rcp_(lo(sum)), box_rcp_(hi(sum)));
        o = {fmul(res.x, r), fmul(res.y, r), fmul(res.z, r)};
    } else {
        o = {fdiv_safe(res.x, sum), fdiv_safe(res.y, sum), fdiv_safe(res.z, sum)};
    }
    dst[(size_t)Y * dw + XA] = pack_rgba16f(lo(o.x), lo(o.y), lo(o.z), 1.0f);
    if (XB < dw) dst[(size_t)Y * dw + XB] = pack_rgba16f(hi(o.x), hi(o.y), hi(o.z), 1.0f);
}

// A persistent cp.async double-buffered variant of this kernel (130 x 18 source texels of the next tile pair
// staged while the current one is computed, as k_upsample_pair does) was measured SLOWER: 35.2 vs 28.4 us at
// 4K -> 1080p under ncu.  Its long-scoreboard stalls vanish (3.5 -> 0.3) but two CTA barriers per tile pair
// (2.0 stalled warps per issue), the copy address arithmetic (+17 % instructions), 2-way bank conflicts of the
// 8-byte window reads and 3 instead of 4 CTAs per SM (73 KB of shared memory) cost more than the latency hidden.
// Box records as two 128-bit halves ({x, y}, {z, w}: 18 instead of 36 shared-memory instructions per thread) were measured
// equal to slower: 29.8-30.1 vs 29.3-29.5 us.
cudaError_t launch_bloom_downsample_box(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                        uint32_t dh, const Launch& L) {
    const dim3 grid((dw + 63) / 64, (dh + kBH - 1) / kBH);
    Timed timed(L, bloom_class(ASH_KERNEL_BLOOM_DOWNSAMPLE, dw, dh));
    if (L.fast) k_downsample_box<true><<<grid, kBoxThreads, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh,
                                                                       static_cast<uint2*>(dst), (int)dw, (int)dh);
    else k_downsample_box<false><<<grid, kBoxThreads, 0, L.s>>>(static_cast<const uint2*>(src), (int)sw, (int)sh,
                                                                static_cast<uint2*>(dst), (int)dw, (int)dh);
    return cudaGetLastError();
}

}  // namespace ash
