// depth_pyramid.cu — single-pass Hi-Z pyramid + 16x16 tile depth min/max.
//
// Replaces the reference's one-dispatch-per-mip shaders/hiz_generate.comp:23-41 (each mip re-read
// from memory) and the depth min/max phase of shaders/light_culling.comp:122-143 (a second full
// read of the depth buffer): depth is read from HBM exactly once.
//
// One CTA owns a 64x64 depth tile (16 KB): 8 warps, each a 32x16 pixel patch, each thread a 4x4
// pixel block loaded as four 128-bit streaming loads (full 128 B lines per warp row).  Mips 1-2
// reduce in registers, 3-4 by warp shuffles, 5-6 through 64 B of shared memory.  The 16x16 light
// tile coincides with the mip-4 footprint, so its min/max (on float bits, like the shader's
// atomicMin/atomicMax on floatBitsToUint) falls out of the same shuffles.  The remaining tiny levels
// (mip 7 and up, <= 60x33 texels at 4K) are reduced by k_pyramid_tail, one CTA ping-ponging through
// shared memory, launched right behind on the same stream.
//
// Algorithmic bytes per launch: 4*W*H (depth) + 4*sum_{L>=1} W_L*H_L (mips) + 8*tiles.
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_stream_f1(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
// hiz_generate.comp:34-38 — gather order x=(i0,j1) y=(i1,j1) z=(i1,j0) w=(i0,j0)
__device__ __forceinline__ float hiz_max4(float i0j0, float i1j0, float i0j1, float i1j1) {
    return gmax(gmax(i0j1, i1j1), gmax(i1j0, i0j0));
}

template <bool VEC>
__global__ void __launch_bounds__(256)
k_depth_pyramid(Pyramid p, uint32_t* __restrict__ tile_minmax, uint32_t tiles_x,
                uint32_t* __restrict__ depth_range, uint32_t* __restrict__ ticket) {
    __shared__ float    s4[4][4];
    __shared__ uint32_t s_mn[8], s_mx[8];

    const uint32_t W = p.w[0], H = p.h[0];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tx = lane & 7, ty = lane >> 3;
    const uint32_t X = blockIdx.x * 64u + (warp & 1) * 32u + tx * 4u;
    const uint32_t Y = blockIdx.y * 64u + (warp >> 1) * 16u + ty * 4u;
    const float* __restrict__ depth = p.mip0;

    float v[4][4];
    uint32_t mn = 0xFFFFFFFFu, mx = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t y = Y + r;
        if (VEC) {
            if (y < H && X < W) {
                float4 q = ld_stream_f4(depth + (size_t)y * W + X);
                v[r][0] = q.x; v[r][1] = q.y; v[r][2] = q.z; v[r][3] = q.w;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t b = __float_as_uint(v[r][c]);
                    mn = min(mn, b); mx = max(mx, b);
                }
            } else {
                v[r][0] = v[r][1] = v[r][2] = v[r][3] = 0.0f;
            }
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (y < H && X + c < W) {
                    v[r][c] = ld_stream_f1(depth + (size_t)y * W + X + c);
                    uint32_t b = __float_as_uint(v[r][c]);
                    mn = min(mn, b); mx = max(mx, b);
                } else {
                    v[r][c] = 0.0f;
                }
            }
        }
    }

    const int F = p.fast_levels;
    // ---- mip 1 (2x2 per thread) and mip 2 (1 per thread), in registers
    float m1[2][2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 2; ++i)
            m1[j][i] = hiz_max4(v[2 * j][2 * i], v[2 * j][2 * i + 1], v[2 * j + 1][2 * i], v[2 * j + 1][2 * i + 1]);
    if (F >= 1) {
        float* d = p.mips + p.off[1];
        const uint32_t w1 = p.w[1], h1 = p.h[1], x1 = X >> 1, y1 = Y >> 1;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (y1 + j < h1) {
                if (VEC) {   // W%4==0 -> w1 even, x1 even, level base 16 B aligned
                    if (x1 + 1 < w1) *reinterpret_cast<float2*>(d + (size_t)(y1 + j) * w1 + x1) = make_float2(m1[j][0], m1[j][1]);
                } else {
                    if (x1 < w1) d[(size_t)(y1 + j) * w1 + x1] = m1[j][0];
                    if (x1 + 1 < w1) d[(size_t)(y1 + j) * w1 + x1 + 1] = m1[j][1];
                }
            }
        }
    }
    const float m2 = hiz_max4(m1[0][0], m1[0][1], m1[1][0], m1[1][1]);
    if (F >= 2) {
        const uint32_t x2 = X >> 2, y2 = Y >> 2;
        if (x2 < p.w[2] && y2 < p.h[2]) p.mips[p.off[2] + (size_t)y2 * p.w[2] + x2] = m2;
    }
    // ---- mip 3: 2x2 threads (lane^1 = +x, lane^8 = +y)
    const float m2x = __shfl_xor_sync(0xFFFFFFFFu, m2, 1);
    const float m2y = __shfl_xor_sync(0xFFFFFFFFu, m2, 8);
    const float m2xy = __shfl_xor_sync(0xFFFFFFFFu, m2, 9);
    const float m3 = hiz_max4(m2, m2x, m2y, m2xy);          // meaningful in (tx,ty) even/even lanes
    if (F >= 3 && !(tx & 1) && !(ty & 1)) {
        const uint32_t x3 = X >> 3, y3 = Y >> 3;
        if (x3 < p.w[3] && y3 < p.h[3]) p.mips[p.off[3] + (size_t)y3 * p.w[3] + x3] = m3;
    }
    // ---- mip 4: 4x4 threads (lane^2 = +x, lane^16 = +y) == one 16x16 light tile
    const float m3x = __shfl_xor_sync(0xFFFFFFFFu, m3, 2);
    const float m3y = __shfl_xor_sync(0xFFFFFFFFu, m3, 16);
    const float m3xy = __shfl_xor_sync(0xFFFFFFFFu, m3, 18);
    const float m4 = hiz_max4(m3, m3x, m3y, m3xy);          // meaningful in lanes 0 and 4
    uint32_t tmn = mn, tmx = mx;
#pragma unroll
    for (int sft : {1, 2, 8, 16}) {
        tmn = min(tmn, __shfl_xor_sync(0xFFFFFFFFu, tmn, sft));
        tmx = max(tmx, __shfl_xor_sync(0xFFFFFFFFu, tmx, sft));
    }
    if ((lane & 27) == 0) {   // lanes 0 and 4
        const uint32_t x4 = X >> 4, y4 = Y >> 4;
        if (F >= 4 && x4 < p.w[4] && y4 < p.h[4]) p.mips[p.off[4] + (size_t)y4 * p.w[4] + x4] = m4;
        s4[warp >> 1][(warp & 1) * 2 + (tx >> 2)] = m4;
        if (X < W && Y < H)   // light_culling.comp:117-119 tile indexing
            *reinterpret_cast<uint2*>(tile_minmax + 2 * ((size_t)y4 * tiles_x + x4)) = make_uint2(tmn, tmx);
    }
    const uint32_t wmn = min(tmn, __shfl_xor_sync(0xFFFFFFFFu, tmn, 4));
    const uint32_t wmx = max(tmx, __shfl_xor_sync(0xFFFFFFFFu, tmx, 4));
    if (lane == 0) { s_mn[warp] = wmn; s_mx[warp] = wmx; }
    __syncthreads();

    // ---- mips 5 and 6 from the 4x4 mip-4 texels of this tile
    if (warp == 0) {
        const int i = lane & 1, j = (lane >> 1) & 1;
        const float m5 = hiz_max4(s4[2 * j][2 * i], s4[2 * j][2 * i + 1], s4[2 * j + 1][2 * i], s4[2 * j + 1][2 * i + 1]);
        if (F >= 5 && lane < 4) {
            const uint32_t x5 = blockIdx.x * 2u + i, y5 = blockIdx.y * 2u + j;
            if (x5 < p.w[5] && y5 < p.h[5]) p.mips[p.off[5] + (size_t)y5 * p.w[5] + x5] = m5;
        }
        const float a = __shfl_sync(0xFFFFFFFFu, m5, 0), b = __shfl_sync(0xFFFFFFFFu, m5, 1);
        const float c = __shfl_sync(0xFFFFFFFFu, m5, 2), d = __shfl_sync(0xFFFFFFFFu, m5, 3);
        if (F >= 6 && lane == 0 && blockIdx.x < p.w[6] && blockIdx.y < p.h[6])
            p.mips[p.off[6] + (size_t)blockIdx.y * p.w[6] + blockIdx.x] = hiz_max4(a, b, c, d);
        if (lane == 0) {
            uint32_t cmn = s_mn[0], cmx = s_mx[0];
#pragma unroll
            for (int k = 1; k < 8; ++k) { cmn = min(cmn, s_mn[k]); cmx = max(cmx, s_mx[k]); }
            atomicMin(depth_range, cmn);
            atomicMin(depth_range + 1, ~cmx);   // stored inverted: both words reset to 0xFFFFFFFF
        }
    }

}

// Levels fast_levels+1 .. levels-1 (a few thousand texels in total): one CTA; the first of them is
// gathered from L2 into shared memory and the rest never leave it, except for their stores.
constexpr int kTailCap = 4096;     // texels of the first tail level that fit the shared-memory path
__global__ void __launch_bounds__(1024)
k_pyramid_tail(Pyramid p) {
    __shared__ float s_a[kTailCap], s_b[kTailCap / 2];     // a level with one extent clamped to 1 only HALVES (4096x1 -> 2048x1)
    const int F = p.fast_levels;
    const float* src_g = (F == 0) ? p.mip0 : p.mips + p.off[F];
    int sw = (int)p.w[F], sh = (int)p.h[F];
    const bool in_smem = sw * sh <= kTailCap;
    float* cur = s_a; float* nxt = s_b;
    if (in_smem) {
        for (int idx = threadIdx.x; idx < sw * sh; idx += blockDim.x) cur[idx] = __ldcg(src_g + idx);
        __syncthreads();
    }
    for (int L = F + 1; L < p.levels; ++L) {
        float* dst = p.mips + p.off[L];
        const int dw = (int)p.w[L], dh = (int)p.h[L];
        for (int idx = threadIdx.x; idx < dw * dh; idx += blockDim.x) {
            const int x = idx % dw, y = idx / dw;
            const int i0 = min(2 * x, sw - 1), i1 = min(2 * x + 1, sw - 1);      // PIN-5 edge clamp
            const int j0 = min(2 * y, sh - 1), j1 = min(2 * y + 1, sh - 1);
            float v;
            if (in_smem) v = hiz_max4(cur[j0 * sw + i0], cur[j0 * sw + i1], cur[j1 * sw + i0], cur[j1 * sw + i1]);
            else v = hiz_max4(__ldcg(src_g + (size_t)j0 * sw + i0), __ldcg(src_g + (size_t)j0 * sw + i1),
                              __ldcg(src_g + (size_t)j1 * sw + i0), __ldcg(src_g + (size_t)j1 * sw + i1));
            dst[idx] = v;
            if (in_smem) nxt[idx] = v;
        }
        __syncthreads();
        if (in_smem) { float* t = cur; cur = nxt; nxt = t; }
        else src_g = dst;
        sw = dw; sh = dh;
    }
}

// Running these levels in the LAST CTA of k_depth_pyramid to finish (a ticket behind a __threadfence) instead of a second
// launch was measured: one node less per view, 0.205 -> 0.202 ms of single-view latency, and -1.7 % views/s with other
// views' kernels beside it — each of the 2040 CTAs holds its slot until its fence has drained.  Two launches.
cudaError_t launch_depth_pyramid(const Pyramid& p, uint32_t* tile_minmax, uint32_t tiles_x,
                                 uint32_t* depth_range, uint32_t* ticket, const Launch& L) {
    dim3 grid((p.w[0] + 63u) / 64u, (p.h[0] + 63u) / 64u);
    const bool vec = (p.w[0] % 4u == 0u) && ((reinterpret_cast<uintptr_t>(p.mip0) & 15u) == 0u);
    Timed timed(L, ASH_KERNEL_DEPTH_PYRAMID);
    if (vec) k_depth_pyramid<true><<<grid, 256, 0, L.s>>>(p, tile_minmax, tiles_x, depth_range, ticket);
    else     k_depth_pyramid<false><<<grid, 256, 0, L.s>>>(p, tile_minmax, tiles_x, depth_range, ticket);
    if (p.levels - 1 > p.fast_levels) {
        ++*L.launches;                       // accounted to the same class as the main pass
        k_pyramid_tail<<<1, 1024, 0, L.s>>>(p);
    }
    return cudaGetLastError();
}

}  // namespace ash
