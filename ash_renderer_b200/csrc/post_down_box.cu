// post_down_box.cu — the 2:1 Karis bloom downsample of post_down_box.cuh as a one-shot kernel: one CTA per tile pair.
#include "post_down_box.cuh"

namespace ash {

#ifndef ASH_DOWN_BOX_MINB
#define ASH_DOWN_BOX_MINB 4
#endif
template <bool TOL>
__global__ void __launch_bounds__(kBoxThreads, ASH_DOWN_BOX_MINB * 256 / kBoxThreads)
k_downsample_box(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int dw, int dh) {
    __shared__ __align__(16) BoxTile B;
    down_box_tile<TOL, true>(src, sw, sh, dst, dw, dh, blockIdx.x * 64, blockIdx.y * kBH, B);
}

// A persistent cp.async double-buffered variant of this kernel (130 x 18 source texels of the next tile pair
// staged while the current one is computed, as k_upsample_pair does) was measured SLOWER: 35.2 vs 28.4 us at
// 4K -> 1080p under ncu.  Its long-scoreboard stalls vanish (3.5 -> 0.3) but two CTA barriers per tile pair
// (2.0 stalled warps per issue), the copy address arithmetic (+17 % instructions), 2-way bank conflicts of the
// 8-byte window reads and 3 instead of 4 CTAs per SM (73 KB of shared memory) cost more than the latency hidden.
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
