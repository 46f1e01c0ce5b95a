// post_tail.cu — the SMALL levels of the bloom chain in ONE launch (bloom_downsample.frag:16-108, bloom_upsample.frag:16-41,
// PIN-9): the downsamples into every level below 2^20 texels and the upsamples back up to the first of those levels.
//
// At 4K with five mips that is D_1 -> D_2 -> D_3 -> D_4 -> D_5, then U_4, U_3, U_2: seven launches of 6-9 us each, every one
// of them latency bound (a level is 0.07 - 4 MB), i.e. a quarter of the post chain's critical path for 3 % of its bytes.
// Here they are seven STAGES of one persistent kernel.  The level kernels themselves are unchanged: a stage runs the tile
// function of the kernel the staged chain would launch for that pair of extents (post_down_box.cuh, post_fast.cuh,
// post_math.cuh), so every stored texel is the one the separate launches store — bit for bit in the exact mode, the same
// tolerance kernels in the fast mode.
//
// Scheduling.  Tiles are numbered stage by stage and handed out through one atomic ticket counter; a CTA that draws a
// tile of stage s first waits until done[s - 1] has reached the tile count of stage s - 1.  Tickets are drawn in order and
// only by CTAs that are running, so every tile a CTA can wait for is finished or in the hands of a resident CTA, which
// itself only waits for lower tickets: no deadlock, whatever share of the grid is resident (the kernel usually runs
// beside other views' kernels) — no cooperative launch, no co-residency assumption.  Release / acquire: the CTA's stores,
// bar.sync, then thread 0's __threadfence + atomicAdd(done[s]); the waiting CTA's thread 0 polls done[s - 1] with
// ld.acquire.gpu, then bar.sync.  Texels written earlier in the same launch are read with plain coherent loads (NC = false),
// never through the non-coherent path.  The last CTA to leave resets the counters for the next launch (same stream: ordered).
#include "post_down_box.cuh"
#include "post_fast.cuh"

namespace ash {

enum : int { kTailDownBox = 0, kTailDownFast = 1, kTailDownAny = 2, kTailUpFast = 3, kTailUpAny = 4 };
constexpr int kTailMaxStages = 2 * ASH_MAX_BLOOM_MIPS;
static_assert(kTailSyncWords >= 2 + kTailMaxStages, "ash_internal.cuh: kTailSyncWords");

struct TailStage {
    const uint2* src; uint2* dst;
    int sw, sh, dw, dh;
    int kind, tiles_x, tiles, first;       // first = ticket of the stage's tile 0
};
struct TailArgs {
    TailStage st[kTailMaxStages];
    int n_stages, total;
    float intensity;
};

#ifndef ASH_TAIL_TEXELS        // levels below this many texels are stages of the tail launch
#define ASH_TAIL_TEXELS (1u << 18)
#endif
#ifndef ASH_TAIL_CTAS          // persistent CTAs per SM, at most
#define ASH_TAIL_CTAS 2
#endif
#ifndef ASH_TAIL_NAP
#define ASH_TAIL_NAP 32
#endif
// sync[0] = ticket, sync[1] = CTAs that have left, sync[2 + s] = finished tiles of stage s
__device__ __forceinline__ uint32_t ld_acquire_(const uint32_t* p) {
    uint32_t v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}

template <bool TOL>
__global__ void __launch_bounds__(256, 3)
k_bloom_tail(const __grid_constant__ TailArgs a, uint32_t* __restrict__ sync) {
    __shared__ __align__(16) union { BoxTile B; uint2 s_t[20][20]; } sm;
    __shared__ int s_ticket;
    int next = 0;                                         // thread 0: the ticket drawn one tile ahead (its latency hides behind the tile)
    if (threadIdx.x == 0) next = (int)atomicAdd(&sync[0], 1u);
    for (;;) {
        if (threadIdx.x == 0) s_ticket = next;
        __syncthreads();
        const int t = s_ticket;
        if (t >= a.total) break;
        if (threadIdx.x == 0) next = (int)atomicAdd(&sync[0], 1u);
        int s = 0;
        while (s + 1 < a.n_stages && t >= a.st[s + 1].first) ++s;
        const TailStage& S = a.st[s];
        if (s > 0) {
            if (threadIdx.x == 0) {
                const uint32_t need = (uint32_t)a.st[s - 1].tiles;
                while (ld_acquire_(&sync[2 + s - 1]) < need) __nanosleep(ASH_TAIL_NAP);
            }
            __syncthreads();
        }
        const int tile = t - S.first, ty = tile / S.tiles_x, tx = tile - ty * S.tiles_x;
        const int px = tx * 32 + (int)(threadIdx.x & 31), py = ty * 8 + (int)(threadIdx.x >> 5);
        switch (S.kind) {
            case kTailDownBox:
                down_box_tile<TOL, false>(S.src, S.sw, S.sh, S.dst, S.dw, S.dh, tx * 64, ty * kBH, sm.B);
                break;
            case kTailDownFast:
                if (px < S.dw && py < S.dh) S.dst[(size_t)py * S.dw + px] = down_fast_texel<false>(S.src, S.sw, S.sh, px, py);
                break;
            case kTailDownAny:
                if (px < S.dw && py < S.dh) S.dst[(size_t)py * S.dw + px] = bloom_downsample_texel<false>(Tex16{S.src, S.sw, S.sh}, px, py, S.dw, S.dh);
                break;
            case kTailUpFast:
                up_fast_tile<false, true, false>(S.src, S.sw, S.sh, S.dst, a.intensity, nullptr, nullptr, AshPostProcessPushConstants{},
                                                 tx * 16, ty * 16, sm.s_t);
                break;
            default:
                if (px < S.dw && py < S.dh)
                    S.dst[(size_t)py * S.dw + px] = bloom_upsample_add_texel<false>(Tex16{S.src, S.sw, S.sh}, S.dst[(size_t)py * S.dw + px],
                                                                                   px, py, S.dw, S.dh, a.intensity);
                break;
        }
        __syncthreads();                                  // the tile's stores are issued; sm and s_ticket may be reused
        if (threadIdx.x == 0) { __threadfence(); atomicAdd(&sync[2 + s], 1u); }
    }
    if (threadIdx.x == 0) {
        __threadfence();                                  // this CTA's done[] increments are performed before it is counted out
        if (atomicAdd(&sync[1], 1u) == gridDim.x - 1) {   // every CTA has drawn its last ticket: nobody reads the counters any more
            for (int s = 0; s < kTailSyncWords; ++s) sync[s] = 0u;
            __threadfence();
        }
    }
}

// which kernel the staged chain launches for a level (launch_bloom_downsample_fast / launch_bloom_upsample_fast)
static int down_kind(const BloomChain& b, uint32_t k) {
    if (!b.down_fast[k]) return kTailDownAny;
    return (b.w[k] >= 64u && b.w[k] % 32u == 0u && kBH == 8) ? kTailDownBox : kTailDownFast;
}

// The tail launch makes D_k for k in [first_down, mips] and U_k for k in [first_down, mips - 1], where first_down is the
// first level below 2^20 texels (its callers launch D_1 .. D_{first_down - 1} before and U_{first_down - 1} .. after it);
// 0 = no tail: it would replace fewer than three launches.
uint32_t bloom_tail_first_level(const BloomChain& b, uint32_t mips) {
    uint32_t k = 1;
    while (k <= mips && (uint64_t)b.w[k] * b.h[k] >= ASH_TAIL_TEXELS) ++k;
    if (k > mips || 2 * (mips - k) + 1 < 3) return 0;
    return k;
}

cudaError_t launch_bloom_tail(const BloomChain& b, uint32_t mips, uint32_t first_down, float intensity, uint32_t* sync, const Launch& L) {
    TailArgs a{};
    int n = 0, ticket = 0;
    auto push = [&](uint32_t from, uint32_t to, int kind) {
        TailStage& S = a.st[n++];
        S.src = static_cast<const uint2*>(b.d[from]); S.dst = static_cast<uint2*>(b.d[to]);
        S.sw = (int)b.w[from]; S.sh = (int)b.h[from]; S.dw = (int)b.w[to]; S.dh = (int)b.h[to];
        S.kind = kind;
        int tw = 32, th = 8;                                                  // destination pixels per tile
        if (kind == kTailDownBox) { tw = 64; th = kBH; }
        if (kind == kTailUpFast) { tw = 32; th = 32; }
        S.tiles_x = (S.dw + tw - 1) / tw;
        S.tiles = S.tiles_x * ((S.dh + th - 1) / th);
        S.first = ticket; ticket += S.tiles;
    };
    for (uint32_t k = first_down; k <= mips; ++k) push(k - 1, k, down_kind(b, k));
    for (int k = (int)mips - 1; k >= (int)first_down; --k) push((uint32_t)k + 1, (uint32_t)k, b.up_fast[k] ? kTailUpFast : kTailUpAny);
    a.n_stages = n; a.total = ticket; a.intensity = intensity;
    const int grid = ticket < 148 * ASH_TAIL_CTAS ? ticket : 148 * ASH_TAIL_CTAS;
    Timed timed(L, ASH_KERNEL_BLOOM_TAIL);
    if (L.fast) k_bloom_tail<true><<<grid, 256, 0, L.s>>>(a, sync);
    else k_bloom_tail<false><<<grid, 256, 0, L.s>>>(a, sync);
    return cudaGetLastError();
}

}  // namespace ash
