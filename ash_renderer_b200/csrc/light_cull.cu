// light_cull.cu — 16x16 tiled light culling into per-tile index lists.
//
// Replaces shaders/light_culling.comp:111-203.  The shader re-reads depth per tile, transforms
// every light to view space once per tile (tiles x lights mat-vecs) and appends through a shared
// atomic in arbitrary order.  Here:
//   k_light_prepass (1 CTA per 1024 lights, ordered across CTAs): view-space transform once per light
//     per view (:181 is tile independent) and an exact global prefilter: a light that fails the
//     near/far planes (:96-97) evaluated with the frame's global [min,max] depth fails them in every
//     tile (rounding is monotone and tile min >= global min, tile max <= global max), so only
//     survivors ("candidates") are kept, in ascending light index.
//   k_tile_planes (1 thread per tile): the four side planes of every tile frustum (:71-98), once.
//   k_cell_planes (cached with the tile planes): per 8x8-tile cell, what a conservative prefilter
//     needs from the planes alone.
//   k_light_cell_tiles (1 CTA per cell): the prefilter — candidates that provably fail the shader's
//     test in every tile of the cell are dropped — then, over the survivors held in shared memory,
//     the shader's exact predicate (:101-109) per tile; tile min/max come from the depth pyramid
//     kernel (depth is not read again); lists are appended by ballot + popc prefix, which yields
//     ascending order and first-256-on-overflow (PIN-7) without atomics.
//
// Algorithmic bytes per launch group: 64*L (lights) + 208 (camera) + 8*tiles (tile min/max)
// + 4*sum_tiles(1+count).
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

constexpr int kPrepassThreads = 1024;

// One candidate = 32 bytes: {view x, y, z, radius} {dot((0,0,-1),c), dot((0,0,1),c), light index bits, 0}.
// The two z dots are the tile-independent parts of planes 4/5 (:96-97, :104), evaluated exactly as the
// shader does (so non-finite light positions behave identically) once per light instead of once per tile.
struct CandZ { float zneg, zpos; uint32_t index, pad; };

// Ordered compaction across CTAs.  CTA b (b = the order in which CTAs START, drawn from a ticket, so that every CTA a
// running CTA waits for is itself running) owns lights [1024 b, 1024 b + 1024): one coalesced load per thread, the global
// z-slab test, a block scan of the 32 warp ballots, then it PUBLISHES its survivor count (bit 31 = valid) and sums the
// published counts of CTAs 0 .. b - 1 (threads poll one predecessor each) for its base slot: survivors land in
// ascending light index.  The single-CTA form of this kernel ran 25 k warp-instructions on one SM: 11.8 us under ncu for
// 8192 lights, the longest link of the front chain.  sync: [0] start ticket, [1] CTAs that have left, [2 + b] counts.
__global__ void __launch_bounds__(kPrepassThreads)
k_light_prepass(const AshGpuLight* __restrict__ lights, uint32_t light_count,
                const AshCullingCameraData* __restrict__ cam, uint32_t* __restrict__ depth_range,
                LightCand* __restrict__ cand, uint32_t* __restrict__ cand_idx,
                uint32_t* __restrict__ cand_count, AshViewStats* __restrict__ stats, uint32_t tiles,
                AshCullingStats* __restrict__ cstats, uint32_t* __restrict__ sync) {
    __shared__ uint32_t s_cnt[kPrepassThreads / 32];     // per warp survivors, then exclusive offsets
    __shared__ uint32_t s_red[kPrepassThreads / 32];
    __shared__ uint32_t s_bid, s_total, s_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) s_bid = atomicAdd(&sync[0], 1u);
    __syncthreads();
    const uint32_t bid = s_bid, nblk = gridDim.x;
    const M4 view = load_m4(cam->view);
    const uint32_t gmn = depth_range[0], gmx = ~depth_range[1];
    // uint order == float order only for non-negative, non-NaN depth (PIN-2); otherwise keep all
    const bool sane = gmx <= 0x7F800000u && gmn <= gmx;
    const float gmin_d = __uint_as_float(gmn), gmax_d = __uint_as_float(gmx);
    CandZ* candz = reinterpret_cast<CandZ*>(cand_idx);

    const uint32_t i = bid * kPrepassThreads + threadIdx.x;
    float4 q = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float enabled = 0.0f;
    if (i < light_count) {
        q = __ldg(reinterpret_cast<const float4*>(lights[i].position));
        enabled = __ldg(&lights[i].params[3]);
    }
    bool keep = i < light_count && !(enabled < 0.5f);                                      // :178
    const V4 pv = mul4(view, q.x, q.y, q.z, 1.0f);                                         // :181
    if (keep && sane) {
        // planes 4/5 with the global slab: a light failing them here fails them in every tile
        const float d4 = fadd(dot3(v3(0.0f, 0.0f, -1.0f), v3(pv.x, pv.y, pv.z)), -gmin_d);
        const float d5 = fadd(dot3(v3(0.0f, 0.0f, 1.0f), v3(pv.x, pv.y, pv.z)), gmax_d);
        if (d4 < -q.w || d5 < -q.w) keep = false;
    }
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) s_cnt[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {                                       // exclusive scan of the 32 warp counts; publish the CTA's total
        const uint32_t v = s_cnt[lane];
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        s_cnt[lane] = incl - v;
        if (lane == 31) { s_total = incl; atomicExch(&sync[2 + bid], incl | 0x80000000u); }
    }
    // base slot = survivors of CTAs 0 .. bid - 1 (they started earlier and publish without waiting for anybody)
    uint32_t part = 0;
    for (uint32_t b = threadIdx.x; b < bid; b += kPrepassThreads) {
        uint32_t v;
        do { v = *reinterpret_cast<volatile uint32_t*>(&sync[2 + b]); } while (!(v & 0x80000000u));
        part += v & 0x7FFFFFFFu;
    }
#pragma unroll
    for (int sft = 16; sft > 0; sft >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, sft);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (warp == 0) {
        uint32_t v = s_red[lane];
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, sft);
        if (lane == 0) s_base = v;
    }
    __syncthreads();
    if (keep) {
        const uint32_t slot = s_base + s_cnt[warp] + __popc(ballot & ((1u << lane) - 1u));
        *reinterpret_cast<float4*>(cand + slot) = make_float4(pv.x, pv.y, pv.z, q.w);
        CandZ cz;
        cz.zneg = dot3(v3(0.0f, 0.0f, -1.0f), v3(pv.x, pv.y, pv.z));
        cz.zpos = dot3(v3(0.0f, 0.0f, 1.0f), v3(pv.x, pv.y, pv.z));
        cz.index = i; cz.pad = 0;
        *reinterpret_cast<uint4*>(candz + slot) = *reinterpret_cast<const uint4*>(&cz);
    }
    if (threadIdx.x == 0) {
        if (bid == nblk - 1u) {                            // the CTA of the last lights knows the total
            const uint32_t total = s_base + s_total;
            *cand_count = total;
            stats->tiles = tiles;
            stats->candidate_lights = total;
            stats->sum_tile_lights = 0;
            stats->max_tile_lights = 0;
            stats->overflow_tiles = 0;
            // CullingStats::lights_culled (view_run: what k_culling_stats_finish would store at the end of the front chain)
            if (cstats) cstats->lights_culled = light_count - min(light_count, total);
        }
        __threadfence();
        if (atomicAdd(&sync[1], 1u) == nblk - 1u) {        // last CTA out: every CTA has read depth_range and every count it needs
            depth_range[0] = 0xFFFFFFFFu;                  // consumed: reset for the next frame's depth kernel
            depth_range[1] = 0xFFFFFFFFu;
            sync[0] = 0u; sync[1] = 0u;
            for (uint32_t b = 0; b < nblk; ++b) sync[2 + b] = 0u;
        }
    }
}

// screenToView(uv, 0.0), :64-68
__device__ __forceinline__ V3 screen_to_view0(const M4& ip, float u, float v) {
    const V4 q = mul4(ip, fsub(fmul(u, 2.0f), 1.0f), fsub(fmul(v, 2.0f), 1.0f), 0.0f, 1.0f);
    return {fdiv(q.x, q.w), fdiv(q.y, q.w), fdiv(q.z, q.w)};
}

// getTileFrustum, :71-98: the four side planes of tile (tx,ty): normals n[0..3] (left, right, bottom, top;
// zero vector when |cross| <= 1e-4, :92-95) and the cross-product lengths (for the cell prefilter).
__device__ __forceinline__ void tile_side_planes(const M4& ip, uint32_t tx, uint32_t ty, float sw, float sh,
                                                 V3 n[4], float len[4]) {
    const float ts = 16.0f;
    const float min_x = fsub(fmul(fdiv(fmul((float)tx, ts), sw), 2.0f), 1.0f);
    const float min_y = fsub(fmul(fdiv(fmul((float)ty, ts), sh), 2.0f), 1.0f);
    const float max_x = fsub(fmul(fdiv(fmul((float)(tx + 1u), ts), sw), 2.0f), 1.0f);
    const float max_y = fsub(fmul(fdiv(fmul((float)(ty + 1u), ts), sh), 2.0f), 1.0f);
    const float u0 = fadd(fmul(min_x, 0.5f), 0.5f), u1 = fadd(fmul(max_x, 0.5f), 0.5f);
    const float v0 = fadd(fmul(min_y, 0.5f), 0.5f), v1 = fadd(fmul(max_y, 0.5f), 0.5f);
    const V3 c0 = screen_to_view0(ip, u0, v0), c1 = screen_to_view0(ip, u1, v0);
    const V3 c2 = screen_to_view0(ip, u1, v1), c3 = screen_to_view0(ip, u0, v1);
    const V3 cr[4] = {cross3(c3, c0), cross3(c1, c2), cross3(c0, c1), cross3(c2, c3)};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        len[k] = length3(cr[k]);
        n[k] = (len[k] > 0.0001f) ? divs3(cr[k], len[k]) : v3(0.0f, 0.0f, 0.0f);   // normalize(v) = v / length(v)
    }
}

// ---- pass 2: one thread per tile builds its side planes once (they depend on the camera only)
struct TilePlanes { float4 a, b, c; };      // n0.xyz n1.x | n1.yz n2.xy | n2.z n3.xyz

__global__ void __launch_bounds__(128)
k_tile_planes(const AshCullingCameraData* __restrict__ cam, uint32_t W, uint32_t H, uint32_t tiles_x,
              uint32_t tiles, TilePlanes* __restrict__ planes) {
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles) return;
    const M4 ip = load_m4(cam->inv_projection);
    V3 n[4]; float len[4];
    tile_side_planes(ip, tile % tiles_x, tile / tiles_x, (float)W, (float)H, n, len);
    TilePlanes p;
    p.a = make_float4(n[0].x, n[0].y, n[0].z, n[1].x);
    p.b = make_float4(n[1].y, n[1].z, n[2].x, n[2].y);
    p.c = make_float4(n[2].z, n[3].x, n[3].y, n[3].z);
    planes[tile] = p;
}

// ---- pass 3: conservative prefilter per cell of 8x8 tiles (128x128 px), one CTA per cell.
// A candidate is dropped from a cell only if it provably fails the shader's test in EVERY tile of it:
//  * planes 4/5: fadd(zdot, w) is monotone in w and the cell's [min,max] depth contains every tile's,
//    so failing with the cell's bounds implies failing with each tile's (exact, no margin);
//  * side planes, per family f (left / right / bottom / top): with a, b the family's normals at the two
//    ends of the cell, every tile's ACTUAL normal (read back from k_tile_planes' output, so the bound
//    holds for any matrix) is decomposed as n_t = al*a + be*b + e with al, be >= 0.  Then
//        n_t.c <= (al+be)*max(a.c, b.c) + |e||c|
//    and with s_min = min_t(al+be), E = max_t|e|, a light with max(a.c,b.c) < 0 and
//        s_min*max(a.c,b.c) + E|c| < -r
//    has n_t.c < -r in every tile.  For a perspective camera the family's normals lie on one great
//    circle, so E is only the rounding noise of the shader's cross products (~1e-4) and s_min ~ 1.
//    A tile whose normal the shader zeroed (:92-95) gets al = be = 0, hence s_min = 0 and nothing is
//    rejected by that family; non-finite normals disable the family.
constexpr int kCellTiles = 8;
constexpr int kCellWarps = 4;

__device__ __forceinline__ V3 plane_n(const TilePlanes& p, int f) {
    switch (f) {
        case 0: return {p.a.x, p.a.y, p.a.z};
        case 1: return {p.a.w, p.b.x, p.b.y};
        case 2: return {p.b.z, p.b.w, p.c.x};
        default: return {p.c.y, p.c.z, p.c.w};
    }
}
__device__ __forceinline__ float pdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// What pass 3 needs from the tile side planes alone — per family the two end normals a, b, s_min, E and whether the
// family may reject at all — depends on invProjection and the extent only, like the planes themselves: computed once
// per projection (k_cell_planes, cached with k_tile_planes' output) instead of once per view (it was 2/3 of k_light_cells).
struct CellPlanes { float fa[4][3], fb[4][3], smin[4], emax[4]; uint32_t fvalid, pad[3]; };

__global__ void __launch_bounds__(kCellWarps * 32)
k_cell_planes(uint32_t tiles_x, uint32_t tiles_y, uint32_t cells_x, uint32_t cells,
              const TilePlanes* __restrict__ planes, CellPlanes* __restrict__ out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cell = blockIdx.x * kCellWarps + warp;
    if (cell >= cells) return;
    const uint32_t cx = cell % cells_x, cy = cell / cells_x;
    const uint32_t tx0 = cx * kCellTiles, ty0 = cy * kCellTiles;
    const uint32_t tx1 = min(tx0 + kCellTiles - 1, tiles_x - 1), ty1 = min(ty0 + kCellTiles - 1, tiles_y - 1);
    const TilePlanes p00 = planes[(size_t)ty0 * tiles_x + tx0];
    const TilePlanes p10 = planes[(size_t)ty0 * tiles_x + tx1];
    const TilePlanes p01 = planes[(size_t)ty1 * tiles_x + tx0];
    V3 fa[4], fb[4];
    float smin[4], emax[4];
    bool bad[4];
#pragma unroll
    for (int f = 0; f < 4; ++f) {
        fa[f] = plane_n(p00, f);
        fb[f] = plane_n(f < 2 ? p10 : p01, f);
        smin[f] = 3.0e38f; emax[f] = 0.0f; bad[f] = false;
    }
    for (uint32_t t = lane; t < (uint32_t)(kCellTiles * kCellTiles); t += 32) {
        const uint32_t tx = tx0 + (t % kCellTiles), ty = ty0 + (t / kCellTiles);
        if (tx > tx1 || ty > ty1) continue;
        const TilePlanes p = planes[(size_t)ty * tiles_x + tx];
#pragma unroll
        for (int f = 0; f < 4; ++f) {
            const V3 n = plane_n(p, f), a = fa[f], b = fb[f];
            const float aa = pdot(a, a), bb = pdot(b, b), ab = pdot(a, b), na = pdot(n, a), nb = pdot(n, b);
            const float det = aa * bb - ab * ab;
            float al, be;
            if (det > 1e-6f) { al = (na * bb - nb * ab) / det; be = (nb * aa - na * ab) / det; }
            else             { al = aa > 0.5f ? na / aa : 0.0f; be = 0.0f; }
            al = al > 0.0f ? al : 0.0f;                              // NaN -> 0 (the residual then carries the NaN)
            be = be > 0.0f ? be : 0.0f;
            const V3 e = {n.x - al * a.x - be * b.x, n.y - al * a.y - be * b.y, n.z - al * a.z - be * b.z};
            const float el = sqrtf(pdot(e, e));
            if (!(el < 1.0f) || !(al + be < 4.0f)) bad[f] = true;    // NaN / inf / nonsense
            smin[f] = fminf(smin[f], al + be);
            emax[f] = fmaxf(emax[f], el);
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1)
#pragma unroll
        for (int f = 0; f < 4; ++f) {
            smin[f] = fminf(smin[f], __shfl_xor_sync(0xFFFFFFFFu, smin[f], s));
            emax[f] = fmaxf(emax[f], __shfl_xor_sync(0xFFFFFFFFu, emax[f], s));
        }
    uint32_t valid = 0;
#pragma unroll
    for (int f = 0; f < 4; ++f)
        if (!__any_sync(0xFFFFFFFFu, bad[f]) && smin[f] > 0.0f) valid |= 1u << f;
    if (lane == 0) {
        CellPlanes c;
#pragma unroll
        for (int f = 0; f < 4; ++f) {
            c.fa[f][0] = fa[f].x; c.fa[f][1] = fa[f].y; c.fa[f][2] = fa[f].z;
            c.fb[f][0] = fb[f].x; c.fb[f][1] = fb[f].y; c.fb[f][2] = fb[f].z;
            c.smin[f] = smin[f];
            c.emax[f] = emax[f] + 4e-6f;      // rounding of e itself and of the shader's own dot product
        }
        c.fvalid = valid; c.pad[0] = c.pad[1] = c.pad[2] = 0u;
        out[cell] = c;
    }
}

// ---- passes 3 + 4 in ONE launch, one CTA of 128 threads per cell.
// Pass 3 (prefilter): a round takes 4 x 128 candidates — thread t owns candidates start + it * 128 + t, every load of the
// round is issued up front — and one scan over the 4 x 4 warp ballots gives each survivor its slot in ascending candidate
// order (= ascending light index), as in k_light_prepass.  The survivors' records (view-space sphere, z dots, light
// index) go straight to SHARED memory, not to a per-cell list in global memory.
// Pass 4 (the shader's exact test, :101-109): each warp runs 16 of the cell's 64 tiles over the round's survivors, 32 per
// step, appending by ballot + popc prefix: ascending order and first-256-on-overflow (PIN-7) without atomics.  The
// tiles' side planes and depth bounds are staged once per cell.
// History: one warp per cell + one warp per tile (round 1: 18 + 15 us at 4K, and the tile kernel alone filled the GPU's
// thread slots for 3.4 waves) -> one CTA per cell for each pass -> fused.  What this buys is not latency but footprint:
// with other views' kernels running beside it (bench.py) the light group cost 11 us of GPU time per view
// (tools/front_cost.sh), most of it thread slots held by warps waiting on the list -> candidate -> test chain of
// dependent loads: 6350 -> 6510 views/s.  Survivors are buffered across rounds (kFuseBuf) so that a scene which keeps
// most of its lights as candidates but few per cell runs the exact pass once, not once per round; the cell's prefilter
// planes are re-read from shared memory each round instead of living in 32 registers across the exact pass.
// tools/light_stress.py (7760 candidates at 4K; prefilter + exact pass, event-timed): 228 lights per tile 84 -> 88 us,
// 3 lights per tile 57 -> 62 us — the price of 256 instead of 256 + 512 threads per cell; 128 threads: 95 / 59 us.
#ifndef ASH_FUSE_THREADS
#define ASH_FUSE_THREADS 256
#endif
#ifndef ASH_FUSE_ITERS
#define ASH_FUSE_ITERS 2
#endif
constexpr int kFuseThreads = ASH_FUSE_THREADS, kFuseWarps = kFuseThreads / 32, kFuseIters = ASH_FUSE_ITERS, kFuseRound = kFuseThreads * kFuseIters;
constexpr int kFuseBuf = 1024;                                 // survivor records held in shared memory between two exact passes
constexpr int kFuseTilesPerWarp = kCellTiles * kCellTiles / kFuseWarps;
static_assert(kFuseIters * kFuseWarps <= 32, "one count per lane in the scan");
static_assert(kFuseRound <= kFuseBuf, "a round's survivors must fit the buffer");
#ifndef ASH_FUSE_MINB
#define ASH_FUSE_MINB 4
#endif
__global__ void __launch_bounds__(kFuseThreads, ASH_FUSE_MINB)
k_light_cell_tiles(uint32_t tiles_x, uint32_t tiles_y, uint32_t cells_x, uint32_t light_count,
                   const CellPlanes* __restrict__ cell_planes, const TilePlanes* __restrict__ planes,
                   const uint32_t* __restrict__ tile_minmax, const LightCand* __restrict__ cand,
                   const uint32_t* __restrict__ cand_idx, const uint32_t* __restrict__ cand_count,
                   uint32_t* __restrict__ tile_data, AshViewStats* __restrict__ stats) {
    __shared__ float4 s_q[kFuseBuf];                           // buffered survivors: view-space centre, radius
    __shared__ float2 s_z[kFuseBuf];                           //   CandZ.zneg, .zpos
    __shared__ uint32_t s_li[kFuseBuf];                        //   light index
    __shared__ TilePlanes s_planes[kCellTiles * kCellTiles];
    __shared__ uint2 s_mm[kCellTiles * kCellTiles];            // tile depth bounds as float bits
    __shared__ uint32_t s_in[2];                               // bit t: tile t of the cell lies inside the image
    __shared__ uint32_t s_tcnt[kCellTiles * kCellTiles];       // lights that pass in each tile so far (owned by one warp each)
    __shared__ CellPlanes s_cp;                                // re-read per round: 32 registers less across the exact pass
    __shared__ uint32_t s_cnt[kFuseIters * kFuseWarps], s_total;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cell = blockIdx.x;
    const uint32_t cx = cell % cells_x, cy = cell / cells_x;
    const uint32_t tx0 = cx * kCellTiles, ty0 = cy * kCellTiles;
    if (threadIdx.x < kCellTiles * kCellTiles) {
        const uint32_t tx = tx0 + (threadIdx.x % kCellTiles), ty = ty0 + (threadIdx.x / kCellTiles);
        uint2 mm = make_uint2(0u, 0u);
        const bool inside = tx < tiles_x && ty < tiles_y;
        if (inside && light_count != 0u) {
            const uint32_t tile = ty * tiles_x + tx;
            s_planes[threadIdx.x] = planes[tile];
            mm = __ldg(reinterpret_cast<const uint2*>(tile_minmax) + tile);
        }
        s_mm[threadIdx.x] = mm;
        s_tcnt[threadIdx.x] = 0u;
        const uint32_t in = __ballot_sync(0xFFFFFFFFu, inside);
        if (lane == 0) s_in[warp] = in;
    } else if (light_count != 0u && threadIdx.x - kCellTiles * kCellTiles < sizeof(CellPlanes) / 4) {
        const uint32_t k = threadIdx.x - kCellTiles * kCellTiles;
        reinterpret_cast<uint32_t*>(&s_cp)[k] = __ldg(reinterpret_cast<const uint32_t*>(cell_planes + cell) + k);
    }
    __syncthreads();
    if (light_count != 0u) {                                                     // :160-165: no lights, every count is zero
        uint32_t mn = 0xFFFFFFFFu, mx = 0u, worst = 0u;                          // the cell's depth bounds, per warp
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint2 mm = s_mm[lane + 32 * h];
            if ((s_in[h] >> lane) & 1u) { mn = min(mn, mm.x); mx = max(mx, mm.y); worst = max(worst, max(mm.x, mm.y)); }
        }
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) {
            mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, sft));
            mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, sft));
            worst = max(worst, __shfl_xor_sync(0xFFFFFFFFu, worst, sft));
        }
        // uint order == float order only when every tile bound is a non-negative non-NaN float
        const bool zsane = worst <= 0x7F800000u && mn <= mx;
        const float cw4 = -__uint_as_float(mn), cw5 = __uint_as_float(mx);
        const uint32_t nc = *cand_count;
        uint32_t n_buf = 0u;
        for (uint32_t start = 0; start < nc; start += kFuseRound) {
            // ---- pass 3 on candidates [start, start + kFuseRound)
            float4 q[kFuseIters]; uint4 zq[kFuseIters];
#pragma unroll
            for (int it = 0; it < kFuseIters; ++it) {
                const uint32_t j = start + it * kFuseThreads + threadIdx.x;
                q[it] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); zq[it] = make_uint4(0u, 0u, 0u, 0u);
                if (j < nc) {
                    q[it] = __ldg(reinterpret_cast<const float4*>(cand) + j);
                    zq[it] = __ldg(reinterpret_cast<const uint4*>(cand_idx) + j);              // CandZ
                }
            }
            uint32_t ballots[kFuseIters];
            {
                V3 fa[4], fb[4];
                float smin[4], emax[4];
                bool fvalid[4];
#pragma unroll
                for (int f = 0; f < 4; ++f) {
                    fa[f] = {s_cp.fa[f][0], s_cp.fa[f][1], s_cp.fa[f][2]};
                    fb[f] = {s_cp.fb[f][0], s_cp.fb[f][1], s_cp.fb[f][2]};
                    smin[f] = s_cp.smin[f]; emax[f] = s_cp.emax[f]; fvalid[f] = (s_cp.fvalid >> f) & 1u;
                }
#pragma unroll
                for (int it = 0; it < kFuseIters; ++it) {
                    const uint32_t j = start + it * kFuseThreads + threadIdx.x;
                    bool keep = false;
                    if (j < nc) {
                        const V3 c = {q[it].x, q[it].y, q[it].z};
                        const float nr = -q[it].w;
                        bool reject = false;
                        if (zsane) reject = (fadd(__uint_as_float(zq[it].x), cw4) < nr) || (fadd(__uint_as_float(zq[it].y), cw5) < nr);
                        const float clen = sqrtf(pdot(c, c)) * 1.000001f;
                        const float slack = 1e-6f * (fabsf(q[it].w) + clen);
#pragma unroll
                        for (int f = 0; f < 4; ++f) {
                            const float m = fmaxf(pdot(fa[f], c), pdot(fb[f], c));
                            // every comparison is false on NaN / inf operands, so those candidates are kept
                            if (fvalid[f] && m < 0.0f && smin[f] * m + emax[f] * clen + slack < nr) reject = true;
                        }
                        keep = !reject;
                    }
                    ballots[it] = __ballot_sync(0xFFFFFFFFu, keep);
                    if (lane == 0) s_cnt[it * kFuseWarps + warp] = __popc(ballots[it]);
                }
            }
            __syncthreads();
            if (warp == 0) {                               // exclusive scan of the kFuseIters x kFuseWarps counts, one per lane
                const uint32_t v = lane < kFuseIters * kFuseWarps ? s_cnt[lane] : 0u;
                uint32_t incl = v;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
                if (lane < kFuseIters * kFuseWarps) s_cnt[lane] = incl - v;
                if (lane == 31) s_total = incl;
            }
            __syncthreads();
#pragma unroll
            for (int it = 0; it < kFuseIters; ++it)
                if ((ballots[it] >> lane) & 1u) {
                    const uint32_t slot = n_buf + s_cnt[it * kFuseWarps + warp] + __popc(ballots[it] & ((1u << lane) - 1u));
                    s_q[slot] = q[it];
                    s_z[slot] = make_float2(__uint_as_float(zq[it].x), __uint_as_float(zq[it].y));
                    s_li[slot] = zq[it].z;
                }
            n_buf += s_total;
            __syncthreads();                               // survivors appended; s_cnt / s_total may be rewritten
            // ---- pass 4 over the buffered survivors, once the buffer could not take another round (or at the end)
            if (start + kFuseRound < nc && n_buf + kFuseRound <= kFuseBuf) continue;
            for (int tt = 0; tt < kFuseTilesPerWarp; ++tt) {
                const int t = warp * kFuseTilesPerWarp + tt;
                if (!((s_in[t >> 5] >> (t & 31)) & 1u)) continue;       // outside the image (uniform per warp)
                const uint2 mm = s_mm[t];
                const TilePlanes p = s_planes[t];
                const V3 n0 = {p.a.x, p.a.y, p.a.z}, n1 = {p.a.w, p.b.x, p.b.y}, n2 = {p.b.z, p.b.w, p.c.x}, n3 = {p.c.y, p.c.z, p.c.w};
                const float w4 = -__uint_as_float(mm.x), w5 = __uint_as_float(mm.y);       // :96-97
                uint32_t* out = tile_data + (size_t)((ty0 + t / kCellTiles) * tiles_x + tx0 + t % kCellTiles) * ASH_TILE_STRIDE;   // :119
                uint32_t count = s_tcnt[t];
                for (uint32_t k0 = 0; k0 < n_buf; k0 += 32) {
                    const uint32_t k = k0 + lane;
                    bool pass = false;
                    if (k < n_buf) {
                        const float4 sq = s_q[k];
                        const float2 z = s_z[k];
                        const V3 c = {sq.x, sq.y, sq.z};
                        const float nr = -sq.w;
                        // sphereInFrustum, :101-109: dist = dot(plane.xyz, c) + plane.w ; reject if dist < -radius.
                        // Side planes have w = 0: (d + 0.0) < nr  <=>  d < nr.  Planes 4/5: the dot is precomputed per light.
                        pass = !(dot3(n0, c) < nr) && !(dot3(n1, c) < nr) && !(dot3(n2, c) < nr) && !(dot3(n3, c) < nr) &&
                               !(fadd(z.x, w4) < nr) && !(fadd(z.y, w5) < nr);
                    }
                    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, pass);
                    if (pass) {
                        const uint32_t slot = count + __popc(ballot & ((1u << lane) - 1u));
                        if (slot < ASH_MAX_LIGHTS_PER_TILE) out[1 + slot] = s_li[k];         // :186-189
                    }
                    count += __popc(ballot);
                }
                __syncwarp();
                if (lane == 0) s_tcnt[t] = count;
                __syncwarp();
            }
            n_buf = 0u;
            __syncthreads();                               // the buffer is rewritten by the next round
        }
    }
    __syncthreads();
    if (threadIdx.x < kCellTiles * kCellTiles) {
        if ((s_in[threadIdx.x >> 5] >> (threadIdx.x & 31)) & 1u)
            tile_data[(size_t)((ty0 + threadIdx.x / kCellTiles) * tiles_x + tx0 + threadIdx.x % kCellTiles) * ASH_TILE_STRIDE] =
                min(s_tcnt[threadIdx.x], ASH_MAX_LIGHTS_PER_TILE);                                                              // :196-197
    }
    if (warp == 0 && light_count != 0u) {   // one set of stats atomics per cell (64 tiles), not per tile
        uint32_t sum = 0, mx = 0, ovf = 0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t c = s_tcnt[lane + 32 * h];
            if ((s_in[h] >> lane) & 1u) { const uint32_t m = min(c, ASH_MAX_LIGHTS_PER_TILE); sum += m; mx = max(mx, m); ovf += c > ASH_MAX_LIGHTS_PER_TILE ? 1u : 0u; }
        }
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) {
            sum += __shfl_xor_sync(0xFFFFFFFFu, sum, sft);
            mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, sft));
            ovf += __shfl_xor_sync(0xFFFFFFFFu, ovf, sft);
        }
        if (lane == 0) {
            if (sum) atomicAdd(&stats->sum_tile_lights, sum);
            if (mx) atomicMax(&stats->max_tile_lights, mx);
            if (ovf) atomicAdd(&stats->overflow_tiles, ovf);
        }
    }
}

static size_t g_cell_planes_off_dummy;
size_t light_cull_scratch_bytes(uint32_t W, uint32_t H, uint32_t max_lights, size_t* planes_off, size_t* list_off,
                                size_t* count_off, size_t* cell_planes_off) {
    if (!cell_planes_off) cell_planes_off = &g_cell_planes_off_dummy;
    const uint32_t tiles_x = (W + 15u) / 16u, tiles_y = (H + 15u) / 16u;
    const uint32_t cells = ((tiles_x + kCellTiles - 1) / kCellTiles) * ((tiles_y + kCellTiles - 1) / kCellTiles);
    auto up = [](size_t v) { return (v + 255) & ~size_t(255); };
    size_t off = 0;
    *planes_off = off; off = up(off + (size_t)tiles_x * tiles_y * sizeof(TilePlanes));
    // (the per-cell survivor lists lived here until the prefilter and the exact pass were fused)
    *list_off = off; *count_off = off;  off = up(off + (2 + (max_lights + kPrepassThreads - 1) / kPrepassThreads + 1) * sizeof(uint32_t));   // k_light_prepass sync words (zeroed once)
    *cell_planes_off = off; off = up(off + (size_t)cells * sizeof(CellPlanes));
    return off;
}

cudaError_t launch_light_cull(const AshGpuLight* lights, const AshLightCullingPushConstants& pc,
                              const AshCullingCameraData* cam_dev, const uint32_t* tile_minmax,
                              uint32_t* depth_range, LightCand* cand, uint32_t* cand_idx,
                              uint32_t* cand_count, uint8_t* scratch, uint32_t max_lights,
                              uint32_t* tile_data, AshViewStats* stats, bool build_planes, AshCullingStats* cstats_or_null,
                              const Launch& L) {
    const uint32_t W = pc.screen_size[0], H = pc.screen_size[1];
    const uint32_t tiles_x = (W + 15u) / 16u, tiles_y = (H + 15u) / 16u, tiles = tiles_x * tiles_y;
    const uint32_t cells_x = (tiles_x + kCellTiles - 1) / kCellTiles, cells_y = (tiles_y + kCellTiles - 1) / kCellTiles;
    const uint32_t cells = cells_x * cells_y;
    size_t po, lo, co, cpo;
    light_cull_scratch_bytes(W, H, max_lights, &po, &lo, &co, &cpo);
    CellPlanes* cell_planes = reinterpret_cast<CellPlanes*>(scratch + cpo);
    TilePlanes* planes = reinterpret_cast<TilePlanes*>(scratch + po);
    cudaError_t e;
    {
        Timed timed(L, ASH_KERNEL_LIGHT_PREPASS);
        const uint32_t nblk = pc.light_count ? (pc.light_count + kPrepassThreads - 1) / kPrepassThreads : 1u;
        k_light_prepass<<<nblk, kPrepassThreads, 0, L.s>>>(lights, pc.light_count, cam_dev, depth_range, cand, cand_idx, cand_count,
                                                         stats, tiles, cstats_or_null, reinterpret_cast<uint32_t*>(scratch + co));
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (pc.light_count != 0u) {
            if (build_planes) {          // the side planes depend on invProjection and the extent only: kept across views
                k_tile_planes<<<(tiles + 127) / 128, 128, 0, L.s>>>(cam_dev, W, H, tiles_x, tiles, planes);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
                k_cell_planes<<<(cells + kCellWarps - 1) / kCellWarps, kCellWarps * 32, 0, L.s>>>(tiles_x, tiles_y, cells_x, cells, planes, cell_planes);
                if ((e = cudaGetLastError()) != cudaSuccess) return e;
                *L.launches += 2;
            }
        }
    }
    Timed timed(L, ASH_KERNEL_LIGHT_TILES);
    k_light_cell_tiles<<<cells, kFuseThreads, 0, L.s>>>(tiles_x, tiles_y, cells_x, pc.light_count, cell_planes, planes, tile_minmax,
                                                       cand, cand_idx, cand_count, tile_data, stats);
    return cudaGetLastError();
}

// =====================================================================================================
// Tile light lists in CSR form for host consumers: the reference's table is tileData[tile * 257] = {count,
// 256 slots} (light_culling.comp:43-45, 33 MB at 4K, almost all of it unused slots); the packed form is
// offsets[tiles + 1] (exclusive scan of the counts) + the concatenated light indices in tile order — the same
// information in about 1 MB.  offsets[tiles] is the total; entries beyond `capacity` are not written, so the
// host detects a too-small destination from the total and falls back to the full table.
// =====================================================================================================
constexpr int kPackThreads = 1024;
__global__ void __launch_bounds__(kPackThreads)
k_tile_pack_offsets(const uint32_t* __restrict__ tile_data, uint32_t tiles, uint32_t* __restrict__ offsets) {
    __shared__ uint32_t s_warp[kPackThreads / 32];
    const uint32_t per = (tiles + kPackThreads - 1) / kPackThreads;          // consecutive tiles per thread
    const uint32_t t0 = min(threadIdx.x * per, tiles), t1 = min(t0 + per, tiles);
    uint32_t sum = 0;
    for (uint32_t t = t0; t < t1; ++t) sum += __ldg(tile_data + (size_t)t * ASH_TILE_STRIDE);
    uint32_t incl = sum;                                                     // warp inclusive scan
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if ((threadIdx.x & 31) >= d) incl += v; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        const uint32_t w = s_warp[threadIdx.x];
        uint32_t wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, wi, d); if (threadIdx.x >= d) wi += v; }
        s_warp[threadIdx.x] = wi - w;                                        // exclusive warp bases
    }
    __syncthreads();
    uint32_t run = s_warp[threadIdx.x >> 5] + incl - sum;
    for (uint32_t t = t0; t < t1; ++t) { offsets[t] = run; run += __ldg(tile_data + (size_t)t * ASH_TILE_STRIDE); }
    if (threadIdx.x == kPackThreads - 1) offsets[tiles] = run;              // the last thread's chunk ends the table
}

__global__ void __launch_bounds__(256)
k_tile_pack_lists(const uint32_t* __restrict__ tile_data, uint32_t tiles, const uint32_t* __restrict__ offsets,
                  uint32_t* __restrict__ lights, uint32_t capacity) {
    const uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (t >= tiles) return;
    const uint32_t* src = tile_data + (size_t)t * ASH_TILE_STRIDE;
    const uint32_t n = __ldg(src), o = __ldg(offsets + t);
    for (uint32_t i = lane; i < n; i += 32) if (o + i < capacity) lights[o + i] = __ldg(src + 1 + i);
}

cudaError_t launch_tile_pack(const uint32_t* tile_data, uint32_t tiles, uint32_t* offsets, uint32_t* lights,
                             uint32_t capacity, const Launch& L) {
    Timed timed(L, ASH_KERNEL_LIGHT_TILES);
    k_tile_pack_offsets<<<1, kPackThreads, 0, L.s>>>(tile_data, tiles, offsets);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    k_tile_pack_lists<<<(tiles + 7) / 8, 256, 0, L.s>>>(tile_data, tiles, offsets, lights, capacity);
    ++*L.launches;
    return cudaGetLastError();
}

}  // namespace ash
