// light_cull.cu — 16x16 tiled light culling into per-tile index lists.
//
// Replaces shaders/light_culling.comp:111-203.  The shader re-reads depth per tile, transforms
// every light to view space once per tile (tiles x lights mat-vecs) and appends through a shared
// atomic in arbitrary order.  Here:
//   k_light_prepass (1 CTA): view-space transform once per light per view (:181 is tile
//     independent) and an exact global prefilter: a light that fails the near/far planes (:96-97)
//     evaluated with the frame's global [min,max] depth fails them in every tile (rounding is
//     monotone and tile min >= global min, tile max <= global max), so only survivors
//     ("candidates") are kept, in ascending light index.
//   k_light_tiles (1 warp per tile): tile min/max come from the depth pyramid kernel (depth is not
//     read again); 4 lanes build the tile frustum (:71-98); lanes test 32 candidates at a time
//     with the shader's exact predicate (:101-109) and append by ballot + popc prefix, which
//     yields ascending order and first-256-on-overflow (PIN-7) without atomics.
//
// Algorithmic bytes per launch pair: 64*L (lights) + 208 (camera) + 8*tiles (tile min/max)
// + 4*sum_tiles(1+count).
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

constexpr int kPrepassThreads = 1024;
constexpr int kTileWarps = 16;

// One candidate = 32 bytes: {view x, y, z, radius} {dot((0,0,-1),c), dot((0,0,1),c), light index bits, 0}.
// The two z dots are the tile-independent parts of planes 4/5 (:96-97, :104), evaluated exactly as the
// shader does (so non-finite light positions behave identically) once per light instead of once per tile.
struct CandZ { float zneg, zpos; uint32_t index, pad; };

__global__ void __launch_bounds__(kPrepassThreads)
k_light_prepass(const AshGpuLight* __restrict__ lights, uint32_t light_count,
                const AshCullingCameraData* __restrict__ cam, uint32_t* __restrict__ depth_range,
                LightCand* __restrict__ cand, uint32_t* __restrict__ cand_idx,
                uint32_t* __restrict__ cand_count, AshViewStats* __restrict__ stats, uint32_t tiles) {
    __shared__ uint32_t s_warp[kPrepassThreads / 32];
    __shared__ uint32_t s_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const M4 view = load_m4(cam->view);
    const uint32_t gmn = depth_range[0], gmx = ~depth_range[1];
    // uint order == float order only for non-negative, non-NaN depth (PIN-2); otherwise keep all
    const bool sane = gmx <= 0x7F800000u && gmn <= gmx;
    const float gmin_d = __uint_as_float(gmn), gmax_d = __uint_as_float(gmx);
    CandZ* candz = reinterpret_cast<CandZ*>(cand_idx);
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    // blocked assignment keeps ascending light order with ONE block scan per 32K lights:
    // thread t owns lights [start + t*chunk, start + (t+1)*chunk), chunk <= 32 (keep bits in a register)
    for (uint32_t start = 0; start < light_count; start += kPrepassThreads * 32u) {
        const uint32_t span = min(light_count - start, kPrepassThreads * 32u);
        const uint32_t chunk = (span + kPrepassThreads - 1) / kPrepassThreads;
        const uint32_t first = start + threadIdx.x * chunk;
        uint32_t keep_bits = 0;
        for (uint32_t j = 0; j < chunk; ++j) {
            const uint32_t i = first + j;
            if (i < start + span) {
                const float4 pos = __ldg(reinterpret_cast<const float4*>(lights[i].position));
                const float enabled = __ldg(&lights[i].params[3]);
                bool keep = !(enabled < 0.5f);                                           // :178
                if (keep && sane) {
                    const float z = mrow(view.c[0].z, view.c[1].z, view.c[2].z, view.c[3].z, pos.x, pos.y, pos.z, 1.0f);
                    const float x = mrow(view.c[0].x, view.c[1].x, view.c[2].x, view.c[3].x, pos.x, pos.y, pos.z, 1.0f);
                    const float y = mrow(view.c[0].y, view.c[1].y, view.c[2].y, view.c[3].y, pos.x, pos.y, pos.z, 1.0f);
                    // planes 4/5 with the global slab: a light failing them here fails them in every tile
                    const float d4 = fadd(dot3(v3(0.0f, 0.0f, -1.0f), v3(x, y, z)), -gmin_d);
                    const float d5 = fadd(dot3(v3(0.0f, 0.0f, 1.0f), v3(x, y, z)), gmax_d);
                    if (d4 < -pos.w || d5 < -pos.w) keep = false;
                }
                if (keep) keep_bits |= 1u << j;
            }
        }
        // block-wide exclusive scan of the per-thread counts
        const uint32_t mine = __popc(keep_bits);
        uint32_t incl = mine;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, s);
            if (lane >= s) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t off = s_base, total = 0;
#pragma unroll
        for (int w = 0; w < kPrepassThreads / 32; ++w) {
            const uint32_t n = s_warp[w];
            if (w < warp) off += n;
            total += n;
        }
        uint32_t slot = off + incl - mine;
        while (keep_bits) {
            const uint32_t j = __ffs(keep_bits) - 1;
            keep_bits &= keep_bits - 1;
            const uint32_t i = first + j;
            const float4 pos = __ldg(reinterpret_cast<const float4*>(lights[i].position));
            const V4 pv = mul4(view, pos.x, pos.y, pos.z, 1.0f);                         // :181
            *reinterpret_cast<float4*>(cand + slot) = make_float4(pv.x, pv.y, pv.z, pos.w);
            CandZ cz;
            cz.zneg = dot3(v3(0.0f, 0.0f, -1.0f), v3(pv.x, pv.y, pv.z));
            cz.zpos = dot3(v3(0.0f, 0.0f, 1.0f), v3(pv.x, pv.y, pv.z));
            cz.index = i; cz.pad = 0;
            *reinterpret_cast<uint4*>(candz + slot) = *reinterpret_cast<const uint4*>(&cz);
            ++slot;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *cand_count = s_base;
        depth_range[0] = 0xFFFFFFFFu;       // consumed: reset for the next frame's depth kernel
        depth_range[1] = 0xFFFFFFFFu;
        stats->tiles = tiles;
        stats->candidate_lights = s_base;
        stats->sum_tile_lights = 0;
        stats->max_tile_lights = 0;
        stats->overflow_tiles = 0;
    }
}

// screenToView(uv, 0.0), :64-68
__device__ __forceinline__ V3 screen_to_view0(const M4& ip, float u, float v) {
    const V4 q = mul4(ip, fsub(fmul(u, 2.0f), 1.0f), fsub(fmul(v, 2.0f), 1.0f), 0.0f, 1.0f);
    return {fdiv(q.x, q.w), fdiv(q.y, q.w), fdiv(q.z, q.w)};
}

__global__ void __launch_bounds__(kTileWarps * 32)
k_light_tiles(const AshCullingCameraData* __restrict__ cam, uint32_t W, uint32_t H, uint32_t tiles_x,
              uint32_t tiles, uint32_t light_count, const uint32_t* __restrict__ tile_minmax,
              const LightCand* __restrict__ cand, const uint32_t* __restrict__ cand_idx,
              const uint32_t* __restrict__ cand_count, uint32_t* __restrict__ tile_data,
              AshViewStats* __restrict__ stats) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __shared__ uint32_t s_cnt[kTileWarps];
    const uint32_t tile = min(blockIdx.x * kTileWarps + warp, tiles - 1u);
    const bool active = blockIdx.x * kTileWarps + warp < tiles;
    uint32_t* out = tile_data + (size_t)tile * ASH_TILE_STRIDE;                  // :119
    if (light_count == 0u) { if (lane == 0 && active) out[0] = 0u; return; }     // :160-165 (uniform)
    const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;

    // getTileFrustum, :71-98 — lane k (<4) computes corner k, then plane k
    const M4 ip = load_m4(cam->inv_projection);
    const float sw = (float)W, sh = (float)H, ts = 16.0f;
    const float min_x = fsub(fmul(fdiv(fmul((float)tx, ts), sw), 2.0f), 1.0f);
    const float min_y = fsub(fmul(fdiv(fmul((float)ty, ts), sh), 2.0f), 1.0f);
    const float max_x = fsub(fmul(fdiv(fmul((float)(tx + 1u), ts), sw), 2.0f), 1.0f);
    const float max_y = fsub(fmul(fdiv(fmul((float)(ty + 1u), ts), sh), 2.0f), 1.0f);
    const int k = lane & 3;
    const float cxn = (k == 1 || k == 2) ? max_x : min_x;       // corners 0..3: (min,min) (max,min) (max,max) (min,max)
    const float cyn = (k >= 2) ? max_y : min_y;
    const V3 corner = screen_to_view0(ip, fadd(fmul(cxn, 0.5f), 0.5f), fadd(fmul(cyn, 0.5f), 0.5f));
    // plane k = cross(corners[a], corners[b]) with (a,b) = (3,0) (1,2) (0,1) (2,3)
    const int ia = (k == 0) ? 3 : (k == 1) ? 1 : (k == 2) ? 0 : 2;
    const int ib = (k == 0) ? 0 : (k == 1) ? 2 : (k == 2) ? 1 : 3;
    const V3 a = {__shfl_sync(0xFFFFFFFFu, corner.x, ia), __shfl_sync(0xFFFFFFFFu, corner.y, ia), __shfl_sync(0xFFFFFFFFu, corner.z, ia)};
    const V3 b = {__shfl_sync(0xFFFFFFFFu, corner.x, ib), __shfl_sync(0xFFFFFFFFu, corner.y, ib), __shfl_sync(0xFFFFFFFFu, corner.z, ib)};
    const V3 cr = cross3(a, b);
    const V3 nk = (length3(cr) > 0.0001f) ? normalize3(cr) : v3(0.0f, 0.0f, 0.0f);
    V3 n[4];
#pragma unroll
    for (int p = 0; p < 4; ++p)
        n[p] = {__shfl_sync(0xFFFFFFFFu, nk.x, p), __shfl_sync(0xFFFFFFFFu, nk.y, p), __shfl_sync(0xFFFFFFFFu, nk.z, p)};
    const uint2 mm = __ldg(reinterpret_cast<const uint2*>(tile_minmax) + tile);
    const float w4 = -__uint_as_float(mm.x), w5 = __uint_as_float(mm.y);       // :96-97

    const uint32_t nc = active ? *cand_count : 0u;
    uint32_t count = 0;
    for (uint32_t base = 0; base < nc; base += 32) {
        const uint32_t j = base + lane;
        bool pass = false;
        uint32_t li = 0;
        if (j < nc) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(cand) + j);
            const uint4 zq = __ldg(reinterpret_cast<const uint4*>(cand_idx) + j);      // CandZ
            li = zq.z;
            const V3 c = {q.x, q.y, q.z};
            const float nr = -q.w;
            // sphereInFrustum, :101-109: dist = dot(plane.xyz, c) + plane.w ; reject if dist < -radius.
            // Side planes have w = 0: (d + 0.0) < nr  <=>  d < nr.  Planes 4/5: the dot is precomputed per light.
            pass = !(dot3(n[0], c) < nr) && !(dot3(n[1], c) < nr) && !(dot3(n[2], c) < nr) && !(dot3(n[3], c) < nr) &&
                   !(fadd(__uint_as_float(zq.x), w4) < nr) && !(fadd(__uint_as_float(zq.y), w5) < nr);
        }
        const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, pass);
        if (pass) {
            const uint32_t slot = count + __popc(ballot & ((1u << lane) - 1u));
            if (slot < ASH_MAX_LIGHTS_PER_TILE) out[1 + slot] = li;              // :186-189
        }
        count += __popc(ballot);
    }
    if (lane == 0) {
        if (active) out[0] = min(count, ASH_MAX_LIGHTS_PER_TILE);                // :196-197
        s_cnt[warp] = count;
    }
    __syncthreads();
    if (warp == 0) {   // one set of stats atomics per CTA (16 tiles), not per tile
        const uint32_t c = lane < kTileWarps ? s_cnt[lane] : 0u;
        uint32_t sum = min(c, ASH_MAX_LIGHTS_PER_TILE), mx = sum, ovf = c > ASH_MAX_LIGHTS_PER_TILE ? 1u : 0u;
#pragma unroll
        for (int sft = 8; sft > 0; sft >>= 1) {
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

cudaError_t launch_light_cull(const AshGpuLight* lights, const AshLightCullingPushConstants& pc,
                              const AshCullingCameraData* cam_dev, const uint32_t* tile_minmax,
                              uint32_t* depth_range, LightCand* cand, uint32_t* cand_idx,
                              uint32_t* cand_count, uint32_t* tile_data, AshViewStats* stats,
                              const Launch& L) {
    const uint32_t W = pc.screen_size[0], H = pc.screen_size[1];
    const uint32_t tiles_x = (W + 15u) / 16u, tiles_y = (H + 15u) / 16u, tiles = tiles_x * tiles_y;
    {
        Timed timed(L, ASH_KERNEL_LIGHT_PREPASS);
        k_light_prepass<<<1, kPrepassThreads, 0, L.s>>>(lights, pc.light_count, cam_dev, depth_range, cand,
                                                      cand_idx, cand_count, stats, tiles);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    Timed timed(L, ASH_KERNEL_LIGHT_TILES);
    k_light_tiles<<<(tiles + kTileWarps - 1) / kTileWarps, kTileWarps * 32, 0, L.s>>>(
        cam_dev, W, H, tiles_x, tiles, pc.light_count, tile_minmax, cand, cand_idx, cand_count,
        tile_data, stats);
    return cudaGetLastError();
}

}  // namespace ash
