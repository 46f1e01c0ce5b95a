// occlusion_cull.cu — per-instance frustum + Hi-Z occlusion culling with ordered compaction.
//
// Replaces shaders/occlusion_cull.comp:184-233 (local_size 64, one global atomicAdd per visible
// object, arbitrary output order).  Here: one thread per object; the 6 frustum planes, which
// depend only on the push constant (:207-208), are built once per CTA; visibility is compacted
// with warp ballots, and draw commands are written in ascending object order (PIN-7) by a second
// tiny kernel that prefix-sums the per-CTA counts — no per-object global atomics at all.
//
// Algorithmic bytes per launch: 112*N (objects) + 4*N (visibility) + 4*N (one Hi-Z texel)
// + 40*V (draw command in + out) + 4.
#include <cstring>
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

constexpr int kCullThreads = 256;

struct CullParams {
    M4       vp;
    float    screen_w, screen_h;
    uint32_t object_count, hiz_levels;
};

// exact ceil(log2(s)) clamped to [0,hi]: never calls log2f (occlusion_cull.comp:163-164)
__device__ __forceinline__ int mip_from_size(float s, int hi) {
    if (!(s > 1.0f)) return 0;
    const uint32_t b = __float_as_uint(s);
    if ((b >> 23) == 0xFFu) return hi;
    const int e = (int)(b >> 23) - 127;
    const int k = (b & 0x7FFFFFu) ? e + 1 : e;
    return k > hi ? hi : k;
}

__device__ __forceinline__ float4 ldg4(const void* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// MODE 0: the reference dispatch.  Two-phase culling (row f-3, ashcuda_view_run with ASH_STAGE_TWO_PHASE):
// MODE 1 = phase 1 against the PREVIOUS frame's pyramid (no_occlusion != 0 on a first frame: frustum only);
// MODE 2 = phase 2 against the pyramid of what phase 1 drew: objects already visible stay visible untested,
// the others are tested and the newly visible ones flagged in late[] (they are drawn, then the final
// pyramid is built).  visibility / vis_bits / cta_counts always describe every object.
template <int MODE>
__global__ void __launch_bounds__(kCullThreads)
k_occlusion_cull(const AshCullObjectData* __restrict__ objs, CullParams cp, Pyramid pyr,
                 uint32_t* __restrict__ visibility, uint32_t* __restrict__ vis_bits,
                 uint32_t* __restrict__ cta_counts, uint32_t no_occlusion, uint32_t* __restrict__ late,
                 AshCullingStats* __restrict__ cstats) {
    __shared__ float4   s_planes[6];
    __shared__ uint32_t s_warp_cnt[kCullThreads / 32];
    __shared__ uint32_t s_warp_aux[kCullThreads / 32];

    // extractFrustumPlanes, :82-100 — built from COLUMNS of viewProj, as the shader does
    if (threadIdx.x < 6) {
        const int i = threadIdx.x >> 1;
        const V4 a = cp.vp.c[3], b = cp.vp.c[i];
        V4 pl;
        if (threadIdx.x & 1) pl = {fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z), fsub(a.w, b.w)};
        else                 pl = {fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z), fadd(a.w, b.w)};
        const float l = length3(v3(pl.x, pl.y, pl.z));
        s_planes[threadIdx.x] = make_float4(fdiv(pl.x, l), fdiv(pl.y, l), fdiv(pl.z, l), fdiv(pl.w, l));
    }
    __syncthreads();

    const uint32_t idx = blockIdx.x * kCullThreads + threadIdx.x;
    bool visible = false;
    bool aux = false;                 // MODE 0/1: passed the frustum test; MODE 2: newly visible (late)
    const bool early = MODE == 2 && idx < cp.object_count && visibility[idx] != 0u;
    if (early) visible = true;
    if (idx < cp.object_count && !early) {
        const char* o = reinterpret_cast<const char*>(objs + idx);
        const float4 tail = ldg4(o + 96);                        // drawIndex, lodBias, flags, pad
        if (__float_as_uint(tail.z) & 1u) {                      // :195-198
            const float4 bc = ldg4(o), be = ldg4(o + 16);
            const float4 c0 = ldg4(o + 32), c1 = ldg4(o + 48), c2 = ldg4(o + 64), c3 = ldg4(o + 80);
            // :201-207 world centre and rotation-ignoring extents
            const V3 wc = {mrow(c0.x, c1.x, c2.x, c3.x, bc.x, bc.y, bc.z, 1.0f),
                           mrow(c0.y, c1.y, c2.y, c3.y, bc.x, bc.y, bc.z, 1.0f),
                           mrow(c0.z, c1.z, c2.z, c3.z, bc.x, bc.y, bc.z, 1.0f)};
            const V3 we = {fmul(be.x, length3(v3(c0.x, c0.y, c0.z))), fmul(be.y, length3(v3(c1.x, c1.y, c1.z))),
                           fmul(be.z, length3(v3(c2.x, c2.y, c2.z)))};
            // frustumCull, :103-116
            bool culled = false;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const float4 pl = s_planes[i];
                const V3 n = {pl.x, pl.y, pl.z};
                const V3 pv = add3(wc, mul3(we, v3(gsign(n.x), gsign(n.y), gsign(n.z))));
                if (fadd(dot3(n, pv), pl.w) < 0.0f) { culled = true; break; }
            }
            if (MODE != 2) aux = !culled;
            // occlusionCull, :123-178
            if (!culled && MODE == 1 && no_occlusion) visible = true;
            else if (!culled) {
                float minx = 1.0f, miny = 1.0f, maxx = 0.0f, maxy = 0.0f, minz = 1.0f;
                bool behind = false;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    // corner order of :125-134: x sign pattern -,+,+,- ; y -,-,+,+ ; z by k>=4
                    const float sx = ((k & 3) == 1 || (k & 3) == 2) ? we.x : -we.x;
                    const float sy = (k & 2) ? we.y : -we.y;
                    const float sz = (k & 4) ? we.z : -we.z;
                    const V4 clip = mul4(cp.vp, fadd(wc.x, sx), fadd(wc.y, sy), fadd(wc.z, sz), 1.0f);
                    if (clip.w <= 0.0f) { behind = true; break; }          // :145-147 -> visible
                    const float nx = fdiv(clip.x, clip.w), ny = fdiv(clip.y, clip.w), nz = fdiv(clip.z, clip.w);
                    const float px = fadd(fmul(nx, 0.5f), 0.5f), py = fadd(fmul(ny, 0.5f), 0.5f);
                    minx = gmin(minx, px); miny = gmin(miny, py);
                    maxx = gmax(maxx, px); maxy = gmax(maxy, py);
                    minz = gmin(minz, nz);
                }
                if (behind) {
                    visible = true;
                } else {
                    minx = gclamp(minx, 0.0f, 1.0f); miny = gclamp(miny, 0.0f, 1.0f);
                    maxx = gclamp(maxx, 0.0f, 1.0f); maxy = gclamp(maxy, 0.0f, 1.0f);
                    const float ssx = fmul(fsub(maxx, minx), cp.screen_w);
                    const float ssy = fmul(fsub(maxy, miny), cp.screen_h);
                    // the image's own mip count bounds the sampler's level as well (textureLod clamps to the last mip):
                    // the reference pushes HIZ_LEVELS = 10 whatever the extent (occlusion_culling.rs:21,262)
                    const int level = mip_from_size(gmax(ssx, ssy), min((int)(cp.hiz_levels - 1u), pyr.levels - 1));
                    const float u = fmul(fadd(minx, maxx), 0.5f), v = fmul(fadd(miny, maxy), 0.5f);
                    const int lw = (int)pyr.w[level], lh = (int)pyr.h[level];
                    const float* base = level == 0 ? pyr.mip0 : pyr.mips + pyr.off[level];
                    const float hiz = __ldg(base + (size_t)nearest_texel(v, lh) * lw + nearest_texel(u, lw));
                    visible = !(minz > hiz);                                  // :173
                }
            }
        }
        if (MODE == 2) aux = visible;
        visibility[idx] = visible ? 1u : 0u;
    }
    if (MODE == 2 && idx < cp.object_count) late[idx] = aux ? 1u : 0u;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, visible);
    const uint32_t ballot_aux = __ballot_sync(0xFFFFFFFFu, aux);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        if (blockIdx.x * kCullThreads + warp * 32 < cp.object_count) vis_bits[idx >> 5] = ballot;
        s_warp_cnt[warp] = __popc(ballot);
        s_warp_aux[warp] = __popc(ballot_aux);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t c = 0, a = 0;
#pragma unroll
        for (int w = 0; w < kCullThreads / 32; ++w) { c += s_warp_cnt[w]; a += s_warp_aux[w]; }
        cta_counts[blockIdx.x] = c;
        // CullingStats (render_stats.rs:54-65): integer atomics, order independent
        if (MODE != 2) { if (a) atomicAdd(&cstats->after_frustum, a); }
        else if (a) atomicAdd(&cstats->late_visible, a);
        if (MODE == 1 && c) atomicAdd(&cstats->early_visible, c);
    }
}

// Ordered compaction (PIN-7): slot = #visible objects with a smaller index.
__global__ void __launch_bounds__(kCullThreads)
k_cull_compact(const AshCullObjectData* __restrict__ objs, const AshIndirectDrawCommand* __restrict__ draws_in,
               uint32_t n_draws, uint32_t object_count, const uint32_t* __restrict__ vis_bits,
               const uint32_t* __restrict__ cta_counts, AshIndirectDrawCommand* __restrict__ draws_out,
               uint32_t* __restrict__ visible_count, AshViewStats* __restrict__ stats,
               AshCullingStats* __restrict__ cstats) {
    __shared__ uint32_t s_red[kCullThreads / 32];
    __shared__ uint32_t s_warp_off[kCullThreads / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // exclusive prefix of this CTA = sum of the counts of all earlier CTAs
    uint32_t part = 0;
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += kCullThreads) part += cta_counts[b];
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, s);
    if (lane == 0) s_red[warp] = part;
    const uint32_t idx = blockIdx.x * kCullThreads + threadIdx.x;
    const uint32_t word = (blockIdx.x * kCullThreads + warp * 32 < object_count) ? vis_bits[idx >> 5] : 0u;
    if (lane == 0) s_warp_off[warp] = __popc(word);
    __syncthreads();
    uint32_t base = 0, woff = 0;
#pragma unroll
    for (int w = 0; w < kCullThreads / 32; ++w) {
        base += s_red[w];
        if (w < warp) woff += s_warp_off[w];
    }
    if ((word >> lane) & 1u) {
        const uint32_t slot = base + woff + __popc(word & ((1u << lane) - 1u));
        const uint32_t di = objs[idx].draw_index;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(draws_in + di);   // :229-232
        uint32_t* dst = reinterpret_cast<uint32_t*>(draws_out + slot);
        // a drawIndex past the template buffer reads zeros, as Vulkan's robustBufferAccess returns for the shader's
        // drawCommandsIn[obj.drawIndex] (an unchecked read would be a sticky illegal-address fault here)
#pragma unroll
        for (int k = 0; k < 5; ++k) dst[k] = di < n_draws ? __ldg(src + k) : 0u;
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
        const uint32_t total = base + cta_counts[blockIdx.x];
        *visible_count = total;
        stats->total_objects = object_count;
        stats->visible_objects = total;
        cstats->objects_submitted = object_count;         // StatsCollector::record_culling, render_stats.rs:213-217
        cstats->after_occlusion = total;
    }
}

static CullParams make_cull_params(const AshCullingPushConstants& pc) {
    CullParams cp;
    for (int c = 0; c < 4; ++c)
        cp.vp.c[c] = {pc.view_proj[c][0], pc.view_proj[c][1], pc.view_proj[c][2], pc.view_proj[c][3]};
    cp.screen_w = pc.screen_params[0];
    cp.screen_h = pc.screen_params[1];
    cp.object_count = pc.object_count;
    cp.hiz_levels = pc.hiz_levels;
    return cp;
}

// graph replay: argument 1 of k_occlusion_cull is the per-view CullParams (every mode of a captured view gets it)
void cull_patch(const AshCullingPushConstants& pc, int mode, PatchArg* out) {
    static_assert(sizeof(CullParams) <= sizeof(out->value), "PatchArg too small");
    const CullParams cp = make_cull_params(pc);
    out->func = mode == 0 ? reinterpret_cast<const void*>(&k_occlusion_cull<0>)
              : mode == 1 ? reinterpret_cast<const void*>(&k_occlusion_cull<1>) : reinterpret_cast<const void*>(&k_occlusion_cull<2>);
    out->arg_index = 1; out->n_args = 9;
    std::memcpy(out->value, &cp, sizeof(cp));
}

// phase: 0 = the reference dispatch + compaction; 1 = phase 1 (no compaction); 2 = phase 2 + compaction
cudaError_t launch_occlusion_cull(const AshCullObjectData* objs, const AshIndirectDrawCommand* draws_in,
                                  uint32_t n_draws, const AshCullingPushConstants& pc, const Pyramid& p,
                                  uint32_t* visibility, uint32_t* vis_bits, uint32_t* cta_counts,
                                  AshIndirectDrawCommand* draws_out, uint32_t* visible_count,
                                  AshViewStats* stats, AshCullingStats* cstats, int phase, bool no_occlusion,
                                  uint32_t* late, const Launch& L) {
    const CullParams cp = make_cull_params(pc);
    const uint32_t grid = (pc.object_count + kCullThreads - 1) / kCullThreads;
    if (phase != 2) {
        cudaError_t e = cudaMemsetAsync(cstats, 0, sizeof(AshCullingStats), L.s);
        if (e != cudaSuccess) return e;
    }
    {
        Timed timed(L, ASH_KERNEL_OCCLUSION_CULL);
        if (phase == 0) k_occlusion_cull<0><<<grid, kCullThreads, 0, L.s>>>(objs, cp, p, visibility, vis_bits, cta_counts, 0u, nullptr, cstats);
        else if (phase == 1) k_occlusion_cull<1><<<grid, kCullThreads, 0, L.s>>>(objs, cp, p, visibility, vis_bits, cta_counts, no_occlusion ? 1u : 0u, nullptr, cstats);
        else k_occlusion_cull<2><<<grid, kCullThreads, 0, L.s>>>(objs, cp, p, visibility, vis_bits, cta_counts, 0u, late, cstats);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || phase == 1) return e;
    Timed timed(L, ASH_KERNEL_CULL_COMPACT);
    k_cull_compact<<<grid, kCullThreads, 0, L.s>>>(objs, draws_in, n_draws, pc.object_count, vis_bits, cta_counts,
                                                 draws_out, visible_count, stats, cstats);
    return cudaGetLastError();
}

// lights_culled and lod_triangles_saved of CullingStats from the view's other statistics (end of view_run)
__global__ void k_culling_stats_finish(const AshViewStats* __restrict__ stats, const AshInstancingStats* __restrict__ inst_stats,
                                       uint32_t light_count, uint32_t with_lights, AshCullingStats* __restrict__ cstats) {
    if (with_lights) cstats->lights_culled = light_count - min(light_count, stats->candidate_lights);
    if (inst_stats) cstats->lod_triangles_saved = inst_stats->triangles_before - inst_stats->triangles_after;   // record_lod_savings, :232-234
}
cudaError_t launch_culling_stats_finish(const AshViewStats* stats, const AshInstancingStats* inst_stats, uint32_t light_count,
                                        bool with_lights, AshCullingStats* cstats, const Launch& L) {
    k_culling_stats_finish<<<1, 1, 0, L.s>>>(stats, inst_stats, light_count, with_lights ? 1u : 0u, cstats);
    *L.launches += 1;
    return cudaGetLastError();
}

}  // namespace ash
