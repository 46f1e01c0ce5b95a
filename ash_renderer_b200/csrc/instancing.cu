// instancing.cu — consumers of the cull result: per-object LOD selection and
// instance batching into per-(batch, LOD) streams with one indirect draw each.
//
// Replaces, per frame and per (visible) object, the reference's CPU managers
//   LodManager::select_lod            src/renderer/lod_system.rs:238-345 (+ :207-235, :347-383)
//   InstancingManager::add_instance   src/renderer/instancing.rs:212-224
//   InstancingManager::finalize       src/renderer/instancing.rs:253-263
// Semantics kept: objects are visited in ascending index, so every bucket's stream is in ascending
// object order (the order add_instance would build it) and a bucket keeps its first
// MAX_INSTANCES_PER_BATCH instances (:219-223).  fp32 arithmetic follows glam 0.30's scalar order
// (Mat4 * Vec4 = ((x_axis*v.x + y_axis*v.y) + z_axis*v.z) + w_axis*v.w, no FMA), PIN-1.
//
//   k_lod_select       1 thread / object: selection, stats, bucket id, per-CTA bucket histogram
//   k_bucket_scan_ctas 1 warp / bucket: exclusive scan of the per-CTA counts
//   k_bucket_offsets   1 CTA: bucket offsets, one indirect draw per bucket, finalize() statistics
//   k_instance_scatter 1 thread / object: ordered slot = CTA base + earlier warps + rank in warp; 96 B copy
//
// Algorithmic bytes per call: N*(4 + 4 + 4) (visibility, batch id, mesh id) + 12 B position per visible
// object + 16 B selection per object + 2*96 B per kept instance + 20 B per bucket.
#include <cstring>
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

constexpr int kInstThreads = 256;
constexpr uint32_t kNoBucket = 0xFFFFFFFFu;

// LodManager::calculate_screen_coverage, lod_system.rs:207-235
__device__ __forceinline__ float screen_coverage(const float* __restrict__ vp /* 16 floats, columns */, float px, float py,
                                                  float pz, float radius, float sw, float sh) {
    const float clip_w = mrow(vp[3], vp[7], vp[11], vp[15], px, py, pz, 1.0f);     // (view_proj * position.extend(1.0)).w
    if (clip_w <= 0.0f) return 0.0f;                                                // behind camera, :218-220
    const float proj_radius = fdiv(radius, clip_w);                                 // :225-226
    const float pixel_radius = fmul(fmul(proj_radius, fmaxf(sw, sh)), 0.5f);        // :229
    const float area = fmul(fmul(3.14159274101257324f, pixel_radius), pixel_radius);   // :232
    return fminf(fdiv(area, fmul(sw, sh)), 1.0f);                                   // :233-235 (f32::min drops a NaN)
}

// Rust `x.clamp(lo, hi)` then `as usize` (saturating, NaN -> 0)
__device__ __forceinline__ uint32_t clamp_to_index(float x, float hi) {
    if (x < 0.0f) x = 0.0f;
    if (x > hi) x = hi;
    if (!(x == x)) return 0u;
    return (uint32_t)x;          // x in [0, hi], hi <= 7
}

struct LodParams {
    float vp[16];
    float cam[3];
    float sw, sh;
    AshLodConfig cfg;
};

__device__ __forceinline__ AshLodSelection select_lod(const AshLodMesh& m, float px, float py, float pz, const LodParams& P,
                                                      uint32_t& stat_level /* objects_per_lod index or 0xFF */,
                                                      uint32_t& tri_after, bool& counted_after) {
    const uint32_t n = m.level_count < ASH_MAX_LOD_LEVELS ? m.level_count : ASH_MAX_LOD_LEVELS;
    const uint32_t last = n ? n - 1u : 0u;                       // levels.len().saturating_sub(1)
    stat_level = 0xFFu; tri_after = 0u; counted_after = false;
    AshLodSelection s;
    if (P.cfg.mode == ASH_LOD_MODE_FORCE) {                      // :255-268
        const uint32_t level = P.cfg.force_level < last ? P.cfg.force_level : last;
        if (n) { tri_after = m.levels[level].triangle_count; counted_after = true; stat_level = level; }
        s.level = level; s.blend = 0.0f; s.culled = 0u; s.screen_coverage = 1.0f;
        return s;
    }
    const float cov = screen_coverage(P.vp, px, py, pz, m.bounding_radius, P.sw, P.sh);     // :271-277
    if (cov < P.cfg.cull_threshold) {                            // :280-288
        s.level = last; s.blend = 0.0f; s.culled = 1u; s.screen_coverage = cov;
        return s;
    }
    uint32_t level = last; float blend = 0.0f;
    if (P.cfg.mode == ASH_LOD_MODE_SCREEN_SIZE) {                // select_by_screen_size, :347-369
        for (uint32_t i = 0; i < n; ++i) {
            const float th = m.levels[i].screen_threshold;
            if (cov >= th) {
                if (P.cfg.smooth_transitions && i + 1u < n) {
                    const float range = fsub(th, m.levels[i + 1u].screen_threshold);
                    if (range > 0.0f) {
                        float b = fdiv(fdiv(fsub(th, cov), range), P.cfg.transition_width);
                        if (b < 0.0f) b = 0.0f;                   // f32::clamp(0.0, 1.0): NaN stays NaN
                        if (b > 1.0f) b = 1.0f;
                        blend = b;
                    }
                }
                level = i;
                break;
            }
        }
    } else {                                                     // select_by_distance, :371-378
        const float dx = fsub(px, P.cam[0]), dy = fsub(py, P.cam[1]), dz = fsub(pz, P.cam[2]);
        const float dist = fsqrt(fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz)));
        for (uint32_t i = 0; i < n; ++i)
            if (dist <= m.levels[i].distance_threshold) { level = i; break; }
    }
    // :303-305  (level as f32 + global_bias + mesh.lod_bias).clamp(0, len-1) as usize
    const uint32_t biased = clamp_to_index(fadd(fadd((float)level, P.cfg.global_bias), m.lod_bias), (float)last);
    if (biased < n) { tri_after = m.levels[biased].triangle_count; counted_after = true; stat_level = biased; }   // :308-313
    s.level = biased; s.blend = blend; s.culled = 0u; s.screen_coverage = cov;
    return s;
}

__global__ void __launch_bounds__(kInstThreads)
k_lod_select(const AshInstanceData* __restrict__ inst, const uint32_t* __restrict__ batch_id,
             const uint32_t* __restrict__ mesh_id, const AshLodMesh* __restrict__ meshes, uint32_t n_meshes,
             const uint32_t* __restrict__ visibility, uint32_t n, LodParams P, uint32_t n_buckets,
             AshLodSelection* __restrict__ sel, uint32_t* __restrict__ bucket_of, uint32_t* __restrict__ cta_hist,
             AshInstancingStats* __restrict__ stats) {
    extern __shared__ uint32_t s_hist[];                 // n_buckets
    __shared__ uint32_t s_lod[ASH_MAX_LOD_LEVELS + 1];   // objects_per_lod + culled
    __shared__ unsigned long long s_tri[2];
    for (uint32_t b = threadIdx.x; b < n_buckets; b += kInstThreads) s_hist[b] = 0u;
    if (threadIdx.x <= ASH_MAX_LOD_LEVELS) s_lod[threadIdx.x] = 0u;
    if (threadIdx.x < 2) s_tri[threadIdx.x] = 0ull;
    __syncthreads();
    const uint32_t i = blockIdx.x * kInstThreads + threadIdx.x;
    if (i < n) {
        uint32_t bucket = kNoBucket;
        AshLodSelection s = {ASH_LOD_NOT_VISIBLE, 0.0f, 0u, 0.0f};
        const bool vis = P.cfg.use_visibility ? (__ldg(visibility + i) != 0u) : true;
        const uint32_t mid = __ldg(mesh_id + i);
        if (vis && mid < n_meshes) {
            const AshLodMesh& m = meshes[mid];
            const float4 pos = __ldg(reinterpret_cast<const float4*>(inst[i].model_row3));     // position(), instancing.rs:68-70
            uint32_t stat_level, tri_after; bool counted;
            s = select_lod(m, pos.x, pos.y, pos.z, P, stat_level, tri_after, counted);
            if (m.level_count) atomicAdd(&s_tri[0], (unsigned long long)m.levels[0].triangle_count);   // :250-252
            if (counted) atomicAdd(&s_tri[1], (unsigned long long)tri_after);
            if (stat_level < ASH_MAX_LOD_LEVELS) atomicAdd(&s_lod[stat_level], 1u);
            if (s.culled) atomicAdd(&s_lod[ASH_MAX_LOD_LEVELS], 1u);
            else {
                const uint32_t b = __ldg(batch_id + i) * ASH_MAX_LOD_LEVELS + s.level;
                if (s.level < ASH_MAX_LOD_LEVELS && b < n_buckets) { bucket = b; atomicAdd(&s_hist[b], 1u); }
            }
        }
        sel[i] = s;
        bucket_of[i] = bucket;
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < n_buckets; b += kInstThreads) cta_hist[(size_t)blockIdx.x * n_buckets + b] = s_hist[b];
    if (threadIdx.x < ASH_MAX_LOD_LEVELS && s_lod[threadIdx.x]) atomicAdd(&stats->objects_per_lod[threadIdx.x], s_lod[threadIdx.x]);
    if (threadIdx.x == ASH_MAX_LOD_LEVELS && s_lod[ASH_MAX_LOD_LEVELS]) atomicAdd(&stats->objects_culled, s_lod[ASH_MAX_LOD_LEVELS]);
    if (threadIdx.x == 0 && s_tri[0]) atomicAdd(reinterpret_cast<unsigned long long*>(&stats->triangles_before), s_tri[0]);
    if (threadIdx.x == 1 && s_tri[1]) atomicAdd(reinterpret_cast<unsigned long long*>(&stats->triangles_after), s_tri[1]);
}

// Per-bucket exclusive scan over the CTAs: one warp per bucket, lanes stride over CTAs (n_ctas x n_buckets is
// small: 391 x 128 at 100 k objects / 16 batches).  cta_base[c][b] = slots taken by CTAs < c in bucket b.
__global__ void __launch_bounds__(256)
k_bucket_scan_ctas(const uint32_t* __restrict__ cta_hist, uint32_t n_ctas, uint32_t n_buckets,
                   uint32_t* __restrict__ cta_base, uint32_t* __restrict__ totals) {
    const uint32_t b = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    if (b >= n_buckets) return;
    uint32_t running = 0;
    for (uint32_t c0 = 0; c0 < n_ctas; c0 += 32u) {
        const uint32_t c = c0 + lane;
        const uint32_t v = c < n_ctas ? cta_hist[(size_t)c * n_buckets + b] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if ((int)lane >= d) incl += t;
        }
        if (c < n_ctas) cta_base[(size_t)c * n_buckets + b] = running + incl - v;
        running += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    if (lane == 0) totals[b] = running;
}

// One CTA, thread b = bucket b (n_buckets <= 1024): bucket offsets (kept = min(total, MAX_INSTANCES_PER_BATCH)),
// one indirect draw per bucket {index_count, instance_count, first_index, vertex_offset = 0, first_instance},
// and InstancingManager::finalize's statistics.
__global__ void __launch_bounds__(1024)
k_bucket_offsets(const uint32_t* __restrict__ totals, uint32_t n_buckets, const uint32_t* __restrict__ batch_mesh,
                 const AshLodMesh* __restrict__ meshes, uint32_t n_meshes, uint32_t* __restrict__ offsets,
                 AshIndirectDrawCommand* __restrict__ draws, AshInstancingStats* __restrict__ stats) {
    __shared__ uint32_t s_scan[1024];
    __shared__ uint32_t s_nonempty, s_dropped;
    const uint32_t b = threadIdx.x;
    if (b == 0) { s_nonempty = 0u; s_dropped = 0u; }
    const uint32_t total = b < n_buckets ? totals[b] : 0u;
    const uint32_t kept = total < ASH_MAX_INSTANCES_PER_BATCH ? total : ASH_MAX_INSTANCES_PER_BATCH;
    s_scan[b] = kept;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {            // Hillis-Steele inclusive scan
        const uint32_t v = b >= d ? s_scan[b - d] : 0u;
        __syncthreads();
        s_scan[b] += v;
        __syncthreads();
    }
    if (b < n_buckets) {
        const uint32_t off = s_scan[b] - kept;
        offsets[b] = off;
        if (kept) atomicAdd(&s_nonempty, 1u);
        if (total > kept) atomicAdd(&s_dropped, total - kept);
        const uint32_t batch = b / ASH_MAX_LOD_LEVELS, level = b % ASH_MAX_LOD_LEVELS;
        AshIndirectDrawCommand d = {0u, kept, 0u, 0, off};
        const uint32_t mid = batch_mesh[batch];
        if (mid < n_meshes && level < meshes[mid].level_count) {
            d.index_count = meshes[mid].levels[level].index_count;
            d.first_index = meshes[mid].levels[level].index_offset;
        }
        draws[b] = d;
    }
    if (b == n_buckets - 1u) offsets[n_buckets] = s_scan[b];
    __syncthreads();
    if (b == 0) {                                         // InstancingManager::finalize, instancing.rs:253-263
        const uint32_t all = s_scan[n_buckets - 1u];
        stats->total_instances = all;
        stats->batch_count = s_nonempty;
        stats->avg_instances_per_batch = s_nonempty ? fdiv((float)all, (float)s_nonempty) : 0.0f;
        stats->instances_dropped = s_dropped;
    }
}

__global__ void __launch_bounds__(kInstThreads)
k_instance_scatter(const AshInstanceData* __restrict__ inst, const uint32_t* __restrict__ bucket_of, uint32_t n,
                   const uint32_t* __restrict__ cta_base, uint32_t n_buckets, const uint32_t* __restrict__ offsets,
                   AshInstanceData* __restrict__ out) {
    extern __shared__ uint32_t s_warp[];                 // [8 warps][n_buckets]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t k = threadIdx.x; k < 8u * n_buckets; k += kInstThreads) s_warp[k] = 0u;
    __syncthreads();
    const uint32_t i = blockIdx.x * kInstThreads + threadIdx.x;
    const uint32_t bucket = i < n ? bucket_of[i] : kNoBucket;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, bucket);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    if (bucket != kNoBucket && rank == 0u) s_warp[warp * n_buckets + bucket] = __popc(peers);
    __syncthreads();
    if (bucket == kNoBucket) return;
    uint32_t slot = cta_base[(size_t)blockIdx.x * n_buckets + bucket] + rank;
    for (int w = 0; w < warp; ++w) slot += s_warp[w * n_buckets + bucket];
    if (slot >= ASH_MAX_INSTANCES_PER_BATCH) return;     // instancing.rs:219-223: the batch is full
    const uint4* src = reinterpret_cast<const uint4*>(inst + i);
    uint4* dst = reinterpret_cast<uint4*>(out + offsets[bucket] + slot);
    uint4 v[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) v[k] = __ldg(src + k);
#pragma unroll
    for (int k = 0; k < 6; ++k) dst[k] = v[k];
}

// batch -> mesh map: the mesh of the lowest-index object of each batch (a BatchKey names one mesh)
__global__ void k_batch_mesh(const uint32_t* __restrict__ batch_id, const uint32_t* __restrict__ mesh_id, uint32_t n,
                             uint32_t n_batches, uint32_t* __restrict__ first_obj) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t b = batch_id[i];
    if (b < n_batches) atomicMin(&first_obj[b], i);
}
__global__ void k_batch_mesh_resolve(const uint32_t* __restrict__ mesh_id, uint32_t n_batches, uint32_t* __restrict__ first_obj) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_batches) return;
    const uint32_t i = first_obj[b];
    first_obj[b] = i == 0xFFFFFFFFu ? 0xFFFFFFFFu : mesh_id[i];
}

cudaError_t launch_batch_mesh(const uint32_t* batch_id, const uint32_t* mesh_id, uint32_t n, uint32_t n_batches,
                              uint32_t* batch_mesh, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(batch_mesh, 0xFF, (size_t)n_batches * 4, s);
    if (e != cudaSuccess) return e;
    if (n) k_batch_mesh<<<(n + 255) / 256, 256, 0, s>>>(batch_id, mesh_id, n, n_batches, batch_mesh);
    k_batch_mesh_resolve<<<(n_batches + 127) / 128, 128, 0, s>>>(mesh_id, n_batches, batch_mesh);
    return cudaGetLastError();
}

size_t instancing_hist_bytes(uint32_t max_objects, uint32_t n_buckets) {
    return (size_t)((max_objects + kInstThreads - 1) / kInstThreads) * n_buckets * sizeof(uint32_t) * 2;   // counts + bases
}

static LodParams make_lod_params(const AshInstancingParams& ip, const AshLodConfig& cfg) {
    LodParams P;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) P.vp[c * 4 + r] = ip.view_proj[c][r];
    P.cam[0] = ip.camera_pos[0]; P.cam[1] = ip.camera_pos[1]; P.cam[2] = ip.camera_pos[2];
    P.sw = ip.screen_size[0]; P.sh = ip.screen_size[1];
    P.cfg = cfg;
    return P;
}

// graph replay: argument 7 of k_lod_select is the per-view LodParams
void lod_patch(const AshInstancingParams& ip, const AshLodConfig& cfg, PatchArg* out) {
    static_assert(sizeof(LodParams) <= sizeof(out->value), "PatchArg too small");
    const LodParams P = make_lod_params(ip, cfg);
    out->func = reinterpret_cast<const void*>(&k_lod_select);
    out->arg_index = 7; out->n_args = 13;
    std::memcpy(out->value, &P, sizeof(P));
}

cudaError_t launch_instancing(const AshInstanceData* inst, const uint32_t* batch_id, const uint32_t* mesh_id,
                              const AshLodMesh* meshes, uint32_t n_meshes, const uint32_t* batch_mesh,
                              const uint32_t* visibility, const AshInstancingParams& ip, const AshLodConfig& cfg,
                              uint32_t n_buckets, AshLodSelection* sel, uint32_t* bucket_of, uint32_t* cta_hist,
                              uint32_t* offsets, uint32_t* totals, AshInstanceData* out, AshIndirectDrawCommand* draws,
                              AshInstancingStats* stats, const Launch& L) {
    const uint32_t n = ip.object_count;
    const uint32_t n_ctas = (n + kInstThreads - 1) / kInstThreads;
    const LodParams P = make_lod_params(ip, cfg);
    Timed timed(L, ASH_KERNEL_INSTANCING);
    cudaError_t e = cudaMemsetAsync(stats, 0, sizeof(AshInstancingStats), L.s);
    if (e != cudaSuccess) return e;
    if (n_ctas) {
        k_lod_select<<<n_ctas, kInstThreads, n_buckets * sizeof(uint32_t), L.s>>>(inst, batch_id, mesh_id, meshes, n_meshes, visibility,
                                                                                   n, P, n_buckets, sel, bucket_of, cta_hist, stats);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    uint32_t* cta_base = cta_hist + (size_t)n_ctas * n_buckets;      // second half of the histogram allocation
    k_bucket_scan_ctas<<<(n_buckets + 7) / 8, 256, 0, L.s>>>(cta_hist, n_ctas, n_buckets, cta_base, totals);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    k_bucket_offsets<<<1, 1024, 0, L.s>>>(totals, n_buckets, batch_mesh, meshes, n_meshes, offsets, draws, stats);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (n_ctas) k_instance_scatter<<<n_ctas, kInstThreads, 8u * n_buckets * sizeof(uint32_t), L.s>>>(inst, bucket_of, n, cta_base, n_buckets, offsets, out);
    *L.launches += n_ctas ? 3u : 1u;                 // the Timed guard counted one of the (up to) four kernels
    return cudaGetLastError();
}

}  // namespace ash
