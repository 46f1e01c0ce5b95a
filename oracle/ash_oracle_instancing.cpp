// ash_oracle_instancing.cpp — CPU ORACLE (test infrastructure, NOT product code) for SURVEY §8.6 row f-2.
//
// A serial C++ restatement of the reference's CPU managers, object by object in ascending index:
//   LodManager::calculate_screen_coverage   src/renderer/lod_system.rs:207-235
//   LodManager::select_lod                  src/renderer/lod_system.rs:238-345
//   LodManager::select_by_screen_size / _distance                    :347-378
//   InstancingManager::add_instance         src/renderer/instancing.rs:212-224
//   InstancingManager::finalize             src/renderer/instancing.rs:253-263
// Unlike the shader oracle this part restates real CPU code, and it is PINNED by replaying the reference's
// own unit tests (instancing.rs:283-333, lod_system.rs:398-428) in tests/test_oracle_instancing.py.  Those
// tests are few and loose (a coverage "in (0,1)", counts), so the arithmetic order is additionally
// fixed by glam 0.30 (Cargo.toml:24; not vendored in the reference tree), whose published scalar / SSE2
// code evaluates Mat4 * Vec4 as ((x_axis*v.x + y_axis*v.y) + z_axis*v.z) + w_axis*v.w and Vec3::length
// as sqrt((x*x + y*y) + z*z), without FMA on the default x86-64 target.
//
// The composition of the two managers is ours (the reference never calls them together): a (batch, LOD
// level) pair is one BatchKey, i.e. one draw call with that level's index range.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../include/ashcuda.h"

namespace {

// Rust f32::max / f32::min: if one operand is NaN the other is returned
inline float rmax(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? b : a)); }
inline float rmin(float a, float b) { return (a != a) ? b : ((b != b) ? a : (b < a ? b : a)); }
// Rust f32::clamp (NaN stays NaN)
inline float rclamp(float x, float lo, float hi) { if (x < lo) x = lo; if (x > hi) x = hi; return x; }
// Rust `f as usize`: truncation, saturating, NaN -> 0
inline uint32_t as_usize(float x) {
    if (!(x == x) || x <= 0.0f) return 0u;
    if (x >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)x;
}

float screen_coverage(const float vp[4][4], const float pos[3], float radius, float sw, float sh) {   // :207-235
    // clip = view_proj * position.extend(1.0); only clip.w is used (the ndc at :222 is dead)
    const float clip_w = ((vp[0][3] * pos[0] + vp[1][3] * pos[1]) + vp[2][3] * pos[2]) + vp[3][3] * 1.0f;
    if (clip_w <= 0.0f) return 0.0f;
    const float dist = clip_w;
    const float proj_radius = radius / dist;
    const float pixel_radius = proj_radius * rmax(sw, sh) * 0.5f;
    const float area = 3.14159274101257324f * pixel_radius * pixel_radius;      // std::f32::consts::PI
    const float screen_area = sw * sh;
    return rmin(area / screen_area, 1.0f);
}

struct LodStats {                     // lod_system.rs:146-157
    uint32_t objects_per_lod[ASH_MAX_LOD_LEVELS] = {};
    uint32_t objects_culled = 0;
    uint64_t triangles_before = 0, triangles_after = 0;
};

AshLodSelection select_lod(LodStats& st, const AshLodConfig& cfg, const AshLodMesh& mesh, const float pos[3],
                           const float vp[4][4], const float cam[3], float sw, float sh) {            // :238-345
    const uint32_t n = mesh.level_count;
    const uint32_t last = n ? n - 1u : 0u;                                    // len().saturating_sub(1)
    if (n) st.triangles_before += mesh.levels[0].triangle_count;              // :250-252
    if (cfg.mode == ASH_LOD_MODE_FORCE) {                                     // :255-268
        const uint32_t level = cfg.force_level < last ? cfg.force_level : last;
        if (n) { st.triangles_after += mesh.levels[level].triangle_count; st.objects_per_lod[level] += 1; }
        return {level, 0.0f, 0u, 1.0f};
    }
    const float cov = screen_coverage(vp, pos, mesh.bounding_radius, sw, sh); // :271-277
    if (cov < cfg.cull_threshold) {                                           // :280-288
        st.objects_culled += 1;
        return {last, 0.0f, 1u, cov};
    }
    uint32_t level = last; float blend = 0.0f;
    if (cfg.mode == ASH_LOD_MODE_SCREEN_SIZE) {                               // :347-369
        for (uint32_t i = 0; i < n; ++i) {
            const AshLodLevel& l = mesh.levels[i];
            if (cov >= l.screen_threshold) {
                if (cfg.smooth_transitions && i + 1 < n) {
                    const float next_threshold = mesh.levels[i + 1].screen_threshold;
                    const float range = l.screen_threshold - next_threshold;
                    if (range > 0.0f) {
                        const float pos_in_range = (l.screen_threshold - cov) / range;
                        blend = rclamp(pos_in_range / cfg.transition_width, 0.0f, 1.0f);
                    }
                }
                level = i;
                break;
            }
        }
    } else {                                                                  // :371-378
        const float d[3] = {pos[0] - cam[0], pos[1] - cam[1], pos[2] - cam[2]};
        const float distance = sqrtf((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]);
        for (uint32_t i = 0; i < n; ++i)
            if (distance <= mesh.levels[i].distance_threshold) { level = i; break; }
    }
    // :303-305
    const uint32_t biased = as_usize(rclamp((float)level + cfg.global_bias + mesh.lod_bias, 0.0f, (float)last));
    if (biased < n) {                                                         // :308-313
        st.triangles_after += mesh.levels[biased].triangle_count;
        if (biased < ASH_MAX_LOD_LEVELS) st.objects_per_lod[biased] += 1;
    }
    return {biased, blend, 0u, cov};
}

}  // namespace

extern "C" {

float ash_oracle_screen_coverage(const float* vp16, const float* pos3, float radius, float sw, float sh) {
    float vp[4][4];
    std::memcpy(vp, vp16, sizeof(vp));
    return screen_coverage(vp, pos3, radius, sw, sh);
}

// Whole frame: for i in 0..object_count (ascending): if visible: select_lod; if !culled: add_instance(key(batch, level)).
// Outputs as the CUDA path lays them out: sel[N]; out = bucket streams back to back (bucket = batch*8 + level);
// offsets[n_buckets+1]; draws[n_buckets]; stats.
void ash_oracle_instancing(const AshInstanceData* inst, const uint32_t* batch_id, const uint32_t* mesh_id,
                           const AshLodMesh* meshes, uint32_t n_meshes, const uint32_t* visibility_or_null,
                           const AshInstancingParams* ip, const AshLodConfig* cfg, uint32_t n_batches,
                           AshLodSelection* sel, AshInstanceData* out, uint32_t* offsets,
                           AshIndirectDrawCommand* draws, AshInstancingStats* stats) {
    const uint32_t n_buckets = n_batches * ASH_MAX_LOD_LEVELS;
    std::vector<std::vector<uint32_t>> batches(n_buckets);       // InstancingManager::batches, by dense key
    std::vector<uint32_t> batch_mesh(n_batches, 0xFFFFFFFFu);
    LodStats lod;
    uint32_t total_instances = 0, dropped = 0;
    for (uint32_t i = 0; i < ip->object_count; ++i) {
        if (batch_id[i] < n_batches && batch_mesh[batch_id[i]] == 0xFFFFFFFFu) batch_mesh[batch_id[i]] = mesh_id[i];
        sel[i] = {ASH_LOD_NOT_VISIBLE, 0.0f, 0u, 0.0f};
        const bool vis = (cfg->use_visibility && visibility_or_null) ? visibility_or_null[i] != 0u : true;
        if (!vis || mesh_id[i] >= n_meshes) continue;
        const AshLodMesh& m = meshes[mesh_id[i]];
        const float pos[3] = {inst[i].model_row3[0], inst[i].model_row3[1], inst[i].model_row3[2]};   // position(), instancing.rs:68-70
        sel[i] = select_lod(lod, *cfg, m, pos, ip->view_proj, ip->camera_pos, ip->screen_size[0], ip->screen_size[1]);
        if (sel[i].culled || sel[i].level >= ASH_MAX_LOD_LEVELS || batch_id[i] >= n_batches) continue;
        std::vector<uint32_t>& b = batches[batch_id[i] * ASH_MAX_LOD_LEVELS + sel[i].level];
        if (b.size() < ASH_MAX_INSTANCES_PER_BATCH) { b.push_back(i); total_instances += 1; }      // add_instance, :219-223
        else dropped += 1;
    }
    // the batch -> mesh map looks at every instance that was set, not only the first object_count
    // (the CUDA path builds it once in ashcuda_set_instances); callers pass object_count == n here.
    uint32_t off = 0, batch_count = 0;
    for (uint32_t b = 0; b < n_buckets; ++b) {
        offsets[b] = off;
        for (uint32_t i : batches[b]) out[off++] = inst[i];
        if (!batches[b].empty()) batch_count += 1;                                                  // finalize: retain non-empty, :255
        AshIndirectDrawCommand d = {0u, (uint32_t)batches[b].size(), 0u, 0, offsets[b]};
        const uint32_t mid = batch_mesh[b / ASH_MAX_LOD_LEVELS], level = b % ASH_MAX_LOD_LEVELS;
        if (mid < n_meshes && level < meshes[mid].level_count) {
            d.index_count = meshes[mid].levels[level].index_count;
            d.first_index = meshes[mid].levels[level].index_offset;
        }
        draws[b] = d;
    }
    offsets[n_buckets] = off;
    std::memset(stats, 0, sizeof(*stats));
    std::memcpy(stats->objects_per_lod, lod.objects_per_lod, sizeof(lod.objects_per_lod));
    stats->objects_culled = lod.objects_culled;
    stats->triangles_before = lod.triangles_before;
    stats->triangles_after = lod.triangles_after;
    stats->total_instances = total_instances;
    stats->batch_count = batch_count;
    stats->avg_instances_per_batch = batch_count ? (float)total_instances / (float)batch_count : 0.0f;   // :258-262
    stats->instances_dropped = dropped;
}

}  // extern "C"
