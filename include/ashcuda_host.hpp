// ashcuda_host.hpp — C++17 host-side mirror of the reference's Rust CPU state for the per-frame
// compute path, sitting directly above the C ABI (ashcuda.h).  Header-only.  Same names, argument
// meaning, defaults and error behaviour as the Rust items cited (paths relative to the reference):
//
//   CullBoundingBox / CullObjectData / OcclusionStats / OcclusionCulling   src/renderer/occlusion_culling.rs:26-272
//   GpuLight helpers / LightCullingPass                                     src/renderer/features/light_culling.rs:28-243
//   create_camera_data                                                      src/renderer/light_culling_integration.rs:112-123
//   HaltonSequence / TaaConfig / TemporalAA                                 src/renderer/temporal_aa.rs:14-207
//   BloomConfig / TonemappingConfig                                         features/bloom.rs:9-33, tonemapping.rs:22-44
//   RenderFeature / FeatureManager                                          features/feature_trait.rs:25-98
//   InstanceData helpers / BatchKey / InstancingStats / InstancingManager   src/renderer/instancing.rs:19-277  (row f-2)
//   LodLevel / LodMesh / LodConfig / LodSelection / LodStats / LodManager   src/renderer/lod_system.rs:15-394   (row f-2)
//   CudaComputeFeature — the plugin a maintainer registers; the Rust twin is rust-shim/src/feature.rs
//
// Matrices are `float[4][4]` with m[c] = column c (glam `to_cols_array_2d`).  The Rust toolchain is
// absent from the build image, so this is the compiled host layer that is actually built and tested
// (tests/cpp/test_host.cpp).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <map>
#include <typeindex>
#include <utility>
#include <vector>

#include "ashcuda.h"

namespace ashhost {

using Mat4 = float[4][4];

// AshError (src/error.rs:12-29) as an exception for the C++ layer; hooks of RenderFeature swallow
// and log it, exactly like the Rust hooks that return ().
struct AshError : std::runtime_error {
    int code;
    AshError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

// ---------------------------------------------------------------- occlusion culling
constexpr size_t MAX_CULLABLE_OBJECTS = 65536;   // occlusion_culling.rs:18
constexpr size_t HIZ_LEVELS = 10;                // occlusion_culling.rs:21

inline AshCullBoundingBox bounding_box_from_min_max(const float mn[3], const float mx[3]) {   // :34-42
    AshCullBoundingBox b{};
    for (int i = 0; i < 3; ++i) { b.center[i] = (mn[i] + mx[i]) * 0.5f; b.extents[i] = (mx[i] - mn[i]) * 0.5f; }
    return b;
}
inline AshCullBoundingBox bounding_box_from_center_extents(const float c[3], const float e[3]) {
    AshCullBoundingBox b{};
    for (int i = 0; i < 3; ++i) { b.center[i] = c[i]; b.extents[i] = e[i]; }
    return b;
}
inline AshCullObjectData cull_object_new(const AshCullBoundingBox& bounds, const Mat4 model, uint32_t draw_index) {   // :95-108
    AshCullObjectData o{};
    o.bounds = bounds;
    std::memcpy(o.model, model, sizeof(o.model));
    o.draw_index = draw_index; o.lod_bias = 0.0f; o.flags = 1; o._padding = 0;
    return o;
}

struct OcclusionStats {                            // :152-185
    uint32_t total_objects = 0, after_frustum = 0, after_occlusion = 0, draws_culled = 0;
    uint64_t triangles_culled = 0;
    double cull_rate() const { return total_objects == 0 ? 0.0 : 1.0 - (double)after_occlusion / (double)total_objects; }
    std::string format() const {
        char buf[160];
        std::snprintf(buf, sizeof buf, "Occlusion: %u/%u visible (%.1f%% culled), %u draws saved", after_occlusion,
                      total_objects, cull_rate() * 100.0, draws_culled);
        return buf;
    }
};

class OcclusionCulling {                           // :188-272
public:
    explicit OcclusionCulling(size_t max_objects = MAX_CULLABLE_OBJECTS, uint32_t hiz_levels = HIZ_LEVELS)
        : max_objects_(max_objects), hiz_levels_(hiz_levels) { objects_.reserve(1024); }
    void set_enabled(bool e) { enabled_ = e; }
    bool is_enabled() const { return enabled_; }
    void set_frustum_only(bool f) { frustum_only_ = f; }
    void begin_frame() { objects_.clear(); stats_ = OcclusionStats{}; }
    void add_object(const AshCullBoundingBox& bounds, const Mat4 model, uint32_t draw_index) {
        if (objects_.size() < max_objects_) objects_.push_back(cull_object_new(bounds, model, draw_index));
        else std::fprintf(stderr, "[warn] Occlusion culling: max object limit reached\n");
    }
    const std::vector<AshCullObjectData>& object_data() const { return objects_; }
    size_t object_count() const { return objects_.size(); }
    uint32_t hiz_levels() const { return hiz_levels_; }
    AshCullingPushConstants push_constants(const Mat4 view_proj, uint32_t width, uint32_t height) const {
        AshCullingPushConstants pc{};
        std::memcpy(pc.view_proj, view_proj, sizeof(pc.view_proj));
        pc.screen_params[0] = (float)width; pc.screen_params[1] = (float)height;
        pc.screen_params[2] = 1.0f / (float)width; pc.screen_params[3] = 1.0f / (float)height;
        pc.object_count = (uint32_t)objects_.size();
        pc.hiz_levels = hiz_levels_;
        return pc;
    }
    void update_stats(uint32_t visible_count) {
        stats_.total_objects = (uint32_t)objects_.size();
        stats_.after_occlusion = visible_count;
        stats_.draws_culled = stats_.total_objects > visible_count ? stats_.total_objects - visible_count : 0;
    }
    const OcclusionStats& stats() const { return stats_; }
private:
    bool enabled_ = true, frustum_only_ = false;
    size_t max_objects_;
    uint32_t hiz_levels_;
    std::vector<AshCullObjectData> objects_;
    OcclusionStats stats_;
};

// ---------------------------------------------------------------- light culling
constexpr size_t MAX_LIGHTS = 1024;                // features/light_culling.rs:17
constexpr uint32_t TILE_SIZE = 16;                 // :23

struct PointLight { float position[3] = {0, 0, 0}, color[3] = {1, 1, 1}, intensity = 1.0f, radius = 10.0f; };   // lighting.rs:22-39
struct DirectionalLight { float direction[3] = {0, -1, 0}, color[3] = {1, 1, 1}, intensity = 1.0f; };          // lighting.rs:6-20

inline AshGpuLight gpu_light_from_point_light(const PointLight& l) {              // :41-53
    return AshGpuLight{{l.position[0], l.position[1], l.position[2], l.radius},
                       {l.color[0], l.color[1], l.color[2], l.intensity}, {0, 0, 0, 0}, {0, 0, 1, 1}};
}
inline AshGpuLight gpu_light_from_directional_light(const DirectionalLight& l) {  // :56-63
    return AshGpuLight{{0, 0, 0, std::numeric_limits<float>::max()},
                       {l.color[0], l.color[1], l.color[2], l.intensity},
                       {l.direction[0], l.direction[1], l.direction[2], 2.0f}, {0, 0, 1, 1}};
}

class LightCullingPass {                           // :121-243
public:
    explicit LightCullingPass(size_t max_lights = MAX_LIGHTS) : max_lights_(max_lights) { lights_.reserve(max_lights); }
    bool enabled = true, debug_tiles = false;
    void update_lights(const std::vector<PointLight>& point, const std::vector<DirectionalLight>& directional) {
        lights_.clear();
        for (const auto& l : point) {
            if (lights_.size() >= max_lights_) { std::fprintf(stderr, "[warn] Light culling: exceeded max lights (%zu)\n", max_lights_); break; }
            lights_.push_back(gpu_light_from_point_light(l));
        }
        for (const auto& l : directional) {
            if (lights_.size() >= max_lights_) break;
            lights_.push_back(gpu_light_from_directional_light(l));
        }
    }
    void calculate_tiles(uint32_t width, uint32_t height) {
        if (width != last_w_ || height != last_h_) {
            tiles_x_ = (width + TILE_SIZE - 1) / TILE_SIZE; tiles_y_ = (height + TILE_SIZE - 1) / TILE_SIZE;
            last_w_ = width; last_h_ = height;
        }
    }
    void get_dispatch_dimensions(uint32_t out[3]) const { out[0] = tiles_x_; out[1] = tiles_y_; out[2] = 1; }
    AshLightCullingPushConstants get_push_constants(uint32_t w, uint32_t h) const {
        return AshLightCullingPushConstants{{w, h}, (uint32_t)lights_.size(), 0};
    }
    const std::vector<AshGpuLight>& get_light_buffer_data() const { return lights_; }
    size_t light_count() const { return lights_.size(); }
    size_t get_tile_buffer_size() const { return (size_t)tiles_x_ * tiles_y_ * (ASH_MAX_LIGHTS_PER_TILE + 1) * sizeof(uint32_t); }
    bool is_enabled() const { return enabled && !lights_.empty(); }
    uint32_t tiles_x() const { return tiles_x_; }
    uint32_t tiles_y() const { return tiles_y_; }
private:
    size_t max_lights_;
    std::vector<AshGpuLight> lights_;
    uint32_t tiles_x_ = 0, tiles_y_ = 0, last_w_ = 0, last_h_ = 0;
};

// ---------------------------------------------------------------- small column-major matrix helpers
inline void mat4_mul(const Mat4 a, const Mat4 b, Mat4 out) {       // out = a * b (glam Mat4 * Mat4)
    float r[4][4];
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            double s = 0.0;
            for (int k = 0; k < 4; ++k) s += (double)a[k][row] * (double)b[c][k];
            r[c][row] = (float)s;
        }
    std::memcpy(out, r, sizeof r);
}
inline bool mat4_inverse(const Mat4 m, Mat4 out) {                  // Gauss-Jordan in double, rounded once
    double a[4][8];
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) { a[r][c] = m[c][r]; a[r][c + 4] = (r == c) ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int p = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[p][i])) p = r;
        if (a[p][i] == 0.0) return false;
        if (p != i) for (int c = 0; c < 8; ++c) std::swap(a[p][c], a[i][c]);
        const double d = a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] /= d;
        for (int r = 0; r < 4; ++r) if (r != i) { const double f = a[r][i]; for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[c][r] = (float)a[r][c + 4];
    return true;
}

// LightCullingIntegration::create_camera_data (light_culling_integration.rs:112-123)
inline AshCullingCameraData create_camera_data(const Mat4 view, const Mat4 projection, const float camera_pos[3]) {
    AshCullingCameraData cd{};
    std::memcpy(cd.view, view, sizeof cd.view);
    std::memcpy(cd.projection, projection, sizeof cd.projection);
    mat4_inverse(projection, cd.inv_projection);
    cd.camera_pos[0] = camera_pos[0]; cd.camera_pos[1] = camera_pos[1]; cd.camera_pos[2] = camera_pos[2]; cd.camera_pos[3] = 1.0f;
    return cd;
}

// ---------------------------------------------------------------- TAA (temporal_aa.rs)
struct TaaConfig { bool enabled = true; float blend_factor = 0.9f; bool color_clamp = true, velocity_rejection = true; float jitter_scale = 1.0f; };

class HaltonSequence {                              // :42-83
public:
    static float halton(uint32_t index, uint32_t base) {
        float f = 1.0f, r = 0.0f;
        while (index > 0) { f /= (float)base; r += f * (float)(index % base); index /= base; }
        return r;
    }
    void next_jitter(float out[2]) {
        out[0] = halton(index_ + 1, 2) - 0.5f; out[1] = halton(index_ + 1, 3) - 0.5f;
        index_ = (index_ + 1) % 16;
    }
    void reset() { index_ = 0; }
    uint32_t index() const { return index_; }
private:
    uint32_t index_ = 0;
};

class TemporalAA {                                  // :111-207
public:
    explicit TemporalAA(TaaConfig cfg = TaaConfig{}) : config_(cfg) {}
    void begin_frame() {
        prev_[0] = cur_[0]; prev_[1] = cur_[1];
        float j[2]; halton_.next_jitter(j);
        cur_[0] = j[0] * config_.jitter_scale; cur_[1] = j[1] * config_.jitter_scale;
        ++frame_index_;
    }
    void jitter_projection(const Mat4 projection, uint32_t width, uint32_t height, Mat4 out) const {
        std::memcpy(out, projection, sizeof(Mat4));
        if (!config_.enabled) return;
        out[3][0] += cur_[0] * 2.0f / (float)width;        // w_axis.x
        out[3][1] += cur_[1] * 2.0f / (float)height;       // w_axis.y
    }
    AshTaaPushConstants push_constants(uint32_t width, uint32_t height) const {
        AshTaaPushConstants pc{};
        pc.screen_params[0] = (float)width; pc.screen_params[1] = (float)height;
        pc.screen_params[2] = 1.0f / (float)width; pc.screen_params[3] = 1.0f / (float)height;
        pc.params[0] = config_.blend_factor; pc.params[1] = config_.color_clamp ? 1.0f : 0.0f;
        pc.params[2] = config_.velocity_rejection ? 1.0f : 0.0f; pc.params[3] = 0.0f;
        pc.jitter[0] = cur_[0]; pc.jitter[1] = cur_[1]; pc.prev_jitter[0] = prev_[0]; pc.prev_jitter[1] = prev_[1];
        return pc;
    }
    const float* current_jitter() const { return cur_; }
    bool is_enabled() const { return config_.enabled; }
    TaaConfig& config_mut() { return config_; }
    const TaaConfig& config() const { return config_; }
    uint64_t frame_index() const { return frame_index_; }
    void reset_history() { halton_.reset(); cur_[0] = cur_[1] = prev_[0] = prev_[1] = 0.0f; }
private:
    TaaConfig config_;
    HaltonSequence halton_;
    float cur_[2] = {0, 0}, prev_[2] = {0, 0};
    uint64_t frame_index_ = 0;
};

// ---------------------------------------------------------------- bloom / tonemap configs
struct BloomConfig {                                // features/bloom.rs:9-33 (soft_knee is ignored by the shader, PIN-10)
    float threshold = 0.8f, intensity = 0.5f; uint32_t mip_count = 5; float soft_knee = 0.5f; bool enabled = true;
    AshBloomParams params() const { return AshBloomParams{threshold, 1.0f, mip_count, 0}; }
};
enum class TonemapOperator : uint32_t { Aces = ASH_TONEMAP_ACES, Reinhard = ASH_TONEMAP_REINHARD,   // features/tonemapping.rs:9-20
                                        Uncharted2 = ASH_TONEMAP_UNCHARTED2, None = ASH_TONEMAP_NONE };
struct TonemappingConfig {                          // features/tonemapping.rs:22-44 (Aces has the shader; the others: PIN-13, ashcuda.h)
    TonemapOperator op = TonemapOperator::Aces;
    float exposure = 1.0f, gamma = 2.2f;
    bool enabled = true;
    AshPostProcessPushConstants push_constants(float bloom_intensity) const { return AshPostProcessPushConstants{exposure, gamma, bloom_intensity, 0.0f}; }
};

// ---------------------------------------------------------------- plugin API (features/feature_trait.rs)
struct Device {};                                   // stands in for ash::Device: the CUDA backend needs nothing from it
struct FeatureFrameContext { const Device* device = nullptr; float elapsed_seconds = 0.0f; };   // :10-16 (fields the path can use)
struct FeatureRenderContext { const Device* device = nullptr; };                                 // :18-23

class RenderFeature {                               // :25-34
public:
    virtual ~RenderFeature() = default;
    virtual const char* name() const = 0;
    virtual void on_added(const Device&) {}
    virtual void before_frame(FeatureFrameContext&) {}
    virtual void render(const FeatureRenderContext&) {}
    virtual void on_removed(const Device&) {}
};

class FeatureManager {                              // :36-98: one instance per type, insertion order
public:
    void set_device(const Device* d) { device_ = d; }
    template <class F>
    F* add_feature(std::unique_ptr<F> f) {
        for (auto& e : features_) if (std::type_index(typeid(*e)) == std::type_index(typeid(F))) return nullptr;
        F* raw = f.get();
        if (device_) raw->on_added(*device_);
        features_.push_back(std::move(f));
        return raw;
    }
    void before_frame(FeatureFrameContext& ctx) { for (auto& f : features_) f->before_frame(ctx); }
    void render(const FeatureRenderContext& ctx) { for (auto& f : features_) f->render(ctx); }
    void cleanup() { if (device_) for (auto& f : features_) f->on_removed(*device_); features_.clear(); }
    size_t len() const { return features_.size(); }
private:
    std::vector<std::unique_ptr<RenderFeature>> features_;
    const Device* device_ = nullptr;
};

// ---------------------------------------------------------------- row f-2: instancing + LOD
constexpr size_t MAX_INSTANCES_PER_BATCH = ASH_MAX_INSTANCES_PER_BATCH;   // instancing.rs:16
constexpr size_t MAX_LOD_LEVELS = ASH_MAX_LOD_LEVELS;                     // lod_system.rs:15

inline AshInstanceData instance_data_from_matrix(const Mat4 model) {      // InstanceData::from_matrix, :37-48
    AshInstanceData d{};
    std::memcpy(d.model_row0, model[0], 16); std::memcpy(d.model_row1, model[1], 16);
    std::memcpy(d.model_row2, model[2], 16); std::memcpy(d.model_row3, model[3], 16);
    d.color[0] = d.color[1] = d.color[2] = d.color[3] = 1.0f;
    return d;
}
inline AshInstanceData instance_data_new(const Mat4 model, const float color[4]) {   // InstanceData::new, :50-61
    AshInstanceData d = instance_data_from_matrix(model);
    std::memcpy(d.color, color, 16);
    return d;
}
inline void instance_position(const AshInstanceData& d, float out[3]) { std::memcpy(out, d.model_row3, 12); }   // :68-70

struct BatchKey {                                    // :74-90
    uint32_t mesh_id = 0, material_id = 0;
    bool operator<(const BatchKey& o) const { return std::make_pair(mesh_id, material_id) < std::make_pair(o.mesh_id, o.material_id); }
    bool operator==(const BatchKey& o) const { return mesh_id == o.mesh_id && material_id == o.material_id; }
};

struct InstancingStats {                             // :155-186
    uint32_t total_instances = 0, batch_count = 0, instances_culled = 0;
    float avg_instances_per_batch = 0.0f;
    float efficiency() const { return batch_count == 0 ? 0.0f : avg_instances_per_batch / (float)MAX_INSTANCES_PER_BATCH; }
    std::string format() const {
        char b[160];
        std::snprintf(b, sizeof b, "Instancing: %u instances in %u batches (avg %.1f), %u culled", total_instances, batch_count,
                      avg_instances_per_batch, instances_culled);
        return b;
    }
};

// InstancingManager (:188-277).  Instances are kept in submission order with the dense id of their BatchKey
// (first-seen order) — the arrays ashcuda_set_instances consumes.
class InstancingManager {
public:
    void begin_frame() { for (auto& c : counts_) c = 0; instances_.clear(); batch_ids_.clear(); stats_ = InstancingStats{}; }   // :203-209
    void add_instance(const BatchKey& key, const AshInstanceData& instance) {                   // :212-224
        auto it = keys_.find(key);
        if (it == keys_.end()) { it = keys_.emplace(key, (uint32_t)keys_.size()).first; counts_.push_back(0); }
        const uint32_t b = it->second;
        if (counts_[b] < MAX_INSTANCES_PER_BATCH) {
            counts_[b] += 1; instances_.push_back(instance); batch_ids_.push_back(b); stats_.total_instances += 1;
        }
    }
    template <class It> void add_instances(const BatchKey& key, It first, It last) { for (; first != last; ++first) add_instance(key, *first); }   // :227-243
    void finalize() {                                                                           // :253-263
        stats_.batch_count = 0;
        for (size_t c : counts_) if (c) stats_.batch_count += 1;
        if (stats_.batch_count > 0) stats_.avg_instances_per_batch = (float)stats_.total_instances / (float)stats_.batch_count;
    }
    const InstancingStats& stats() const { return stats_; }
    void set_frustum_cull(bool e) { frustum_cull_ = e; }
    size_t batch_len(const BatchKey& key) const { auto it = keys_.find(key); return it == keys_.end() ? 0 : counts_[it->second]; }
    size_t key_count() const { return keys_.size(); }
    const std::vector<AshInstanceData>& instance_data() const { return instances_; }
    const std::vector<uint32_t>& batch_ids() const { return batch_ids_; }
private:
    std::map<BatchKey, uint32_t> keys_;
    std::vector<size_t> counts_;
    std::vector<AshInstanceData> instances_;
    std::vector<uint32_t> batch_ids_;
    InstancingStats stats_;
    bool frustum_cull_ = true;
};

using LodLevel = AshLodLevel;                        // lod_system.rs:61-76
using LodSelection = AshLodSelection;                // :126-137

struct LodMesh {                                     // :78-124
    std::vector<LodLevel> levels;
    float bounding_radius = 0.0f, lod_bias = 0.0f;
    static LodMesh single(uint32_t triangle_count, float bounding_radius) {                     // :91-104
        LodMesh m; m.bounding_radius = bounding_radius;
        m.levels.push_back(LodLevel{0.0f, std::numeric_limits<float>::max(), triangle_count, triangle_count, 0u, triangle_count * 3u});
        return m;
    }
    LodMesh& with_level(const LodLevel& l) { levels.push_back(l); return *this; }
    size_t level_count() const { return levels.size(); }
    AshLodMesh record() const {                      // levels flattened to MAX_LOD_LEVELS for the device
        if (levels.size() > MAX_LOD_LEVELS) throw AshError(ASH_ERR_INVALID, "LodMesh: more than MAX_LOD_LEVELS levels");
        AshLodMesh r{};
        for (size_t i = 0; i < levels.size(); ++i) r.levels[i] = levels[i];
        r.level_count = (uint32_t)levels.size(); r.bounding_radius = bounding_radius; r.lod_bias = lod_bias;
        return r;
    }
};

struct LodConfig {                                   // :35-58 (+ LodSelectionMode :17-33)
    uint32_t mode = ASH_LOD_MODE_SCREEN_SIZE, force_level = 0;
    bool smooth_transitions = true;
    float transition_width = 0.1f, global_bias = 0.0f, cull_threshold = 0.001f;
    AshLodConfig record(bool use_visibility = true) const {
        AshLodConfig r{};
        r.mode = mode; r.force_level = force_level; r.smooth_transitions = smooth_transitions ? 1u : 0u;
        r.transition_width = transition_width; r.global_bias = global_bias; r.cull_threshold = cull_threshold;
        r.use_visibility = use_visibility ? 1u : 0u;
        return r;
    }
};

struct LodStats {                                    // :146-183
    uint32_t objects_per_lod[ASH_MAX_LOD_LEVELS] = {};
    uint32_t objects_culled = 0;
    uint64_t triangles_before = 0, triangles_after = 0;
    double reduction_rate() const { return triangles_before == 0 ? 0.0 : 1.0 - (double)triangles_after / (double)triangles_before; }
    void reset() { *this = LodStats{}; }
};

// LodManager (:139-394): the per-object scalar version, fp32 in the reference's evaluation order (the translation
// unit must be built without FMA contraction, as tests/cpp does).  The device stage does the same in bulk.
class LodManager {
public:
    LodManager() = default;
    explicit LodManager(const LodConfig& c) : config_(c) {}
    void begin_frame() { stats_.reset(); }
    static float calculate_screen_coverage(const float position[3], float radius, const Mat4 view_proj, float screen_width,
                                           float screen_height) {                               // :207-235
        volatile float t0 = view_proj[0][3] * position[0], t1 = view_proj[1][3] * position[1], t2 = view_proj[2][3] * position[2];
        volatile float a = t0 + t1; volatile float b = a + t2; volatile float clip_w = b + view_proj[3][3] * 1.0f;
        if (clip_w <= 0.0f) return 0.0f;
        volatile float proj_radius = radius / clip_w;
        const float mx = rmax(screen_width, screen_height);
        volatile float pr = proj_radius * mx; volatile float pixel_radius = pr * 0.5f;
        volatile float ar = 3.14159274101257324f * pixel_radius; volatile float area = ar * pixel_radius;
        volatile float screen_area = screen_width * screen_height;
        volatile float cov = area / screen_area;
        return rmin(cov, 1.0f);
    }
    LodSelection select_lod(const LodMesh& mesh, const float world_position[3], const Mat4 view_proj, const float camera_position[3],
                            float screen_width, float screen_height) {                         // :238-345
        const size_t n = mesh.levels.size(), last = n ? n - 1 : 0;
        if (n) stats_.triangles_before += mesh.levels[0].triangle_count;
        if (config_.mode == ASH_LOD_MODE_FORCE) {
            const size_t level = config_.force_level < last ? config_.force_level : last;
            if (n) { stats_.triangles_after += mesh.levels[level].triangle_count; stats_.objects_per_lod[level] += 1; }
            return LodSelection{(uint32_t)level, 0.0f, 0u, 1.0f};
        }
        const float cov = calculate_screen_coverage(world_position, mesh.bounding_radius, view_proj, screen_width, screen_height);
        if (cov < config_.cull_threshold) { stats_.objects_culled += 1; return LodSelection{(uint32_t)last, 0.0f, 1u, cov}; }
        size_t level = last; float blend = 0.0f;
        if (config_.mode == ASH_LOD_MODE_SCREEN_SIZE) {
            for (size_t i = 0; i < n; ++i) {
                const float th = mesh.levels[i].screen_threshold;
                if (cov >= th) {
                    if (config_.smooth_transitions && i + 1 < n) {
                        volatile float range = th - mesh.levels[i + 1].screen_threshold;
                        if (range > 0.0f) {
                            volatile float d = th - cov; volatile float pos = d / range; volatile float q = pos / config_.transition_width;
                            float bl = q; if (bl < 0.0f) bl = 0.0f; if (bl > 1.0f) bl = 1.0f;
                            blend = bl;
                        }
                    }
                    level = i; break;
                }
            }
        } else {
            volatile float dx = world_position[0] - camera_position[0], dy = world_position[1] - camera_position[1], dz = world_position[2] - camera_position[2];
            volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz; volatile float s1 = xx + yy; volatile float s2 = s1 + zz;
            const float dist = std::sqrt((float)s2);
            for (size_t i = 0; i < n; ++i) if (dist <= mesh.levels[i].distance_threshold) { level = i; break; }
        }
        volatile float x0 = (float)level + config_.global_bias; volatile float x1 = x0 + mesh.lod_bias;
        float x = x1; if (x < 0.0f) x = 0.0f; if (x > (float)last) x = (float)last;
        const size_t biased = (x == x && x > 0.0f) ? (size_t)x : 0;
        if (biased < n) {
            stats_.triangles_after += mesh.levels[biased].triangle_count;
            if (biased < MAX_LOD_LEVELS) stats_.objects_per_lod[biased] += 1;
        }
        return LodSelection{(uint32_t)biased, blend, 0u, cov};
    }
    const LodStats& stats() const { return stats_; }
    LodConfig& config_mut() { return config_; }
private:
    static float rmax(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? b : a)); }   // f32::max
    static float rmin(float a, float b) { return (a != a) ? b : ((b != b) ? a : (b < a ? b : a)); }   // f32::min
    LodConfig config_;
    LodStats stats_;
};

// ---------------------------------------------------------------- the drop-in feature
// Per-frame inputs the reference's contexts do not carry (feature_trait.rs:10-23 has no camera,
// depth or colour): handed over through an additive setter before render_frame.
struct FrameInputs {
    Mat4 view{}, projection{};
    float camera_pos[3] = {0, 0, 0};
    const float* depth = nullptr;            // W*H fp32
    const void*  hdr = nullptr;              // W*H RGBA16F, or nullptr: no post chain this frame
    const void*  velocity = nullptr;         // W*H RG16F or nullptr
    int          on_device = 0;              // pointers are device pointers (borrowed)
    uint32_t     view_id = 0;
};

struct CullingStats { uint32_t objects_submitted = 0, after_frustum = 0, after_occlusion = 0, lights_culled = 0; uint64_t lod_triangles_saved = 0; };   // render_stats.rs:54-65
struct StatsCollector {                              // the culling part of render_stats.rs:132-240
    CullingStats culling;
    void record_culling(uint32_t submitted, uint32_t after_frustum, uint32_t after_occlusion) {      // :213-217
        culling.objects_submitted = submitted; culling.after_frustum = after_frustum; culling.after_occlusion = after_occlusion;
    }
    void record_lod_savings(uint64_t triangles_saved) { culling.lod_triangles_saved = triangles_saved; }   // :232-234
};

class CudaComputeFeature : public RenderFeature {
public:
    CudaComputeFeature(uint32_t width, uint32_t height, size_t max_objects = MAX_CULLABLE_OBJECTS,
                       size_t max_lights = MAX_LIGHTS, uint32_t hiz_levels = HIZ_LEVELS, int cuda_device = 0)
        : culling(max_objects, hiz_levels), lights(max_lights), width_(width), height_(height), device_index_(cuda_device) {
        ashcuda_default_config(&cfg_);
        cfg_.max_objects = (uint32_t)max_objects; cfg_.max_lights = (uint32_t)max_lights; cfg_.hiz_levels = hiz_levels;
    }
    ~CudaComputeFeature() override { ashcuda_destroy(ctx_); }
    CudaComputeFeature(const CudaComputeFeature&) = delete;
    CudaComputeFeature& operator=(const CudaComputeFeature&) = delete;

    // the reference's CPU state objects, public like the fields a Rust caller would reach through accessors
    OcclusionCulling  culling;
    LightCullingPass  lights;
    TemporalAA        taa;
    BloomConfig       bloom;
    TonemappingConfig tonemapping;
    std::vector<AshIndirectDrawCommand> draw_templates{AshIndirectDrawCommand{}};
    // row f-2 (optional): one instance + LodMesh id per cull object, in the same order as culling.add_object
    InstancingManager instancing;
    LodManager        lod;
    std::vector<LodMesh>  lod_meshes;
    std::vector<uint32_t> lod_mesh_ids;
    // row f-1 (optional): make the depth and HDR colour buffers on the device from the cull objects and the tile
    // light lists (shaders/frag.frag's BRDF) instead of taking them from FrameInputs
    bool             produce_inputs = false;
    AshForwardParams forward_params = default_forward_params();

    static AshForwardParams default_forward_params() {     // the library's defaults (ashcuda_set_forward_params(ctx, NULL))
        AshForwardParams p{};
        const float dir[4] = {-0.3f, -1.0f, -0.2f, 0.0f}, white[4] = {1, 1, 1, 1}, amb[4] = {0.03f, 0.03f, 0.03f, 1.0f};
        const float mr[4] = {0.0f, 0.5f, 1.0f, 1.0f}, bg[4] = {0.02f, 0.02f, 0.03f, 1.0f};
        std::memcpy(p.light_direction, dir, 16); std::memcpy(p.light_color, white, 16); std::memcpy(p.ambient_color, amb, 16);
        std::memcpy(p.base_color_factor, white, 16); std::memcpy(p.parameters, mr, 16); std::memcpy(p.background, bg, 16);
        p.near_plane = 0.1f; p.radius_scale = 1.0f; p.flags = ASH_FWD_TILE_LIGHTS;
        return p;
    }

    const char* name() const override { return "CudaCompute"; }
    void on_added(const Device&) override {               // LightCullingPipeline::new, :41-243
        cfg_.bloom_mips = bloom.mip_count;
        int rc = ashcuda_create(&ctx_, device_index_, width_, height_, &cfg_);
        if (rc != ASH_OK) { last_error_ = ashcuda_last_error(nullptr); last_code_ = rc; ctx_ = nullptr; log_error("on_added"); return; }
        lights.calculate_tiles(width_, height_);
        scene_dirty_ = true; last_code_ = ASH_OK;
    }
    void on_removed(const Device&) override { ashcuda_destroy(ctx_); ctx_ = nullptr; }
    void set_frame_inputs(const FrameInputs& in) { inputs_ = in; have_inputs_ = true; }
    void mark_scene_dirty() { scene_dirty_ = true; }

    void before_frame(FeatureFrameContext&) override {
        if (!ctx_) { fail(ASH_ERR_NOT_INITIALIZED, "CudaComputeFeature used before on_added (FeatureNotInitialized)"); return; }
        taa.begin_frame();
        if (scene_dirty_) {
            const auto& o = culling.object_data();
            if (!check(ashcuda_set_objects(ctx_, o.data(), (uint32_t)o.size(), draw_templates.data(), (uint32_t)draw_templates.size(), 0), "set_objects")) return;
            const auto& l = lights.get_light_buffer_data();
            if (!check(ashcuda_set_lights(ctx_, l.data(), (uint32_t)l.size(), 0), "set_lights")) return;
            instancing_on_ = false;
            if (!lod_meshes.empty() && !instancing.instance_data().empty() &&
                instancing.instance_data().size() == lod_mesh_ids.size() && lod_mesh_ids.size() == o.size()) {
                std::vector<AshLodMesh> recs;
                for (const LodMesh& m : lod_meshes) recs.push_back(m.record());
                const AshLodConfig k = lod.config_mut().record(true);
                if (!check(ashcuda_set_instances(ctx_, instancing.instance_data().data(), instancing.batch_ids().data(), lod_mesh_ids.data(),
                                                 (uint32_t)lod_mesh_ids.size(), 0), "set_instances")) return;
                if (!check(ashcuda_set_lod_meshes(ctx_, recs.data(), (uint32_t)recs.size(), 0), "set_lod_meshes")) return;
                if (!check(ashcuda_set_lod_config(ctx_, &k), "set_lod_config")) return;
                instancing_on_ = true;
            }
            scene_dirty_ = false;
        }
    }
    void render(const FeatureRenderContext&) override {
        if (!ctx_ || !have_inputs_) { fail(ASH_ERR_NOT_INITIALIZED, "render without context or frame inputs"); return; }
        if (produce_inputs) {
            if (!check(ashcuda_set_forward_params(ctx_, &forward_params), "set_forward_params")) return;
        } else {
            if (!check(ashcuda_set_depth(ctx_, inputs_.depth, inputs_.on_device), "set_depth")) return;
            if (inputs_.hdr && !check(ashcuda_set_color(ctx_, inputs_.hdr, inputs_.velocity, inputs_.on_device), "set_color")) return;
        }
        if (!check(ashcuda_set_tonemap_operator(ctx_, (uint32_t)tonemapping.op), "set_tonemap_operator")) return;
        const AshCudaViewParams vp = view_params();
        check(ashcuda_view_run(ctx_, &vp), "view_run");
    }
    AshCudaViewParams view_params() const {
        AshCudaViewParams vp{};
        Mat4 view_proj;
        mat4_mul(inputs_.projection, inputs_.view, view_proj);             // renderer.rs:1645 projection * view
        vp.cull = culling.push_constants(view_proj, width_, height_);
        const uint32_t built = ctx_ ? (uint32_t)ashcuda_hiz_level_count(ctx_) : culling.hiz_levels();
        if (vp.cull.hiz_levels > built) vp.cull.hiz_levels = built;
        vp.light = lights.get_push_constants(width_, height_);
        vp.camera = create_camera_data(inputs_.view, inputs_.projection, inputs_.camera_pos);
        vp.taa = taa.push_constants(width_, height_);
        vp.bloom = bloom.params();
        vp.post = tonemapping.push_constants(bloom.enabled ? bloom.intensity : 0.0f);
        vp.view_id = inputs_.view_id;
        uint32_t st = ASH_STAGE_HIZ;
        if (culling.is_enabled()) st |= ASH_STAGE_CULL;
        if (lights.is_enabled()) st |= ASH_STAGE_LIGHTS;
        if (instancing_on_ && culling.is_enabled()) st |= ASH_STAGE_INSTANCING;
        if (produce_inputs) st |= ASH_STAGE_FORWARD_DEPTH | ASH_STAGE_FORWARD_SHADE | ASH_STAGE_LIGHTS;
        if (inputs_.hdr || produce_inputs) {
            if (taa.is_enabled()) st |= ASH_STAGE_TAA;
            if (bloom.enabled) st |= ASH_STAGE_BLOOM;
            if (tonemapping.enabled) st |= ASH_STAGE_TONEMAP;
        }
        vp.stages = st;
        return vp;
    }
    // results: syncs the stream (ashcuda_readback); updates OcclusionStats like update_stats(visible)
    uint32_t read_visible_count() {
        uint32_t n = 0;
        if (ctx_ && check(ashcuda_readback(ctx_, ASH_BUF_VISIBLE_COUNT, &n, 4), "readback")) culling.update_stats(n);
        return n;
    }
    // CullingStats through the reference's recorders (render_stats.rs:213-234); syncs the stream
    bool read_culling_stats(StatsCollector* sc) {
        AshCullingStats cs{};
        if (!ctx_ || !check(ashcuda_readback(ctx_, ASH_BUF_CULLING_STATS, &cs, sizeof(cs)), "readback")) return false;
        sc->record_culling(cs.objects_submitted, cs.after_frustum, cs.after_occlusion);
        sc->record_lod_savings(cs.lod_triangles_saved);
        sc->culling.lights_culled = cs.lights_culled;
        return true;
    }
    // LodStats + InstancingStats of the last frame (syncs the stream)
    bool read_instancing_stats(AshInstancingStats* out) { return ctx_ && check(ashcuda_readback(ctx_, ASH_BUF_INSTANCING_STATS, out, sizeof(*out)), "readback"); }
    ashcuda_ctx* context() const { return ctx_; }
    int last_code() const { return last_code_; }
    const std::string& last_error() const { return last_error_; }

private:
    bool check(int rc, const char* what) {
        if (rc == ASH_OK) { last_code_ = ASH_OK; return true; }      // last_code() = status of the latest call
        last_code_ = rc; last_error_ = ashcuda_last_error(ctx_); log_error(what);
        return false;
    }
    void fail(int code, const char* msg) { last_code_ = code; last_error_ = msg; log_error("feature"); }
    void log_error(const char* what) const { std::fprintf(stderr, "[error] CudaComputeFeature %s: %s (%d)\n", what, last_error_.c_str(), last_code_); }

    uint32_t width_, height_;
    int device_index_;
    AshCudaConfig cfg_{};
    ashcuda_ctx* ctx_ = nullptr;
    FrameInputs inputs_{};
    bool have_inputs_ = false, scene_dirty_ = true, instancing_on_ = false;
    int last_code_ = ASH_OK;
    std::string last_error_;
};

}  // namespace ashhost
