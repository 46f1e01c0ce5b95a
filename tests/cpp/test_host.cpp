// C++ host-layer tests (include/ashcuda_host.hpp).  Cases follow the reference's in-file unit tests:
// occlusion_culling.rs:290-314, features/light_culling.rs:257-276, light_culling_integration.rs:216-231,
// temporal_aa.rs:220-238, instancing.rs:279-334, lod_system.rs:396-428, plus FeatureManager ordering
// (feature_trait.rs:57-97).
// Usage: test_host          -> host logic + "no CPU fallback" check (exit 0 on success)
//        test_host --gpu    -> additionally runs one small view through CudaComputeFeature on cuda:0
#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>
#include "ashcuda_host.hpp"

using namespace ashhost;
#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static std::vector<const char*> g_log;
struct FA : RenderFeature { const char* name() const override { return "A"; }
    void on_added(const Device&) override { g_log.push_back("A.added"); } void before_frame(FeatureFrameContext&) override { g_log.push_back("A.before"); }
    void render(const FeatureRenderContext&) override { g_log.push_back("A.render"); } void on_removed(const Device&) override { g_log.push_back("A.removed"); } };
struct FB : RenderFeature { const char* name() const override { return "B"; }
    void on_added(const Device&) override { g_log.push_back("B.added"); } void before_frame(FeatureFrameContext&) override { g_log.push_back("B.before"); }
    void render(const FeatureRenderContext&) override { g_log.push_back("B.render"); } void on_removed(const Device&) override { g_log.push_back("B.removed"); } };

static void identity(Mat4 m) { std::memset(m, 0, sizeof(Mat4)); for (int i = 0; i < 4; ++i) m[i][i] = 1.0f; }
static void perspective_rh(float fovy, float aspect, float n, float f, Mat4 m) {     // glam perspective_rh + Y flip (02_cube.rs:87-91)
    std::memset(m, 0, sizeof(Mat4));
    const float h = std::cos(0.5f * fovy) / std::sin(0.5f * fovy), w = h / aspect, r = f / (n - f);
    m[0][0] = w; m[1][1] = -h; m[2][2] = r; m[2][3] = -1.0f; m[3][2] = r * n;
}

int main(int argc, char** argv) {
    const bool gpu = argc > 1 && std::strcmp(argv[1], "--gpu") == 0;
    static_assert(sizeof(AshCullObjectData) % 16 == 0, "occlusion_culling.rs:311-314");
    {   // from_min_max, cull_rate
        const float mn[3] = {-1, -1, -1}, mx[3] = {1, 1, 1};
        const AshCullBoundingBox b = bounding_box_from_min_max(mn, mx);
        CHECK(b.center[0] == 0 && b.center[1] == 0 && b.center[2] == 0 && b.extents[0] == 1 && b.extents[2] == 1);
        OcclusionStats s; s.total_objects = 100; s.after_occlusion = 30;
        CHECK(std::fabs(s.cull_rate() - 0.7) < 1e-9 && OcclusionStats{}.cull_rate() == 0.0);
    }
    {   // OcclusionCulling state
        OcclusionCulling oc(3);
        Mat4 I; identity(I);
        const float c[3] = {0, 0, 0}, e[3] = {0.5f, 0.5f, 0.5f};
        for (uint32_t i = 0; i < 5; ++i) oc.add_object(bounding_box_from_center_extents(c, e), I, i);
        CHECK(oc.object_count() == 3 && oc.object_data()[2].draw_index == 2 && oc.object_data()[0].flags == 1);
        const AshCullingPushConstants pc = oc.push_constants(I, 1920, 1080);
        CHECK(pc.object_count == 3 && pc.hiz_levels == 10 && pc.screen_params[2] == 1.0f / 1920.0f);
        oc.update_stats(1); CHECK(oc.stats().draws_culled == 2);
        oc.begin_frame(); CHECK(oc.object_count() == 0);
    }
    {   // lights
        PointLight p; p.position[0] = 1; p.position[1] = 2; p.position[2] = 3; p.radius = 10; p.intensity = 2;
        const AshGpuLight g = gpu_light_from_point_light(p);
        CHECK(g.position[3] == 10.0f && g.color[3] == 2.0f && g.params[3] == 1.0f);
        CHECK(gpu_light_from_directional_light(DirectionalLight{}).direction[3] == 2.0f);
        LightCullingPass pass; CHECK(!pass.is_enabled());
        pass.calculate_tiles(1920, 1080); CHECK(pass.tiles_x() == 120 && pass.tiles_y() == 68 && pass.get_tile_buffer_size() == 8388480u);
        pass.update_lights(std::vector<PointLight>(2000), {DirectionalLight{}}); CHECK(pass.light_count() == 1024 && pass.is_enabled());
        Mat4 I, P; identity(I); perspective_rh(0.7853982f, 16.0f / 9.0f, 0.5f, 100.0f, P);
        const float cam[3] = {1, 2, 3};
        const AshCullingCameraData cd = create_camera_data(I, P, cam);
        CHECK(cd.camera_pos[0] == 1 && cd.camera_pos[2] == 3 && cd.camera_pos[3] == 1);
        Mat4 prod; mat4_mul(P, cd.inv_projection, prod);
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) CHECK(std::fabs(prod[i][j] - (i == j ? 1.0f : 0.0f)) < 1e-4f);
    }
    {   // TAA
        TemporalAA taa; Mat4 I, J; identity(I);
        taa.jitter_projection(I, 1920, 1080, J); CHECK(std::memcmp(I, J, sizeof(Mat4)) == 0);
        taa.begin_frame(); const float a0 = taa.current_jitter()[0], a1 = taa.current_jitter()[1];
        taa.begin_frame(); const float b0 = taa.current_jitter()[0], b1 = taa.current_jitter()[1];
        CHECK((a0 != b0 || a1 != b1) && std::fabs(a0) <= 0.5f && std::fabs(b1) <= 0.5f && a0 == 0.0f);
        CHECK(HaltonSequence::halton(1, 2) == 0.5f && HaltonSequence::halton(2, 2) == 0.25f && HaltonSequence::halton(3, 2) == 0.75f);
        HaltonSequence h; float first[2], j[2]; h.next_jitter(first); for (int i = 1; i < 16; ++i) h.next_jitter(j); h.next_jitter(j);
        CHECK(j[0] == first[0] && j[1] == first[1]);                               // period 16
        const AshTaaPushConstants pc = taa.push_constants(1920, 1080);
        CHECK(pc.params[0] == 0.9f && pc.params[1] == 1.0f && pc.jitter[0] == b0 && pc.prev_jitter[1] == a1);
    }
    {   // FeatureManager
        Device dev; FeatureManager fm;
        fm.add_feature(std::make_unique<FA>()); fm.set_device(&dev); fm.add_feature(std::make_unique<FB>());
        CHECK(fm.add_feature(std::make_unique<FA>()) == nullptr && fm.len() == 2);
        FeatureFrameContext fc; FeatureRenderContext rc; fm.before_frame(fc); fm.render(rc); fm.cleanup();
        const char* want[] = {"B.added", "A.before", "B.before", "A.render", "B.render", "A.removed", "B.removed"};
        CHECK(g_log.size() == 7); for (int i = 0; i < 7; ++i) CHECK(std::strcmp(g_log[i], want[i]) == 0);
    }
    {   // instancing.rs:283-320 test_instance_data / test_batching / test_multiple_batches
        Mat4 m; identity(m); m[3][0] = 1.0f; m[3][1] = 2.0f; m[3][2] = 3.0f;                    // Mat4::from_translation
        float p[3]; instance_position(instance_data_from_matrix(m), p);
        CHECK(p[0] == 1.0f && p[1] == 2.0f && p[2] == 3.0f && sizeof(AshInstanceData) == 96);
        InstancingManager mgr; mgr.begin_frame();
        const BatchKey key{1, 1};
        for (int i = 0; i < 100; ++i) { Mat4 t; identity(t); t[3][0] = (float)i; mgr.add_instance(key, instance_data_from_matrix(t)); }
        mgr.finalize();
        CHECK(mgr.stats().total_instances == 100 && mgr.stats().batch_count == 1 && mgr.stats().avg_instances_per_batch == 100.0f);
        InstancingManager m2; m2.begin_frame();
        for (uint32_t mesh = 0; mesh < 5; ++mesh) m2.add_instance(BatchKey{mesh, 0}, AshInstanceData{});
        m2.finalize();
        CHECK(m2.stats().batch_count == 5 && m2.batch_ids().size() == 5 && m2.batch_ids()[4] == 4);
        m2.begin_frame(); m2.add_instance(BatchKey{3, 0}, AshInstanceData{}); m2.finalize();
        CHECK(m2.stats().batch_count == 1 && m2.batch_ids()[0] == 3);                            // keys keep their dense ids
    }
    {   // lod_system.rs:400-425 test_single_lod / test_screen_coverage / test_reduction_rate
        const LodMesh mesh = LodMesh::single(1000, 1.0f);
        CHECK(mesh.level_count() == 1 && mesh.levels[0].index_count == 3000 && mesh.record().level_count == 1);
        Mat4 proj, view, vp; perspective_rh(0.7853982f, 16.0f / 9.0f, 0.1f, 100.0f, proj); proj[1][1] = -proj[1][1];   // no Y flip here
        identity(view); view[3][2] = -10.0f;                                                      // look_at_rh((0,0,10), 0, Y)
        mat4_mul(proj, view, vp);
        const float origin[3] = {0, 0, 0};
        const float cov = LodManager::calculate_screen_coverage(origin, 1.0f, vp, 1920.0f, 1080.0f);
        CHECK(cov > 0.0f && cov < 1.0f && std::fabs(cov - 3.14159265f * 96.0f * 96.0f / (1920.0f * 1080.0f)) < 1e-6f);
        const float behind[3] = {0, 0, 20};
        CHECK(LodManager::calculate_screen_coverage(behind, 1.0f, vp, 1920.0f, 1080.0f) == 0.0f);
        LodStats st; st.triangles_before = 1000; st.triangles_after = 250;
        CHECK(std::fabs(st.reduction_rate() - 0.75) < 0.001);
        // select_lod: three levels, coverage 0.01396 -> level 1; Force; cull threshold
        LodMesh m3; m3.bounding_radius = 1.0f;
        m3.with_level(LodLevel{0.05f, 10.0f, 4000, 2000, 0, 12000}).with_level(LodLevel{0.01f, 50.0f, 1000, 500, 12000, 3000})
          .with_level(LodLevel{0.0f, 1e9f, 100, 50, 15000, 300});
        LodManager lm; lm.begin_frame();
        const float cam[3] = {0, 0, 10};
        LodSelection s = lm.select_lod(m3, origin, vp, cam, 1920.0f, 1080.0f);
        CHECK(s.level == 1 && s.culled == 0 && s.blend == 0.0f && s.screen_coverage == cov);       // (0.01 - cov) / 0.01 < 0 -> clamped to 0
        LodConfig force; force.mode = ASH_LOD_MODE_FORCE; force.force_level = 9;
        LodManager lf(force); s = lf.select_lod(m3, origin, vp, cam, 1920.0f, 1080.0f);
        CHECK(s.level == 2 && s.screen_coverage == 1.0f && lf.stats().objects_per_lod[2] == 1 && lf.stats().triangles_after == 100);
        LodConfig strict; strict.cull_threshold = 0.5f;
        LodManager lc(strict); s = lc.select_lod(m3, origin, vp, cam, 1920.0f, 1080.0f);
        CHECK(s.culled == 1 && s.level == 2 && lc.stats().objects_culled == 1 && lc.stats().triangles_before == 4000);
    }
    const uint32_t W = 320, H = 180;
    CudaComputeFeature f(W, H, 4096, 256, 9);
    FeatureFrameContext fc; FeatureRenderContext rc; Device dev;
    f.before_frame(fc); CHECK(f.last_code() == ASH_ERR_NOT_INITIALIZED);          // FeatureNotInitialized
    f.on_added(dev);
    if (!gpu) {                                                                    // no CUDA device here: loud failure, no fallback
        if (f.context() == nullptr) { CHECK(f.last_code() == ASH_ERR_DEVICE); std::puts("host tests ok (no device: create failed loudly)"); return 0; }
    }
    CHECK(f.context() != nullptr);
    Mat4 I; identity(I);
    const float c[3] = {0, 0, 0}, e[3] = {0.5f, 0.5f, 0.5f};
    f.lod_meshes.push_back(LodMesh::single(1000, 1.0f));
    f.instancing.begin_frame();
    for (int i = 0; i < 64; ++i) {
        Mat4 M; identity(M); M[3][0] = (float)(i % 8) * 2.0f - 7.0f; M[3][1] = (float)(i / 8) * 2.0f - 7.0f;
        f.culling.add_object(bounding_box_from_center_extents(c, e), M, 0);
        f.instancing.add_instance(BatchKey{(uint32_t)(i % 2), 0}, instance_data_from_matrix(M));
        f.lod_mesh_ids.push_back(0);
    }
    f.instancing.finalize();
    std::vector<PointLight> pl(32); for (int i = 0; i < 32; ++i) { pl[i].position[0] = (float)i - 16.0f; pl[i].position[2] = 20.0f; pl[i].radius = 4.0f; }
    f.lights.update_lights(pl, {});
    FrameInputs in; identity(in.view); in.view[3][2] = -22.0f; in.camera_pos[2] = 22.0f;
    perspective_rh(0.7853982f, (float)W / (float)H, 0.5f, 100.0f, in.projection);
    std::vector<float> depth((size_t)W * H, 1.0f); std::vector<uint16_t> hdr((size_t)W * H * 4, 0x3800 /* 0.5h */);
    in.depth = depth.data(); in.hdr = hdr.data();
    f.set_frame_inputs(in); f.before_frame(fc); f.render(rc);
    CHECK(f.last_code() == ASH_OK);
    const uint32_t vis = f.read_visible_count();
    CHECK(vis > 0 && vis <= 64 && f.culling.stats().total_objects == 64);
    AshInstancingStats ist{};
    CHECK(f.read_instancing_stats(&ist) && ist.total_instances + ist.objects_culled == vis && ist.batch_count <= 2 && ist.triangles_before == 1000ull * vis);
    StatsCollector sc;
    CHECK(f.read_culling_stats(&sc) && sc.culling.objects_submitted == 64 && sc.culling.after_occlusion == vis && sc.culling.after_frustum >= vis);
    std::vector<uint8_t> ldr((size_t)W * H * 4);
    CHECK(ashcuda_readback(f.context(), ASH_BUF_LDR8, ldr.data(), ldr.size()) == ASH_OK && ldr[3] == 255 && ldr[0] > 0);
    // row f-1: the same frame with depth and HDR colour produced on the device (no FrameInputs buffers)
    f.produce_inputs = true;
    f.forward_params.near_plane = 0.5f;
    FrameInputs in2 = in; in2.depth = nullptr; in2.hdr = nullptr; in2.view_id = 9;
    f.set_frame_inputs(in2); f.before_frame(fc); f.render(rc);
    CHECK(f.last_code() == ASH_OK);
    CHECK((f.view_params().stages & (ASH_STAGE_FORWARD_DEPTH | ASH_STAGE_FORWARD_SHADE | ASH_STAGE_TONEMAP)) ==
          (ASH_STAGE_FORWARD_DEPTH | ASH_STAGE_FORWARD_SHADE | ASH_STAGE_TONEMAP));
    std::vector<uint64_t> keys((size_t)W * H);
    CHECK(ashcuda_readback(f.context(), ASH_BUF_FORWARD_KEYS, keys.data(), keys.size() * 8) == ASH_OK);
    size_t covered = 0; for (uint64_t k : keys) covered += (uint32_t)k != ASH_FWD_NO_OBJECT;
    CHECK(covered > 0 && covered < keys.size() && f.read_visible_count() > 0);
    f.on_removed(dev);
    std::printf("host tests ok (gpu: %u/64 visible, ldr[0]=%u)\n", vis, ldr[0]);
    return 0;
}
