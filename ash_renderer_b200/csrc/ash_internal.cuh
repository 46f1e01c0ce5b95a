// ash_internal.cuh — context layout and kernel launch prototypes shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/ashcuda.h"

namespace ash {

constexpr int kMaxMips = 16;

// Hi-Z pyramid in HBM: mip 0 is the depth buffer itself (PIN-3: aliased, never copied), mips 1..
// live back to back in one allocation, each level's offset rounded up to 4 floats (16 B).
struct Pyramid {
    const float* mip0;
    float*       mips;              // base of levels >= 1
    uint32_t     w[kMaxMips], h[kMaxMips];
    uint32_t     off[kMaxMips];     // float offset of level l (l >= 1) inside `mips`
    int          levels;            // levels_built (PIN-4)
    int          fast_levels;       // levels 1..fast_levels reduce inside one 64x64 depth tile
};

struct BloomChain {
    void*    d[ASH_MAX_BLOOM_MIPS + 1];   // d[k]: RGBA16F D_k, overwritten in place by U_k
    uint32_t w[ASH_MAX_BLOOM_MIPS + 1], h[ASH_MAX_BLOOM_MIPS + 1];
    uint32_t mips;
    bool     down_fast[ASH_MAX_BLOOM_MIPS + 1];   // [k]: level k-1 -> k may use k_downsample_fast
    bool     up_fast[ASH_MAX_BLOOM_MIPS + 1];     // [k]: level k+1 -> k may use k_upsample_fast
};

// One candidate light after the per-view prepass: view-space centre + radius, original index.
struct LightCand { float x, y, z, r; };

// Per-kernel-class device timing (ashcuda_profile_kernels / ashcuda_kernel_time_ms): CUDA events
// recorded on the ctx stream around each launch of a selected class, resolved lazily on query.
struct Profiler {
    struct Span { int id; cudaEvent_t a, b; };
    uint32_t mask = 0;
    std::vector<Span> open;                 // recorded, not yet resolved
    std::vector<cudaEvent_t> pool;          // free events
    double   ms[ASH_KERNEL_COUNT] = {};
    uint64_t n[ASH_KERNEL_COUNT] = {};
    cudaEvent_t get();
    void resolve();
    void release_all();
};

// What every launch wrapper needs: the stream, the launch counter, the optional profiler.
struct Launch {
    cudaStream_t s;
    uint64_t*    launches;
    Profiler*    prof;
    bool         fast = false;      // AshCudaConfig.precision == ASH_PRECISION_FAST: wrappers may pick the tolerance kernels
};
// RAII: counts the launch and brackets it with events when its class is being profiled.
struct Timed {
    const Launch& L; int id; cudaEvent_t b = nullptr;
    Timed(const Launch& l, int kid) : L(l), id(kid) {
        ++*L.launches;
        if (L.prof && (L.prof->mask >> id & 1u)) {
            cudaEvent_t a = L.prof->get(); b = L.prof->get();
            cudaEventRecord(a, L.s);
            L.prof->open.push_back({id, a, b});
        }
    }
    ~Timed() { if (b) cudaEventRecord(b, L.s); }
};

// bloom level launches are accounted per size class (ASH_KERNEL_BLOOM_TAIL: destination below 2^20 texels)
inline int bloom_class(int full_size_class, uint32_t dw, uint32_t dh) {
    return (uint64_t)dw * dh < (1u << 20) ? ASH_KERNEL_BLOOM_TAIL : full_size_class;
}

// One by-value kernel argument that changes from view to view: which kernel, which argument, its new value.
// CUDA-graph replay (view_graph.cu) patches exactly these nodes before each launch.
struct PatchArg {
    const void* func = nullptr;
    int         arg_index = 0, n_args = 0;
    alignas(16) unsigned char value[256];
};
void cull_patch(const AshCullingPushConstants& pc, int mode, PatchArg* out);                                   // occlusion_cull.cu
void taa_pair_patch(bool thresh, bool vel, const AshTaaPushConstants& pc, PatchArg* out);           // post_pair.cu
// per-view argument block of k_taa_tol (post_taa_tol.cu, precision = FAST).  regular: the host verified that the history
// taps of every column / row are texels clamp(p + d), clamp(p + d + 1) with one weight per axis (exact_taps.cu)
struct TaaTolArgs { AshTaaPushConstants pc; int dx, dy; float fx, fy, blend, threshold; uint32_t regular; };
void taa_tol_patch(bool thresh, const TaaTolArgs& a, PatchArg* out);                                // post_taa_tol.cu
bool taa_in_halo(const AshTaaPushConstants& pc);                                                     // post_fast.cu
void lod_patch(const AshInstancingParams& ip, const AshLodConfig& cfg, PatchArg* out);               // instancing.cu
void forward_setup_patch(const AshCullingCameraData& cam, PatchArg* out);                            // forward.cu
void forward_velocity_patch(const float reproj[4][4], PatchArg* out);                                // forward.cu

// What a captured view depends on besides the per-view values patched at replay (memcmp'd, zero-initialised).
struct GraphKey {
    cudaStream_t stream; const void *depth, *color, *velocity, *objs, *draws_in, *lights, *inst, *inst_batch, *inst_mesh, *lod_meshes;
    uint32_t stages, n_objs, n_draws, n_lights, n_inst, n_lod_meshes, object_count, hiz_levels, lod_use_visibility;
    AshLightCullingPushConstants light;
    float    taa_screen[4];
    float    inv_projection[4][4];                          // a graph without the k_tile_planes node is only valid for this matrix
    AshBloomParams bloom; AshPostProcessPushConstants post;
    uint32_t in_halo, tol_variant, tonemap_op, planes_cached, fwd_velocity, taa_cur, history_valid, taa_ran, have_color, have_depth, have_instancing, concurrent, have_pyramid;
    AshForwardParams fwd;
};
struct ViewGraph {
    cudaGraph_t graph = nullptr; cudaGraphExec_t exec = nullptr;
    cudaGraphNode_t n_store = nullptr, n_cull = nullptr, n_cull2 = nullptr, n_taa = nullptr, n_lod = nullptr, n_fwd = nullptr, n_vel = nullptr;
    GraphKey key{}; bool valid = false;
    uint64_t launches = 0;                                  // kernels per replay
    bool taa_flip = false, history_valid = false, taa_ran = false, bloom_valid = false, have_pyramid = false, have_camera = false;
    bool have_depth = false, have_color = false, have_forward = false;
    const float* depth = nullptr; const void* color = nullptr; const void* velocity = nullptr;
};

}  // namespace ash

struct ashcuda_ctx {
    int          device = 0;
    cudaStream_t stream = nullptr;
    bool         own_stream = false;
    cudaStream_t front_stream = nullptr;      // view_run: Hi-Z + cull + light cull run here (fork / join by events)
    cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;
    uint32_t     W = 0, H = 0, tiles_x = 0, tiles_y = 0;
    AshCudaConfig cfg{};
    std::string  err;
    uint64_t     launches = 0;
    ash::Profiler prof;
    ash::Launch  L() { return ash::Launch{stream, &launches, &prof, cfg.precision == ASH_PRECISION_FAST}; }

    // depth + pyramid
    const float* depth = nullptr;      // current mip 0 (owned copy or borrowed)
    float*       depth_owned = nullptr;
    ash::Pyramid pyr{};
    uint32_t*    tile_minmax = nullptr;   // tiles * 2
    uint32_t*    depth_range = nullptr;   // [min bits, max bits] over the frame
    uint32_t*    counters = nullptr;      // [0] depth-kernel ticket, [1] cull ticket, ...
    bool         have_depth = false, have_pyramid = false;

    // occlusion cull
    const AshCullObjectData*      objs = nullptr;
    AshCullObjectData*            objs_owned = nullptr;
    const AshIndirectDrawCommand* draws_in = nullptr;
    AshIndirectDrawCommand*       draws_in_owned = nullptr;
    uint32_t  n_objs = 0, n_draws = 0;
    uint32_t* visibility = nullptr;
    uint32_t* vis_bits = nullptr;
    uint32_t* cta_counts = nullptr;
    AshIndirectDrawCommand* draws_out = nullptr;
    uint32_t* visible_count = nullptr;
    AshCullingStats* culling_stats = nullptr;

    // light cull
    const AshGpuLight* lights = nullptr;
    AshGpuLight*       lights_owned = nullptr;
    uint32_t  n_lights = 0;
    AshCullingCameraData* camera_dev = nullptr;
    AshCullingCameraData  camera_host{};
    bool      have_camera = false;
    ash::LightCand* cand = nullptr;
    uint32_t* cand_idx = nullptr;
    uint32_t* cand_count = nullptr;
    uint8_t*  light_scratch = nullptr;   // tile planes + per-cell candidate lists (light_cull.cu)
    float     planes_ip[4][4] = {};      // invProjection the tile side planes in light_scratch were built from
    bool      planes_valid = false;

    // instancing / LOD (instancing.cu); allocated on the first ashcuda_set_instances
    const AshInstanceData* inst = nullptr;      AshInstanceData* inst_owned = nullptr;
    const uint32_t* inst_batch = nullptr;       uint32_t* inst_batch_owned = nullptr;
    const uint32_t* inst_mesh = nullptr;        uint32_t* inst_mesh_owned = nullptr;
    const AshLodMesh* lod_meshes = nullptr;     AshLodMesh* lod_meshes_owned = nullptr;
    uint32_t  n_inst = 0, n_lod_meshes = 0, lod_meshes_cap = 0, n_buckets = 0;
    AshLodConfig lod_cfg{};
    uint32_t* batch_mesh = nullptr;             // mesh id of each batch
    AshLodSelection* lod_sel = nullptr;
    uint32_t* bucket_of = nullptr;
    uint32_t* cta_hist = nullptr;
    uint32_t* bucket_offsets = nullptr;
    uint32_t* bucket_totals = nullptr;
    AshInstanceData* inst_out = nullptr;
    AshIndirectDrawCommand* inst_draws = nullptr;
    AshInstancingStats* inst_stats = nullptr;
    bool      have_instancing = false;

    // (f-1) compute forward pass (forward.cu); allocated on the first ashcuda_forward_depth
    AshForwardParams fwd_params{};
    unsigned long long* fwd_keys = nullptr;      // W*H: depth bits << 32 | object index
    float4*   fwd_spheres = nullptr;             // per object: view-space centre, radius
    int4*     fwd_rects = nullptr;               // per object: pixel rectangle [x0,y0,x1,y1)
    uint32_t* fwd_big = nullptr;                 // [0] count, [1..] object indices with a big rectangle
    float4*   fwd_lights_view = nullptr;
    float*    fwd_rays = nullptr;                // W + H ray-table entries
    void*     fwd_consts = nullptr;
    uint32_t* fwd_late = nullptr;                // two-phase culling: objects newly visible in phase 2
    bool      have_forward = false;              // keys / spheres of the current view are valid
    float     fwd_reproj[4][4] = {};             // prevViewProj * view^-1 (columns): forward_shade also writes RG16F velocity
    bool      fwd_has_reproj = false;

    // CUDA-graph replay of view_run (view_graph.cu): one captured graph per TAA ping-pong parity
    ash::ViewGraph vgraph[2];
    bool      graphs_broken = false;             // a capture failed once: stay on the eager path
    uint64_t  graph_replays = 0, graph_captures = 0;
    uint32_t* tile_data = nullptr;
    uint32_t* tile_offsets = nullptr;          // packed (CSR) tile lists: tiles + 1 offsets, allocated on first use
    uint32_t* tile_packed = nullptr;           // ... and tile_packed_cap light indices
    uint32_t  tile_packed_cap = 0;
    AshViewStats* stats = nullptr;

    // post chain (RGBA16F = 8 B / texel)
    const void* color = nullptr;
    void*       color_owned = nullptr;
    const void* velocity = nullptr;
    void*       velocity_owned = nullptr;
    void*       taa[2] = {nullptr, nullptr};   // ping-pong: taa[taa_cur] = latest resolve
    int         taa_cur = 0;
    bool        have_color = false, history_valid = false, taa_ran = false;
    void*       bloom_t = nullptr;             // T = D_0 for the staged path
    ash::BloomChain bloom{};
    bool        bloom_valid = false;
    void*       ldr8 = nullptr;
    uint32_t    tonemap_op = ASH_TONEMAP_ACES;  // TonemappingConfig.operator (features/tonemapping.rs:9-20)
    bool        center_exact = false;          // (x+0.5)/W taps hit texel x exactly (exact_taps.cu)
    float       taa_chk[4] = {0, 0, 0, 0};     // screen_params the TAA tap check was last run for
    bool        taa_exact = false;
    bool warned_general = false;                 // the one-line stderr note of ashcuda.cu: note_general_kernel
    AshCullingStats* prepass_cstats = nullptr;   // view_run: the light prepass stores CullingStats::lights_culled itself (no k_culling_stats_finish launch)
    std::vector<ash::TaaTolArgs> tol_plans;    // precision = FAST: history-tap plans by push constants (a Halton period has 16)
};

namespace ash {

// each returns cudaGetLastError() after launching on `s`; *launches is incremented per kernel
cudaError_t launch_depth_pyramid(const Pyramid& p, uint32_t* tile_minmax, uint32_t tiles_x,
                                 uint32_t* depth_range, uint32_t* ticket, const Launch& L);
cudaError_t launch_occlusion_cull(const AshCullObjectData* objs, const AshIndirectDrawCommand* draws_in,
                                  uint32_t n_draws, const AshCullingPushConstants& pc, const Pyramid& p,
                                  uint32_t* visibility, uint32_t* vis_bits, uint32_t* cta_counts,
                                  AshIndirectDrawCommand* draws_out, uint32_t* visible_count,
                                  AshViewStats* stats, AshCullingStats* cstats, int phase, bool no_occlusion,
                                  uint32_t* late, const Launch& L);
cudaError_t launch_culling_stats_finish(const AshViewStats* stats, const AshInstancingStats* inst_stats, uint32_t light_count,
                                        bool with_lights, AshCullingStats* cstats, const Launch& L);
cudaError_t launch_tile_pack(const uint32_t* tile_data, uint32_t tiles, uint32_t* offsets, uint32_t* lights,
                             uint32_t capacity, const Launch& L);                                            // light_cull.cu
cudaError_t launch_light_cull(const AshGpuLight* lights, const AshLightCullingPushConstants& pc,
                              const AshCullingCameraData* cam_dev, const uint32_t* tile_minmax,
                              uint32_t* depth_range, LightCand* cand, uint32_t* cand_idx,
                              uint32_t* cand_count, uint8_t* scratch, uint32_t max_lights,
                              uint32_t* tile_data, AshViewStats* stats, bool build_planes, AshCullingStats* cstats_or_null, const Launch& L);
cudaError_t launch_store_view(const AshCullingCameraData& cam, bool with_camera, uint32_t view_id, AshCullingCameraData* camera_dev,
                              AshViewStats* stats, const Launch& L);                                  // view_graph.cu
cudaError_t launch_forward_velocity(const unsigned long long* keys, const float4* spheres, const void* consts, const float* ray_x,
                                    const float* ray_y, const float reproj[4][4], uint32_t W, uint32_t H, void* velocity, const Launch& L);
size_t forward_consts_bytes();                                                                       // forward.cu
struct ForwardBuffers {                       // forward.cu: everything the depth producer touches
    float4* spheres; int4* rects; uint32_t* big_list; uint32_t* big_count; float4* lights_view; float* ray_x; float* ray_y;
    void* consts; unsigned long long* keys; float* depth;
};
cudaError_t launch_forward_setup(const AshCullingCameraData& cam, const AshForwardParams& fp, const AshCullObjectData* objs, uint32_t n,
                                 const AshGpuLight* lights, uint32_t n_lights, uint32_t W, uint32_t H, const ForwardBuffers& B, const Launch& L);
cudaError_t launch_forward_splat(uint32_t n, uint32_t W, uint32_t H, const ForwardBuffers& B, const uint32_t* mask_or_null, const Launch& L);
cudaError_t launch_forward_resolve(uint32_t W, uint32_t H, const ForwardBuffers& B, const Launch& L);
cudaError_t launch_forward_shade(const AshForwardParams& fp, const unsigned long long* keys, const float4* spheres,
                                 const AshGpuLight* lights, uint32_t n_lights, const float4* lights_view, const uint32_t* tile_data,
                                 const AshInstanceData* inst, uint32_t n_inst, const void* consts, const float* ray_x,
                                 const float* ray_y, uint32_t W, uint32_t H, void* hdr, const Launch& L);
cudaError_t launch_batch_mesh(const uint32_t* batch_id, const uint32_t* mesh_id, uint32_t n, uint32_t n_batches,
                              uint32_t* batch_mesh, cudaStream_t s);
size_t instancing_hist_bytes(uint32_t max_objects, uint32_t n_buckets);
cudaError_t launch_instancing(const AshInstanceData* inst, const uint32_t* batch_id, const uint32_t* mesh_id,
                              const AshLodMesh* meshes, uint32_t n_meshes, const uint32_t* batch_mesh,
                              const uint32_t* visibility, const AshInstancingParams& ip, const AshLodConfig& cfg,
                              uint32_t n_buckets, AshLodSelection* sel, uint32_t* bucket_of, uint32_t* cta_hist,
                              uint32_t* offsets, uint32_t* totals, AshInstanceData* out, AshIndirectDrawCommand* draws,
                              AshInstancingStats* stats, const Launch& L);
size_t light_cull_scratch_bytes(uint32_t W, uint32_t H, uint32_t max_lights, size_t* planes_off, size_t* list_off,
                                size_t* count_off, size_t* cell_planes_off = nullptr);
// `center`: the exact-tap checks passed, so centre / +-1 texel taps may be fetched directly
cudaError_t launch_taa(const void* cur, const void* hist, const void* vel, void* out,
                       const AshTaaPushConstants& pc, bool center, const Launch& L);
cudaError_t launch_bloom_threshold(const void* hdr, void* out, uint32_t W, uint32_t H, float threshold,
                                   bool center, const Launch& L);
cudaError_t launch_bloom_downsample(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                    uint32_t dh, const Launch& L);
cudaError_t launch_bloom_upsample_add(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                      uint32_t dh, float intensity, const Launch& L);
cudaError_t launch_tonemap(const void* hdr, const void* bloom_or_null, void* ldr8, uint32_t W, uint32_t H,
                           const AshPostProcessPushConstants& pc, bool center, uint32_t op, const Launch& L);

// ---- fp32_intermediates storage mode (post_f32.cu): the general kernels over RGBA16F / RGBA32F textures (flag = 1: RGBA32F)
cudaError_t launch_taa_f(const void* cur, int cur_f32, const void* hist, int hist_f32, const void* vel, void* out, int out_f32,
                         const AshTaaPushConstants& pc, const Launch& L);
cudaError_t launch_bloom_threshold_f(const void* hdr, int hdr_f32, void* out, int out_f32, uint32_t W, uint32_t H, float threshold,
                                     const Launch& L);
cudaError_t launch_bloom_downsample_f(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw, uint32_t dh, int f32, const Launch& L);
cudaError_t launch_bloom_upsample_add_f(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw, uint32_t dh, float intensity,
                                        int f32, const Launch& L);
cudaError_t launch_tonemap_f(const void* hdr, int hdr_f32, const void* bloom_or_null, int bloom_f32, void* ldr8, uint32_t W, uint32_t H,
                             const AshPostProcessPushConstants& pc, uint32_t op, const Launch& L);

// ---- fast post kernels (post_fast.cu): bit-identical to the general ones, valid when the exact-tap
//      checks below hold for the extents involved
cudaError_t launch_taa_fast(const void* cur, const void* hist, const void* vel_or_null, void* out, void* thr_or_null,
                            const AshTaaPushConstants& pc, float threshold, const Launch& L);
cudaError_t launch_taa_pair(const void* cur, const void* hist, const void* vel_or_null, void* out, void* thr_or_null,
                            const AshTaaPushConstants& pc, float threshold, const Launch& L);
cudaError_t launch_taa_tol(const void* cur, const void* hist, void* out, void* thr_or_null, const TaaTolArgs& a,
                           const Launch& L);                                                                 // post_taa_tol.cu (precision = FAST)
cudaError_t launch_bloom_downsample_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                         uint32_t dh, const Launch& L);
cudaError_t launch_bloom_downsample_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                         uint32_t dh, const Launch& L);
cudaError_t launch_bloom_downsample_box(const void* src, uint32_t sw, uint32_t sh, void* dst, uint32_t dw,
                                        uint32_t dh, const Launch& L);                                     // post_down_box.cu
cudaError_t launch_bloom_upsample_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity,
                                       const Launch& L);
cudaError_t launch_upsample_tonemap_fast(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity,
                                         const void* hdr, void* ldr8, const AshPostProcessPushConstants& pc,
                                         bool store_u, const Launch& L);
// post_up_pair.cu: the same two kernels with two source texels per thread on the packed fp32 pipe
bool upsample_pair_applies(uint32_t sw, uint32_t sh);
cudaError_t launch_upsample_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity, bool tol, cudaStream_t s);
cudaError_t launch_upsample_tonemap_pair(const void* src, uint32_t sw, uint32_t sh, void* dst, float intensity, const void* hdr,
                                         void* ldr8, const AshPostProcessPushConstants& pc, bool store_u, bool tol, cudaStream_t s);
cudaError_t launch_tonemap_fast(const void* hdr, const void* bloom_or_null, void* ldr8, uint32_t W, uint32_t H,
                                const AshPostProcessPushConstants& pc, const Launch& L);
// ---- exact-tap validation (exact_taps.cu), host side, O(n) per axis
bool axis_center_exact(int n, float size_f);
bool axis_taa_taps_exact(int n, float size_f, float inv);
bool axis_down_taps_exact(int sn, int dn);
bool axis_up_taps_exact(int sn, int dn);
bool axis_taa_history_regular(int n, float size_f, float jitter_delta, int* d_out, float* f_out);
float taa_uniform_blend(const AshTaaPushConstants& pc);

}  // namespace ash

#define ASH_CUDA(ctx, call)                                                              \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);            \
            return e__ == cudaErrorMemoryAllocation ? ASH_ERR_OUT_OF_MEMORY : ASH_ERR_DEVICE; \
        }                                                                                \
    } while (0)

// view_graph.cu / ashcuda.cu
int  ash_view_run_eager(ashcuda_ctx* c, const AshCudaViewParams* vp);
int  ash_view_run_graph(ashcuda_ctx* c, const AshCudaViewParams* vp, bool* handled);
void ash_view_graphs_destroy(ashcuda_ctx* c);
int  ash_forward_alloc(ashcuda_ctx* c);
const ash::TaaTolArgs& ash_taa_tol_plan(ashcuda_ctx* c, const AshTaaPushConstants& pc, float threshold, ash::TaaTolArgs* scratch);
