// ashcuda.cu — the C ABI (include/ashcuda.h): context, buffers, stage dispatch.
// Plays the role of the reference's LightCullingPipeline (vulkan/light_culling_pipeline.rs:41-351:
// create buffers, upload, dispatch, expose the output buffer, free in Drop) for every stage.
#include <cstdio>
#include <cstring>
#include <new>
#include "ash_internal.cuh"

using namespace ash;

static thread_local std::string g_create_error;

// ---------------------------------------------------------------- Profiler (ash_internal.cuh)
cudaEvent_t ash::Profiler::get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
void ash::Profiler::resolve() {      // caller has synchronised the stream
    for (const Span& sp : open) {
        float t = 0.0f;
        if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) { ms[sp.id] += t; ++n[sp.id]; }
        pool.push_back(sp.a); pool.push_back(sp.b);
    }
    open.clear();
}
void ash::Profiler::release_all() {
    for (const Span& sp : open) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (cudaEvent_t e : pool) cudaEventDestroy(e);
    open.clear(); pool.clear();
}

static int fail(ashcuda_ctx* c, int code, const char* msg) {
    if (c) c->err = msg;
    return code;
}

template <typename T>
static cudaError_t dmalloc(T** p, size_t n) {
    return cudaMalloc(reinterpret_cast<void**>(p), n ? n : 16);
}

static int hiz_level_count(uint32_t W, uint32_t H) {     // PIN-4: floor(log2(max(W,H))) + 1
    uint32_t m = W > H ? W : H;
    int n = 1;
    while (m > 1) { m >>= 1; ++n; }
    return n;
}

extern "C" {

const char* ashcuda_version(void) { return "ashcuda 0.1.0 (sm_100a)"; }

void ashcuda_default_config(AshCudaConfig* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->max_objects = ASH_DEFAULT_MAX_OBJECTS;     // occlusion_culling.rs:18
    cfg->max_lights = ASH_DEFAULT_MAX_LIGHTS;       // features/light_culling.rs:17
    cfg->max_draws = 4096;
    cfg->hiz_levels = ASH_DEFAULT_HIZ_LEVELS;       // occlusion_culling.rs:21
    cfg->bloom_mips = 5;                            // features/bloom.rs:29
    cfg->enable_post = 1;
    cfg->srgb_encode = 0;
    cfg->concurrent_frontend = 1;
    cfg->max_batches = 64;
    cfg->graph_replay = 1;
    cfg->precision = ASH_PRECISION_EXACT;
}

const char* ashcuda_last_error(ashcuda_ctx* ctx) {
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

void ashcuda_destroy(ashcuda_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    ash_view_graphs_destroy(c);
    void* owned[] = {c->depth_owned, c->pyr.mips, c->tile_minmax, c->depth_range, c->counters,
                     c->objs_owned, c->draws_in_owned, c->visibility, c->vis_bits, c->cta_counts,
                     c->draws_out, c->visible_count, c->lights_owned, c->camera_dev, c->cand,
                     c->cand_idx, c->cand_count, c->light_scratch, c->tile_data, c->tile_offsets, c->tile_packed, c->stats, c->color_owned,
                     c->velocity_owned, c->taa[0], c->taa[1], c->bloom_t, c->ldr8,
                     c->inst_owned, c->inst_batch_owned, c->inst_mesh_owned, c->lod_meshes_owned, c->batch_mesh,
                     c->lod_sel, c->bucket_of, c->cta_hist, c->bucket_offsets, c->bucket_totals, c->inst_out,
                     c->inst_draws, c->inst_stats, c->fwd_keys, c->fwd_spheres, c->fwd_rects, c->fwd_big,
                     c->fwd_lights_view, c->fwd_rays, c->fwd_consts, c->fwd_late, c->culling_stats};
    for (void* p : owned) if (p) cudaFree(p);
    for (uint32_t k = 1; k <= ASH_MAX_BLOOM_MIPS; ++k) if (c->bloom.d[k]) cudaFree(c->bloom.d[k]);
    c->prof.release_all();
    if (c->front_stream) cudaStreamSynchronize(c->front_stream);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->front_stream) cudaStreamDestroy(c->front_stream);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

static int create_impl(ashcuda_ctx* c) {
    const uint32_t W = c->W, H = c->H;
    const size_t P = (size_t)W * H;
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
    ASH_CUDA(c, cudaStreamCreateWithFlags(&c->front_stream, cudaStreamNonBlocking));
    ASH_CUDA(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    ASH_CUDA(c, cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    // ---- pyramid geometry (PIN-3/4)
    Pyramid& p = c->pyr;
    p.levels = hiz_level_count(W, H);
    p.w[0] = W; p.h[0] = H; p.off[0] = 0;
    uint32_t off = 0;
    for (int l = 1; l < p.levels; ++l) {
        p.w[l] = p.w[l - 1] / 2 ? p.w[l - 1] / 2 : 1;
        p.h[l] = p.h[l - 1] / 2 ? p.h[l - 1] / 2 : 1;
        p.off[l] = off;
        off += (p.w[l] * p.h[l] + 3u) & ~3u;
    }
    int f = 0;     // levels reducible inside one 64x64 tile: clean 2x halving in both dimensions
    while (f < 6 && f + 1 < p.levels && (W >> (f + 1)) >= 1 && (H >> (f + 1)) >= 1) ++f;
    p.fast_levels = f;
    c->tiles_x = (W + 15u) / 16u;
    c->tiles_y = (H + 15u) / 16u;
    const size_t tiles = (size_t)c->tiles_x * c->tiles_y;
    ASH_CUDA(c, dmalloc(&c->depth_owned, P * 4));
    ASH_CUDA(c, dmalloc(&p.mips, (size_t)off * 4));
    ASH_CUDA(c, dmalloc(&c->tile_minmax, tiles * 8));
    ASH_CUDA(c, dmalloc(&c->depth_range, 8));
    ASH_CUDA(c, dmalloc(&c->counters, 64));
    ASH_CUDA(c, cudaMemsetAsync(c->depth_range, 0xFF, 8, c->stream));
    ASH_CUDA(c, cudaMemsetAsync(c->counters, 0, 64, c->stream));
    // ---- cull
    const size_t N = c->cfg.max_objects;
    ASH_CUDA(c, dmalloc(&c->objs_owned, N * sizeof(AshCullObjectData)));
    ASH_CUDA(c, dmalloc(&c->draws_in_owned, (size_t)c->cfg.max_draws * sizeof(AshIndirectDrawCommand)));
    ASH_CUDA(c, dmalloc(&c->culling_stats, sizeof(AshCullingStats)));
    ASH_CUDA(c, cudaMemsetAsync(c->culling_stats, 0, sizeof(AshCullingStats), c->stream));
    ASH_CUDA(c, dmalloc(&c->visibility, N * 4));
    ASH_CUDA(c, dmalloc(&c->vis_bits, ((N + 31) / 32) * 4));
    ASH_CUDA(c, dmalloc(&c->cta_counts, ((N + 255) / 256) * 4));
    ASH_CUDA(c, dmalloc(&c->draws_out, N * sizeof(AshIndirectDrawCommand)));
    ASH_CUDA(c, dmalloc(&c->visible_count, 4));
    ASH_CUDA(c, cudaMemsetAsync(c->visible_count, 0, 4, c->stream));
    // ---- lights
    const size_t L = c->cfg.max_lights;
    ASH_CUDA(c, dmalloc(&c->lights_owned, L * sizeof(AshGpuLight)));
    ASH_CUDA(c, dmalloc(&c->camera_dev, sizeof(AshCullingCameraData)));
    ASH_CUDA(c, dmalloc(&c->cand, L * sizeof(LightCand)));
    ASH_CUDA(c, dmalloc(&c->cand_idx, L * 16));   // CandZ records (light_cull.cu), 16 B each
    ASH_CUDA(c, dmalloc(&c->cand_count, 4));
    {
        size_t po, lo, co;
        const size_t bytes = light_cull_scratch_bytes(c->W, c->H, c->cfg.max_lights, &po, &lo, &co);
        ASH_CUDA(c, dmalloc(&c->light_scratch, bytes));
        ASH_CUDA(c, cudaMemsetAsync(c->light_scratch, 0, bytes, c->stream));      // the prepass tickets start at zero
    }
    ASH_CUDA(c, dmalloc(&c->tile_data, tiles * ASH_TILE_STRIDE * 4));    // light_culling_pipeline.rs:55-56
    ASH_CUDA(c, dmalloc(&c->stats, sizeof(AshViewStats)));
    ASH_CUDA(c, cudaMemsetAsync(c->stats, 0, sizeof(AshViewStats), c->stream));
    // ---- post chain
    if (c->cfg.enable_post) {
        ASH_CUDA(c, dmalloc(&c->color_owned, P * 8));
        ASH_CUDA(c, dmalloc(&c->velocity_owned, P * 4));
        const size_t texel = c->cfg.fp32_intermediates ? 16 : 8;    // RGBA32F intermediates (PIN-6, second mode) or RGBA16F
        ASH_CUDA(c, dmalloc(&c->taa[0], P * texel));
        ASH_CUDA(c, dmalloc(&c->taa[1], P * texel));
        ASH_CUDA(c, dmalloc(&c->bloom_t, P * texel));
        ASH_CUDA(c, dmalloc(&c->ldr8, P * 4));
        BloomChain& b = c->bloom;
        b.mips = c->cfg.bloom_mips;
        b.w[0] = W; b.h[0] = H; b.d[0] = c->bloom_t;
        for (uint32_t k = 1; k <= b.mips; ++k) {     // PIN-9: extent max(1, floor(prev/2))
            b.w[k] = b.w[k - 1] / 2 ? b.w[k - 1] / 2 : 1;
            b.h[k] = b.h[k - 1] / 2 ? b.h[k - 1] / 2 : 1;
            ASH_CUDA(c, dmalloc(&b.d[k], (size_t)b.w[k] * b.h[k] * texel));
        }
        // which levels may run the fast kernels: every sampler shortcut is validated here, per extent
        c->center_exact = axis_center_exact((int)W, (float)W) && axis_center_exact((int)H, (float)H);
        for (uint32_t k = 0; k <= ASH_MAX_BLOOM_MIPS; ++k) b.down_fast[k] = b.up_fast[k] = false;
        for (uint32_t k = 1; k <= b.mips; ++k) {
            b.down_fast[k] = axis_down_taps_exact((int)b.w[k - 1], (int)b.w[k]) && axis_down_taps_exact((int)b.h[k - 1], (int)b.h[k]);
            b.up_fast[k - 1] = axis_up_taps_exact((int)b.w[k], (int)b.w[k - 1]) && axis_up_taps_exact((int)b.h[k], (int)b.h[k - 1]);
        }
    }
    ASH_CUDA(c, cudaStreamSynchronize(c->stream));
    return ASH_OK;
}

static void lod_config_default(AshLodConfig* k);
static void forward_params_default(AshForwardParams* p);

int ashcuda_create(ashcuda_ctx** out, int device, uint32_t width, uint32_t height,
                   const AshCudaConfig* cfg_or_null) {
    if (!out) { g_create_error = "ashcuda_create: out is NULL"; return ASH_ERR_INVALID; }
    *out = nullptr;
    if (width == 0 || height == 0 || width > 32768 || height > 32768) {
        g_create_error = "ashcuda_create: bad extent";
        return ASH_ERR_INVALID;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || device < 0 || device >= ndev) {   // no CPU fallback: fail loudly
        g_create_error = std::string("ashcuda_create: no usable CUDA device (") +
                         (e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range") + ")";
        return ASH_ERR_DEVICE;
    }
    ashcuda_ctx* c = new (std::nothrow) ashcuda_ctx();
    if (!c) { g_create_error = "out of host memory"; return ASH_ERR_OUT_OF_MEMORY; }
    c->device = device; c->W = width; c->H = height;
    if (cfg_or_null) c->cfg = *cfg_or_null; else ashcuda_default_config(&c->cfg);
    if (c->cfg.max_batches == 0) c->cfg.max_batches = 64;          // configs written before the field existed
    if (c->cfg.bloom_mips > ASH_MAX_BLOOM_MIPS || c->cfg.hiz_levels == 0 || c->cfg.max_batches > 128 ||
        c->cfg.precision > ASH_PRECISION_FAST || c->cfg.fp32_intermediates > 1u) {
        g_create_error = "ashcuda_create: bad config"; delete c; return ASH_ERR_INVALID;
    }
    lod_config_default(&c->lod_cfg);
    forward_params_default(&c->fwd_params);
    int rc = create_impl(c);
    if (rc != ASH_OK) { g_create_error = c->err; ashcuda_destroy(c); return rc; }
    *out = c;
    return ASH_OK;
}

int ashcuda_set_stream(ashcuda_ctx* c, void* cuda_stream) {
    if (!c) return ASH_ERR_INVALID;
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, cudaStreamSynchronize(c->stream));
    ash_view_graphs_destroy(c);                // captured views name the old stream
    if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
    if (cuda_stream) { c->stream = static_cast<cudaStream_t>(cuda_stream); }
    else { ASH_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
    return ASH_OK;
}

// ------------------------------------------------------------------------------ inputs
int ashcuda_set_depth(ashcuda_ctx* c, const float* depth, int on_device) {
    if (!c || !depth) return fail(c, ASH_ERR_INVALID, "set_depth: null");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (on_device) c->depth = depth;
    else {
        ASH_CUDA(c, cudaMemcpyAsync(c->depth_owned, depth, (size_t)c->W * c->H * 4, cudaMemcpyHostToDevice, c->stream));
        c->depth = c->depth_owned;
    }
    c->pyr.mip0 = c->depth;
    c->have_depth = true;
    c->have_pyramid = false;
    c->have_forward = false;            // the object keys of a forward_depth no longer describe this depth buffer
    return ASH_OK;
}

int ashcuda_set_objects(ashcuda_ctx* c, const AshCullObjectData* objs, uint32_t n,
                        const AshIndirectDrawCommand* draws_in, uint32_t n_draws, int on_device) {
    if (!c || (n && !objs) || (n_draws && !draws_in)) return fail(c, ASH_ERR_INVALID, "set_objects: null");
    if (n > c->cfg.max_objects) return fail(c, ASH_ERR_INVALID, "set_objects: n > max_objects");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (on_device) { c->objs = objs; c->draws_in = draws_in; }
    else {
        if (n_draws > c->cfg.max_draws) return fail(c, ASH_ERR_INVALID, "set_objects: n_draws > max_draws");
        if (n) ASH_CUDA(c, cudaMemcpyAsync(c->objs_owned, objs, (size_t)n * sizeof(*objs), cudaMemcpyHostToDevice, c->stream));
        if (n_draws) ASH_CUDA(c, cudaMemcpyAsync(c->draws_in_owned, draws_in, (size_t)n_draws * sizeof(*draws_in), cudaMemcpyHostToDevice, c->stream));
        c->objs = c->objs_owned; c->draws_in = c->draws_in_owned;
    }
    c->n_objs = n; c->n_draws = n_draws;
    return ASH_OK;
}

int ashcuda_set_lights(ashcuda_ctx* c, const AshGpuLight* lights, uint32_t n, int on_device) {
    if (!c || (n && !lights)) return fail(c, ASH_ERR_INVALID, "set_lights: null");
    if (n > c->cfg.max_lights) return fail(c, ASH_ERR_INVALID, "set_lights: n > max_lights");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (on_device) c->lights = lights;
    else {
        if (n) ASH_CUDA(c, cudaMemcpyAsync(c->lights_owned, lights, (size_t)n * sizeof(*lights), cudaMemcpyHostToDevice, c->stream));
        c->lights = c->lights_owned;
    }
    c->n_lights = n;
    return ASH_OK;
}

int ashcuda_set_camera(ashcuda_ctx* c, const AshCullingCameraData* cam) {
    if (!c || !cam) return fail(c, ASH_ERR_INVALID, "set_camera: null");
    ASH_CUDA(c, cudaSetDevice(c->device));
    c->camera_host = *cam;   // pageable source is staged by the runtime before the call returns
    ASH_CUDA(c, cudaMemcpyAsync(c->camera_dev, cam, sizeof(*cam), cudaMemcpyHostToDevice, c->stream));
    c->have_camera = true;
    return ASH_OK;
}

int ashcuda_set_color(ashcuda_ctx* c, const void* hdr, const void* vel, int on_device) {
    if (!c || !hdr) return fail(c, ASH_ERR_INVALID, "set_color: null");
    if (!c->cfg.enable_post) return fail(c, ASH_ERR_NOT_INITIALIZED, "set_color: post chain disabled in config");
    ASH_CUDA(c, cudaSetDevice(c->device));
    const size_t P = (size_t)c->W * c->H;
    if (on_device) { c->color = hdr; c->velocity = vel; }
    else {
        ASH_CUDA(c, cudaMemcpyAsync(c->color_owned, hdr, P * 8, cudaMemcpyHostToDevice, c->stream));
        c->color = c->color_owned;
        if (vel) { ASH_CUDA(c, cudaMemcpyAsync(c->velocity_owned, vel, P * 4, cudaMemcpyHostToDevice, c->stream)); c->velocity = c->velocity_owned; }
        else c->velocity = nullptr;
    }
    c->have_color = true;
    c->taa_ran = false;
    c->bloom_valid = false;
    return ASH_OK;
}

int ashcuda_taa_reset_history(ashcuda_ctx* c) {
    if (!c) return ASH_ERR_INVALID;
    c->history_valid = false;
    return ASH_OK;
}

// ------------------------------------------------------------------------------ stages
int ashcuda_hiz_build(ashcuda_ctx* c) {
    if (!c) return ASH_ERR_INVALID;
    if (!c->have_depth) return fail(c, ASH_ERR_NOT_INITIALIZED, "hiz_build: no depth buffer set");
    ASH_CUDA(c, cudaSetDevice(c->device));
    // the frame's depth range accumulates by atomics: start from the identity
    ASH_CUDA(c, cudaMemsetAsync(c->depth_range, 0xFF, 8, c->stream));
    ASH_CUDA(c, launch_depth_pyramid(c->pyr, c->tile_minmax, c->tiles_x, c->depth_range, c->counters,
                                     c->L()));
    c->have_pyramid = true;
    return ASH_OK;
}

int ashcuda_occlusion_cull(ashcuda_ctx* c, const AshCullingPushConstants* pc) {
    if (!c || !pc) return fail(c, ASH_ERR_INVALID, "occlusion_cull: null");
    if (!c->have_depth) return fail(c, ASH_ERR_NOT_INITIALIZED, "occlusion_cull: no depth buffer set");
    if (pc->object_count > c->n_objs) return fail(c, ASH_ERR_INVALID, "occlusion_cull: objectCount > uploaded objects");
    // hizLevels above the pyramid's own level count is legal (the reference always pushes 10): the sampler clamps
    if (pc->hiz_levels == 0) return fail(c, ASH_ERR_INVALID, "occlusion_cull: hizLevels == 0");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_pyramid) { int rc = ashcuda_hiz_build(c); if (rc) return rc; }
    if (pc->object_count == 0) {     // PIN-8: visibleCount is zeroed before each dispatch
        ASH_CUDA(c, cudaMemsetAsync(c->visible_count, 0, 4, c->stream));
        return ASH_OK;
    }
    ASH_CUDA(c, launch_occlusion_cull(c->objs, c->draws_in, c->n_draws, *pc, c->pyr, c->visibility, c->vis_bits,
                                      c->cta_counts, c->draws_out, c->visible_count, c->stats, c->culling_stats,
                                      0, false, nullptr, c->L()));
    return ASH_OK;
}

int ashcuda_light_cull(ashcuda_ctx* c, const AshLightCullingPushConstants* pc) {
    if (!c || !pc) return fail(c, ASH_ERR_INVALID, "light_cull: null");
    if (!c->have_depth) return fail(c, ASH_ERR_NOT_INITIALIZED, "light_cull: no depth buffer set");
    if (!c->have_camera) return fail(c, ASH_ERR_NOT_INITIALIZED, "light_cull: no camera uploaded");
    if (pc->screen_size[0] != c->W || pc->screen_size[1] != c->H)
        return fail(c, ASH_ERR_INVALID, "light_cull: screenSize differs from the context extent");
    if (pc->light_count > c->n_lights) return fail(c, ASH_ERR_INVALID, "light_cull: lightCount > uploaded lights");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_pyramid) { int rc = ashcuda_hiz_build(c); if (rc) return rc; }
    // getTileFrustum's four side planes (light_culling.comp:71-95) depend on invProjection and the extent only
    const bool build_planes = !c->planes_valid || memcmp(c->planes_ip, c->camera_host.inv_projection, sizeof(c->planes_ip)) != 0;
    ASH_CUDA(c, launch_light_cull(c->lights, *pc, c->camera_dev, c->tile_minmax, c->depth_range, c->cand,
                                  c->cand_idx, c->cand_count, c->light_scratch, c->cfg.max_lights, c->tile_data, c->stats,
                                  build_planes, c->prepass_cstats, c->L()));
    if (pc->light_count != 0u) { memcpy(c->planes_ip, c->camera_host.inv_projection, sizeof(c->planes_ip)); c->planes_valid = true; }
    return ASH_OK;
}

// the general kernels of post.cu are correct for any extent but several times slower: say so ONCE per context when a level
// of 2^20 texels or more runs them (the caller's screenParams / extents do not pass the exact-tap checks of exact_taps.cu)
static void note_general_kernel(ashcuda_ctx* c, const char* what, uint32_t w, uint32_t h) {
    if (c->warned_general || (uint64_t)w * h < (1u << 20)) return;
    c->warned_general = true;
    fprintf(stderr, "ashcuda: %s at %ux%u runs the general one-thread-per-pixel kernel (post.cu): its sampler taps are not "
                    "texel-exact for these extents / screenParams; results are the same, throughput is lower\n", what, w, h);
}

static bool taa_extent_ok(const ashcuda_ctx* c, const AshTaaPushConstants* pc) {
    return (uint32_t)pc->screen_params[0] == c->W && (uint32_t)pc->screen_params[1] == c->H;
}

// the 3x3 taps of taa_resolve.comp:118 depend on the caller's screenParams: validate once per value
static bool taa_taps_exact(ashcuda_ctx* c, const AshTaaPushConstants* pc) {
    if (memcmp(c->taa_chk, pc->screen_params, sizeof(c->taa_chk)) != 0) {
        memcpy(c->taa_chk, pc->screen_params, sizeof(c->taa_chk));
        c->taa_exact = axis_taa_taps_exact((int)c->W, pc->screen_params[0], pc->screen_params[2]) &&
                       axis_taa_taps_exact((int)c->H, pc->screen_params[1], pc->screen_params[3]);
    }
    return c->taa_exact;
}

// thr_or_null != NULL fuses the bloom threshold of the resolved frame into the resolve (view_run)
static int taa_resolve_impl(ashcuda_ctx* c, const AshTaaPushConstants* pc, void* thr_or_null, float threshold, bool* fused) {
    if (!c->have_color) return fail(c, ASH_ERR_NOT_INITIALIZED, "taa_resolve: no colour buffer set");
    if (!taa_extent_ok(c, pc)) return fail(c, ASH_ERR_INVALID, "taa_resolve: screenParams differ from the context extent");
    ASH_CUDA(c, cudaSetDevice(c->device));
    // PIN-12: history(t) = output(t-1); first frame (or after reset_history) history = current colour
    const void* hist = c->history_valid ? c->taa[c->taa_cur] : c->color;
    void* out = c->taa[c->taa_cur ^ 1];
    const bool exact = taa_taps_exact(c, pc);
    *fused = false;
    if (c->cfg.fp32_intermediates) {      // history is RGBA32F once a resolve exists, the colour input (first frame) RGBA16F
        ASH_CUDA(c, launch_taa_f(c->color, 0, hist, c->history_valid ? 1 : 0, c->velocity, out, 1, *pc, c->L()));
    } else if (exact && c->cfg.precision == ASH_PRECISION_FAST && !c->velocity && taa_in_halo(*pc)) {
        TaaTolArgs ta;
        ASH_CUDA(c, launch_taa_tol(c->color, hist, out, thr_or_null, ash_taa_tol_plan(c, *pc, threshold, &ta), c->L()));   // post_taa_tol.cu
        *fused = thr_or_null != nullptr;
    } else if (exact) {
        ASH_CUDA(c, launch_taa_fast(c->color, hist, c->velocity, out, thr_or_null, *pc, threshold, c->L()));
        *fused = thr_or_null != nullptr;
    } else {
        note_general_kernel(c, "TAA resolve", c->W, c->H);
        ASH_CUDA(c, launch_taa(c->color, hist, c->velocity, out, *pc, exact, c->L()));
    }
    return ASH_OK;
}

int ashcuda_taa_resolve(ashcuda_ctx* c, const AshTaaPushConstants* pc) {
    if (!c || !pc) return fail(c, ASH_ERR_INVALID, "taa_resolve: null");
    bool fused;
    int rc = taa_resolve_impl(c, pc, nullptr, 0.0f, &fused);
    if (rc) return rc;
    c->taa_cur ^= 1;
    c->history_valid = true;
    c->taa_ran = true;
    c->bloom_valid = false;
    return ASH_OK;
}

static const void* resolved_color(const ashcuda_ctx* c) { return c->taa_ran ? c->taa[c->taa_cur] : c->color; }

// T = threshold(resolved colour) into D_0; in the fp32_intermediates mode the source is RGBA32F iff a TAA resolve produced it
static cudaError_t bloom_threshold_any(ashcuda_ctx* c, float threshold) {
    BloomChain& b = c->bloom;
    if (c->cfg.fp32_intermediates)
        return launch_bloom_threshold_f(resolved_color(c), c->taa_ran ? 1 : 0, b.d[0], 1, c->W, c->H, threshold, c->L());
    return launch_bloom_threshold(resolved_color(c), b.d[0], c->W, c->H, threshold, c->center_exact, c->L());
}

// D_0 is stored by the TAA kernel and read by the first downsample and the last upsample.  Rebuilding it in those two
// from the resolve (fast mode; bit-identical; commit 'Experiment: D_0 never materialised') removed 132 MB per 4K view and
// was SLOWER: TAA 68.4 -> 63.3 us but downsample 29.6 -> 40.5 us, 6505 -> 6358 views/s.  DESIGN.md section 5b.
// One launch per level.  Collapsing the levels below 2^20 (or 2^18) texels into the stages of ONE persistent kernel — tiles
// handed out by an atomic ticket in stage order, a stage waiting on the finished-tile count of the one before, the same
// tile functions — was built and measured (commit eb64929, 'k_bloom_tail'): bit-identical, 20 -> 14 (16) launches per view,
// and SLOWER on both counts: 6355 -> 5300 (5750..5930) views/s, because CTAs that wait for a stage hold registers the other
// views' kernels would use, and single-view latency 0.206 -> 0.222 (0.217) ms, because a kernel boundary inside a CUDA
// graph (~4.4 us per small level, launch to launch) is cheaper than ticket + acquire poll + fence + release (~6.6 us per
// stage).  DESIGN.md section 5b has the table.
static int bloom_down(ashcuda_ctx* c, uint32_t mips) {          // D_k = downsample(D_{k-1}), PIN-9
    BloomChain& b = c->bloom;
    for (uint32_t k = 1; k <= mips; ++k) {
        if (c->cfg.fp32_intermediates) ASH_CUDA(c, launch_bloom_downsample_f(b.d[k - 1], b.w[k - 1], b.h[k - 1], b.d[k], b.w[k], b.h[k], 1, c->L()));
        else if (b.down_fast[k]) ASH_CUDA(c, launch_bloom_downsample_fast(b.d[k - 1], b.w[k - 1], b.h[k - 1], b.d[k], b.w[k], b.h[k], c->L()));
        else {
            note_general_kernel(c, "bloom downsample", b.w[k], b.h[k]);
            ASH_CUDA(c, launch_bloom_downsample(b.d[k - 1], b.w[k - 1], b.h[k - 1], b.d[k], b.w[k], b.h[k], c->L()));
        }
    }
    return ASH_OK;
}

static int bloom_up(ashcuda_ctx* c, const AshBloomParams* bp, int last) {   // U_k = D_k + upsample(U_{k+1}), k >= last
    BloomChain& b = c->bloom;
    for (int k = (int)bp->mip_count - 1; k >= last; --k) {
        if (c->cfg.fp32_intermediates) ASH_CUDA(c, launch_bloom_upsample_add_f(b.d[k + 1], b.w[k + 1], b.h[k + 1], b.d[k], b.w[k], b.h[k],
                                                                                  bp->upsample_intensity, 1, c->L()));
        else if (b.up_fast[k]) ASH_CUDA(c, launch_bloom_upsample_fast(b.d[k + 1], b.w[k + 1], b.h[k + 1], b.d[k], bp->upsample_intensity, c->L()));
        else {
            note_general_kernel(c, "bloom upsample", b.w[k], b.h[k]);
            ASH_CUDA(c, launch_bloom_upsample_add(b.d[k + 1], b.w[k + 1], b.h[k + 1], b.d[k], b.w[k], b.h[k],
                                                   bp->upsample_intensity, c->L()));
        }
    }
    return ASH_OK;
}

int ashcuda_bloom(ashcuda_ctx* c, const AshBloomParams* bp) {
    if (!c || !bp) return fail(c, ASH_ERR_INVALID, "bloom: null");
    if (!c->have_color) return fail(c, ASH_ERR_NOT_INITIALIZED, "bloom: no colour buffer set");
    if (bp->mip_count > c->bloom.mips) return fail(c, ASH_ERR_INVALID, "bloom: mip_count > config bloom_mips");
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, bloom_threshold_any(c, bp->threshold));
    int rc = bloom_down(c, bp->mip_count);
    if (rc) return rc;
    rc = bloom_up(c, bp, 0);
    if (rc) return rc;
    c->bloom_valid = true;
    return ASH_OK;
}

int ashcuda_tonemap(ashcuda_ctx* c, const AshPostProcessPushConstants* pc) {
    if (!c || !pc) return fail(c, ASH_ERR_INVALID, "tonemap: null");
    if (!c->have_color) return fail(c, ASH_ERR_NOT_INITIALIZED, "tonemap: no colour buffer set");
    ASH_CUDA(c, cudaSetDevice(c->device));
    const void* bloom = c->bloom_valid ? c->bloom.d[0] : nullptr;
    if (c->cfg.fp32_intermediates)
        ASH_CUDA(c, launch_tonemap_f(resolved_color(c), c->taa_ran ? 1 : 0, bloom, 1, c->ldr8, c->W, c->H, *pc, c->tonemap_op, c->L()));
    else if (c->tonemap_op == ASH_TONEMAP_ACES && c->center_exact && ((size_t)c->W * c->H) % 2 == 0)
        ASH_CUDA(c, launch_tonemap_fast(resolved_color(c), bloom, c->ldr8, c->W, c->H, *pc, c->L()));
    else        // the operators without a reference shader (PIN-13) run the general kernel
        ASH_CUDA(c, launch_tonemap(resolved_color(c), bloom, c->ldr8, c->W, c->H, *pc, c->center_exact, c->tonemap_op, c->L()));
    return ASH_OK;
}

int ashcuda_set_tonemap_operator(ashcuda_ctx* c, uint32_t op) {      // TonemappingFeature::set_operator, tonemapping.rs:92-94
    if (!c) return ASH_ERR_INVALID;
    if (op > ASH_TONEMAP_NONE) return fail(c, ASH_ERR_INVALID, "set_tonemap_operator: unknown operator");
    c->tonemap_op = op;
    return ASH_OK;
}

static int view_post(ashcuda_ctx* c, const AshCudaViewParams* vp) {
    int rc = ASH_OK;
    const uint32_t st = vp->stages;
    // ---- post chain: TAA+threshold fused, bloom chain, last upsample + tonemap fused (same results
    //      as the staged calls; falls back to them stage by stage when a fast kernel does not apply)
    const bool want_bloom = (st & ASH_STAGE_BLOOM) != 0;
    if (want_bloom && vp->bloom.mip_count > c->bloom.mips) return fail(c, ASH_ERR_INVALID, "view_run: mip_count > config bloom_mips");
    bool thr_done = false;
    if (st & ASH_STAGE_TAA) {
        rc = taa_resolve_impl(c, &vp->taa, want_bloom ? c->bloom.d[0] : nullptr, vp->bloom.threshold, &thr_done);
        if (rc) return rc;
        c->taa_cur ^= 1; c->history_valid = true; c->taa_ran = true; c->bloom_valid = false;
    }
    bool tonemap_done = false;
    if (want_bloom) {
        if (!c->have_color) return fail(c, ASH_ERR_NOT_INITIALIZED, "bloom: no colour buffer set");
        BloomChain& b = c->bloom;
        if (!thr_done)
            ASH_CUDA(c, bloom_threshold_any(c, vp->bloom.threshold));
        rc = bloom_down(c, vp->bloom.mip_count); if (rc) return rc;
        const bool fuse_last = (st & ASH_STAGE_TONEMAP) && vp->bloom.mip_count >= 1 && b.up_fast[0] && c->tonemap_op == ASH_TONEMAP_ACES &&
                               !c->cfg.fp32_intermediates;
        rc = bloom_up(c, &vp->bloom, fuse_last ? 1 : 0); if (rc) return rc;
        if (fuse_last) {
            const bool keep = (st & ASH_STAGE_KEEP_BLOOM) != 0;
            ASH_CUDA(c, launch_upsample_tonemap_fast(b.d[1], b.w[1], b.h[1], b.d[0], vp->bloom.upsample_intensity,
                                                     resolved_color(c), c->ldr8, vp->post, keep, c->L()));
            c->bloom_valid = keep;          // without KEEP_BLOOM, U_0 is never materialised (d[0] still holds D_0)
            tonemap_done = true;
        } else {
            c->bloom_valid = true;
        }
    }
    if ((st & ASH_STAGE_TONEMAP) && !tonemap_done) { rc = ashcuda_tonemap(c, &vp->post); if (rc) return rc; }
    return ASH_OK;
}

int ashcuda_view_run(ashcuda_ctx* c, const AshCudaViewParams* vp) {
    if (!c || !vp) return fail(c, ASH_ERR_INVALID, "view_run: null");
    bool handled = false;                      // view_graph.cu: replay of a captured view, else the eager body
    int rc = ash_view_run_graph(c, vp, &handled);
    if (rc || handled) return rc;
    return ash_view_run_eager(c, vp);
}

static int two_phase_cull(ashcuda_ctx* c, const AshCudaViewParams* vp);
static int forward_alloc(ashcuda_ctx* c);

}  // extern "C"

int ash_forward_alloc(ashcuda_ctx* c) { return forward_alloc(c); }     // view_graph.cu: allocations cannot be captured

int ash_view_run_eager(ashcuda_ctx* c, const AshCudaViewParams* vp) {
    int rc = ASH_OK;
    const uint32_t st = vp->stages;
    // ---- front end (Hi-Z -> cull, light cull): independent of the post chain of the same view, so it
    //      is forked onto the context's second stream and joined before view_run returns control of
    //      the main stream (inputs uploaded earlier on the main stream are ordered by the fork event)
    const bool front = (st & (ASH_STAGE_HIZ | ASH_STAGE_CULL | ASH_STAGE_LIGHTS | ASH_STAGE_INSTANCING | ASH_STAGE_FORWARD_DEPTH)) != 0;
    const bool post = (st & (ASH_STAGE_TAA | ASH_STAGE_BLOOM | ASH_STAGE_TONEMAP)) != 0;
    // (per-kernel profiling serialises the view so that event pairs time one kernel each; a view whose colour
    //  buffer is shaded from this view's tile lists has nothing to overlap: the post chain waits for it)
    const bool fork = front && post && c->cfg.concurrent_frontend && c->front_stream && c->prof.mask == 0 &&
                      !(st & ASH_STAGE_FORWARD_SHADE);
    cudaStream_t main_stream = c->stream;
    if (fork) {
        ASH_CUDA(c, cudaSetDevice(c->device));
        ASH_CUDA(c, cudaEventRecord(c->ev_fork, main_stream));
        ASH_CUDA(c, cudaStreamWaitEvent(c->front_stream, c->ev_fork, 0));
        c->stream = c->front_stream;
    }
    {   // camera block (light cull) and view id, by value: no host memory is read after this call returns
        const bool with_camera = (st & ASH_STAGE_LIGHTS) != 0;
        cudaError_t e = cudaSetDevice(c->device);
        if (e == cudaSuccess) e = launch_store_view(vp->camera, with_camera, vp->view_id, c->camera_dev, c->stats, c->L());
        if (e != cudaSuccess) { c->err = std::string("view_run: ") + cudaGetErrorString(e); rc = ASH_ERR_DEVICE; }
        else if (with_camera) { c->camera_host = vp->camera; c->have_camera = true; }
    }
    const bool two_phase = (st & ASH_STAGE_TWO_PHASE) != 0;
    if (!rc && two_phase) {
        const uint32_t need = ASH_STAGE_FORWARD_DEPTH | ASH_STAGE_HIZ | ASH_STAGE_CULL;
        if ((st & need) != need) rc = fail(c, ASH_ERR_INVALID, "view_run: ASH_STAGE_TWO_PHASE needs FORWARD_DEPTH | HIZ | CULL");
        else rc = two_phase_cull(c, vp);
    } else if (!rc && (st & ASH_STAGE_FORWARD_DEPTH)) rc = ashcuda_forward_depth(c, &vp->camera, vp->cull.object_count);
    if (!rc && (st & ASH_STAGE_HIZ)) rc = ashcuda_hiz_build(c);
    if (!rc && (st & ASH_STAGE_CULL) && !two_phase) rc = ashcuda_occlusion_cull(c, &vp->cull);
    if (!rc && (st & ASH_STAGE_INSTANCING)) {        // consumes visibility[] of the cull above
        AshInstancingParams ip{};
        std::memcpy(ip.view_proj, vp->cull.view_proj, sizeof(ip.view_proj));
        std::memcpy(ip.camera_pos, vp->camera.camera_pos, sizeof(ip.camera_pos));
        ip.screen_size[0] = vp->cull.screen_params[0]; ip.screen_size[1] = vp->cull.screen_params[1];
        ip.object_count = vp->cull.object_count;
        rc = ashcuda_instancing_finalize(c, &ip);
    }
    // CullingStats: the fields other stages own.  lights_culled needs the candidate count only: the light prepass stores it;
    // lod_triangles_saved (instancing) keeps the one-thread kernel at the end of the chain.
    const bool finish = (st & ASH_STAGE_CULL) && vp->cull.object_count;
    const bool finish_in_prepass = finish && (st & ASH_STAGE_LIGHTS) && !(st & ASH_STAGE_INSTANCING);
    if (!rc && (st & ASH_STAGE_LIGHTS)) {
        c->prepass_cstats = finish_in_prepass ? c->culling_stats : nullptr;
        rc = ashcuda_light_cull(c, &vp->light);
        c->prepass_cstats = nullptr;
    }
    if (!rc && (st & ASH_STAGE_FORWARD_SHADE)) rc = ashcuda_forward_shade(c);
    if (!rc && finish && !finish_in_prepass) {
        cudaError_t e = launch_culling_stats_finish(c->stats, (st & ASH_STAGE_INSTANCING) ? c->inst_stats : nullptr, vp->light.light_count,
                                                    (st & ASH_STAGE_LIGHTS) != 0, c->culling_stats, c->L());
        if (e != cudaSuccess) { c->err = std::string("view_run: ") + cudaGetErrorString(e); rc = ASH_ERR_DEVICE; }
    }
    if (fork) {
        cudaError_t e = cudaEventRecord(c->ev_join, c->front_stream);
        c->stream = main_stream;
        if (!rc && e != cudaSuccess) { c->err = std::string("view_run: ") + cudaGetErrorString(e); rc = ASH_ERR_DEVICE; }
    }
    if (!rc) rc = view_post(c, vp);
    if (fork) {                                                              // join, also on the error path: later main-stream work
                                                                             // must not race kernels still running on front_stream
        cudaError_t e = cudaStreamWaitEvent(c->stream, c->ev_join, 0);
        if (!rc && e != cudaSuccess) { c->err = std::string("view_run: ") + cudaGetErrorString(e); rc = ASH_ERR_DEVICE; }
    }
    return rc;
}

extern "C" {

// ------------------------------------------------------------------------------ (f-2) instancing / LOD
static void forward_params_default(AshForwardParams* p) {
    std::memset(p, 0, sizeof(*p));
    const float dir[4] = {-0.3f, -1.0f, -0.2f, 0.0f}, white[4] = {1.0f, 1.0f, 1.0f, 1.0f}, amb[4] = {0.03f, 0.03f, 0.03f, 1.0f};
    const float mr[4] = {0.0f, 0.5f, 1.0f, 1.0f}, bg[4] = {0.02f, 0.02f, 0.03f, 1.0f};
    std::memcpy(p->light_direction, dir, 16); std::memcpy(p->light_color, white, 16); std::memcpy(p->ambient_color, amb, 16);
    std::memcpy(p->base_color_factor, white, 16); std::memcpy(p->parameters, mr, 16); std::memcpy(p->background, bg, 16);
    p->near_plane = 0.1f; p->radius_scale = 1.0f; p->flags = ASH_FWD_TILE_LIGHTS;
}

static void lod_config_default(AshLodConfig* k) {          // LodConfig::default(), lod_system.rs:47-57
    std::memset(k, 0, sizeof(*k));
    k->mode = ASH_LOD_MODE_SCREEN_SIZE; k->smooth_transitions = 1u; k->transition_width = 0.1f;
    k->global_bias = 0.0f; k->cull_threshold = 0.001f; k->use_visibility = 1u;
}

static int instancing_alloc(ashcuda_ctx* c) {
    if (c->lod_sel) return ASH_OK;
    const size_t N = c->cfg.max_objects;
    c->n_buckets = c->cfg.max_batches * ASH_MAX_LOD_LEVELS;
    ASH_CUDA(c, dmalloc(&c->inst_owned, N * sizeof(AshInstanceData)));
    ASH_CUDA(c, dmalloc(&c->inst_batch_owned, N * 4));
    ASH_CUDA(c, dmalloc(&c->inst_mesh_owned, N * 4));
    ASH_CUDA(c, dmalloc(&c->batch_mesh, (size_t)c->cfg.max_batches * 4));
    ASH_CUDA(c, dmalloc(&c->lod_sel, N * sizeof(AshLodSelection)));
    ASH_CUDA(c, dmalloc(&c->bucket_of, N * 4));
    ASH_CUDA(c, dmalloc(&c->cta_hist, instancing_hist_bytes(c->cfg.max_objects, c->n_buckets)));
    ASH_CUDA(c, dmalloc(&c->bucket_offsets, ((size_t)c->n_buckets + 1) * 4));
    ASH_CUDA(c, dmalloc(&c->bucket_totals, (size_t)c->n_buckets * 4));
    ASH_CUDA(c, dmalloc(&c->inst_out, N * sizeof(AshInstanceData)));
    ASH_CUDA(c, dmalloc(&c->inst_draws, (size_t)c->n_buckets * sizeof(AshIndirectDrawCommand)));
    ASH_CUDA(c, dmalloc(&c->inst_stats, sizeof(AshInstancingStats)));
    ASH_CUDA(c, cudaMemsetAsync(c->inst_stats, 0, sizeof(AshInstancingStats), c->stream));
    return ASH_OK;
}

int ashcuda_set_instances(ashcuda_ctx* c, const AshInstanceData* instances, const uint32_t* batch_id,
                          const uint32_t* lod_mesh_id, uint32_t n, int on_device) {
    if (!c || (n && (!instances || !batch_id || !lod_mesh_id))) return fail(c, ASH_ERR_INVALID, "set_instances: null");
    if (n > c->cfg.max_objects) return fail(c, ASH_ERR_INVALID, "set_instances: n > max_objects");
    ASH_CUDA(c, cudaSetDevice(c->device));
    int rc = instancing_alloc(c);
    if (rc) return rc;
    if (on_device) { c->inst = instances; c->inst_batch = batch_id; c->inst_mesh = lod_mesh_id; }
    else {
        for (uint32_t i = 0; i < n; ++i)
            if (batch_id[i] >= c->cfg.max_batches) return fail(c, ASH_ERR_INVALID, "set_instances: batch id >= max_batches");
        if (n) {
            ASH_CUDA(c, cudaMemcpyAsync(c->inst_owned, instances, (size_t)n * sizeof(*instances), cudaMemcpyHostToDevice, c->stream));
            ASH_CUDA(c, cudaMemcpyAsync(c->inst_batch_owned, batch_id, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
            ASH_CUDA(c, cudaMemcpyAsync(c->inst_mesh_owned, lod_mesh_id, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        }
        c->inst = c->inst_owned; c->inst_batch = c->inst_batch_owned; c->inst_mesh = c->inst_mesh_owned;
    }
    c->n_inst = n;
    // a BatchKey names one mesh: take it from the batch's lowest-index object
    ASH_CUDA(c, launch_batch_mesh(c->inst_batch, c->inst_mesh, n, c->cfg.max_batches, c->batch_mesh, c->stream));
    c->launches += 2 + (n ? 1 : 0);
    c->have_instancing = true;
    return ASH_OK;
}

int ashcuda_set_lod_meshes(ashcuda_ctx* c, const AshLodMesh* meshes, uint32_t n, int on_device) {
    if (!c || !meshes || !n) return fail(c, ASH_ERR_INVALID, "set_lod_meshes: null or empty");
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (on_device) c->lod_meshes = meshes;
    else {
        for (uint32_t i = 0; i < n; ++i)
            if (meshes[i].level_count > ASH_MAX_LOD_LEVELS) return fail(c, ASH_ERR_INVALID, "set_lod_meshes: level_count > MAX_LOD_LEVELS");
        if (n > c->lod_meshes_cap) {
            ASH_CUDA(c, cudaStreamSynchronize(c->stream));
            if (c->lod_meshes_owned) cudaFree(c->lod_meshes_owned);
            c->lod_meshes_owned = nullptr; c->lod_meshes_cap = 0;
            ASH_CUDA(c, dmalloc(&c->lod_meshes_owned, (size_t)n * sizeof(AshLodMesh)));
            c->lod_meshes_cap = n;
        }
        ASH_CUDA(c, cudaMemcpyAsync(c->lod_meshes_owned, meshes, (size_t)n * sizeof(*meshes), cudaMemcpyHostToDevice, c->stream));
        c->lod_meshes = c->lod_meshes_owned;
    }
    c->n_lod_meshes = n;
    return ASH_OK;
}

int ashcuda_set_lod_config(ashcuda_ctx* c, const AshLodConfig* k) {
    if (!c) return fail(c, ASH_ERR_INVALID, "set_lod_config: null");
    if (k) {
        if (k->mode > ASH_LOD_MODE_FORCE) return fail(c, ASH_ERR_INVALID, "set_lod_config: unknown mode");
        c->lod_cfg = *k;
    } else lod_config_default(&c->lod_cfg);
    return ASH_OK;
}

int ashcuda_instancing_finalize(ashcuda_ctx* c, const AshInstancingParams* ip) {
    if (!c || !ip) return fail(c, ASH_ERR_INVALID, "instancing_finalize: null");
    if (!c->have_instancing) return fail(c, ASH_ERR_NOT_INITIALIZED, "instancing_finalize: no instances set");
    if (!c->lod_meshes) return fail(c, ASH_ERR_NOT_INITIALIZED, "instancing_finalize: no LOD meshes set");
    if (ip->object_count > c->n_inst) return fail(c, ASH_ERR_INVALID, "instancing_finalize: object_count > instances set");
    if (c->lod_cfg.use_visibility && ip->object_count > c->n_objs)
        return fail(c, ASH_ERR_INVALID, "instancing_finalize: use_visibility needs a cull result for every object");
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, launch_instancing(c->inst, c->inst_batch, c->inst_mesh, c->lod_meshes, c->n_lod_meshes, c->batch_mesh,
                                  c->visibility, *ip, c->lod_cfg, c->n_buckets, c->lod_sel, c->bucket_of, c->cta_hist,
                                  c->bucket_offsets, c->bucket_totals, c->inst_out, c->inst_draws, c->inst_stats, c->L()));
    return ASH_OK;
}

// ------------------------------------------------------------------------------ (f-1) forward producers
static int forward_alloc(ashcuda_ctx* c) {
    if (c->fwd_keys) return ASH_OK;
    const size_t P = (size_t)c->W * c->H, N = c->cfg.max_objects;
    ASH_CUDA(c, dmalloc(&c->fwd_keys, P * 8));
    ASH_CUDA(c, dmalloc(&c->fwd_spheres, N * sizeof(float4)));
    ASH_CUDA(c, dmalloc(&c->fwd_rects, N * sizeof(int4)));
    ASH_CUDA(c, dmalloc(&c->fwd_big, (N + 1) * 4));
    ASH_CUDA(c, dmalloc(&c->fwd_lights_view, (size_t)c->cfg.max_lights * sizeof(float4)));
    ASH_CUDA(c, dmalloc(&c->fwd_rays, ((size_t)c->W + c->H) * 4));
    ASH_CUDA(c, dmalloc(&c->fwd_consts, forward_consts_bytes()));
    ASH_CUDA(c, dmalloc(&c->fwd_late, N * 4));
    return ASH_OK;
}

static ForwardBuffers forward_buffers(ashcuda_ctx* c) {
    return ForwardBuffers{c->fwd_spheres, c->fwd_rects, c->fwd_big + 1, c->fwd_big, c->fwd_lights_view, c->fwd_rays, c->fwd_rays + c->W,
                          c->fwd_consts, c->fwd_keys, c->depth_owned};
}

int ashcuda_set_forward_params(ashcuda_ctx* c, const AshForwardParams* p) {
    if (!c) return fail(c, ASH_ERR_INVALID, "set_forward_params: null");
    if (p) {
        if (!(p->near_plane > 0.0f) || !(p->radius_scale >= 0.0f)) return fail(c, ASH_ERR_INVALID, "set_forward_params: near_plane <= 0 or radius_scale < 0");
        c->fwd_params = *p;
    } else forward_params_default(&c->fwd_params);
    return ASH_OK;
}

int ashcuda_forward_depth(ashcuda_ctx* c, const AshCullingCameraData* cam, uint32_t object_count) {
    if (!c || !cam) return fail(c, ASH_ERR_INVALID, "forward_depth: null");
    if (object_count > c->n_objs) return fail(c, ASH_ERR_INVALID, "forward_depth: object_count > uploaded objects");
    if (cam->projection[2][3] != -1.0f || cam->projection[3][3] != 0.0f)
        return fail(c, ASH_ERR_INVALID, "forward_depth: projection is not a right-handed perspective (w_clip = -z_view)");
    if (c->W > 65535u || c->H > 65535u) return fail(c, ASH_ERR_INVALID, "forward_depth: extent above 65535 (rectangles are packed in 16 bits)");
    ASH_CUDA(c, cudaSetDevice(c->device));
    int rc = forward_alloc(c);
    if (rc) return rc;
    const ForwardBuffers B = forward_buffers(c);
    ASH_CUDA(c, launch_forward_setup(*cam, c->fwd_params, c->objs, object_count, c->lights, c->n_lights, c->W, c->H, B, c->L()));
    ASH_CUDA(c, launch_forward_splat(object_count, c->W, c->H, B, nullptr, c->L()));
    ASH_CUDA(c, launch_forward_resolve(c->W, c->H, B, c->L()));
    c->depth = c->depth_owned; c->pyr.mip0 = c->depth;
    c->have_depth = true; c->have_pyramid = false; c->have_forward = true;
    return ASH_OK;
}

// Row f-3: two-phase occlusion culling around the device rasteriser (forward_depth's splat).  The pyramid the
// previous view of this context left (mip 0 = the owned depth buffer, untouched until the resolve below)
// is phase 1's occluder set; on a first view phase 1 is the frustum test alone.
static int two_phase_cull(ashcuda_ctx* c, const AshCudaViewParams* vp) {
    const AshCullingPushConstants* pc = &vp->cull;
    if (pc->object_count > c->n_objs) return fail(c, ASH_ERR_INVALID, "two_phase: objectCount > uploaded objects");
    if (pc->object_count == 0) return fail(c, ASH_ERR_INVALID, "two_phase: no objects");
    if (pc->hiz_levels == 0) return fail(c, ASH_ERR_INVALID, "two_phase: hizLevels == 0");
    if (vp->camera.projection[2][3] != -1.0f || vp->camera.projection[3][3] != 0.0f)
        return fail(c, ASH_ERR_INVALID, "two_phase: projection is not a right-handed perspective (w_clip = -z_view)");
    if (c->W > 65535u || c->H > 65535u) return fail(c, ASH_ERR_INVALID, "two_phase: extent above 65535");
    ASH_CUDA(c, cudaSetDevice(c->device));
    int rc = forward_alloc(c);
    if (rc) return rc;
    const ForwardBuffers B = forward_buffers(c);
    const uint32_t n = pc->object_count;
    const bool have_prev = c->have_pyramid && c->depth == c->depth_owned;
    ASH_CUDA(c, launch_forward_setup(vp->camera, c->fwd_params, c->objs, n, c->lights, c->n_lights, c->W, c->H, B, c->L()));
    c->depth = c->depth_owned; c->pyr.mip0 = c->depth;
    // phase 1: against the previous pyramid; draw the survivors; Hi-Z of what was drawn
    ASH_CUDA(c, launch_occlusion_cull(c->objs, c->draws_in, c->n_draws, *pc, c->pyr, c->visibility, c->vis_bits, c->cta_counts, c->draws_out,
                                      c->visible_count, c->stats, c->culling_stats, 1, !have_prev, nullptr, c->L()));
    ASH_CUDA(c, launch_forward_splat(n, c->W, c->H, B, c->visibility, c->L()));
    ASH_CUDA(c, launch_forward_resolve(c->W, c->H, B, c->L()));
    c->have_depth = true; c->have_forward = true;
    rc = ashcuda_hiz_build(c);
    if (rc) return rc;
    // phase 2: everything phase 1 rejected, against the new pyramid; draw the late objects; final depth
    ASH_CUDA(c, launch_occlusion_cull(c->objs, c->draws_in, c->n_draws, *pc, c->pyr, c->visibility, c->vis_bits, c->cta_counts, c->draws_out,
                                      c->visible_count, c->stats, c->culling_stats, 2, false, c->fwd_late, c->L()));
    ASH_CUDA(c, launch_forward_splat(n, c->W, c->H, B, c->fwd_late, c->L()));
    ASH_CUDA(c, launch_forward_resolve(c->W, c->H, B, c->L()));
    c->have_pyramid = false;                      // the caller rebuilds it from the final depth (ASH_STAGE_HIZ follows)
    return ASH_OK;
}

int ashcuda_forward_shade(ashcuda_ctx* c) {
    if (!c) return fail(c, ASH_ERR_INVALID, "forward_shade: null");
    if (!c->cfg.enable_post) return fail(c, ASH_ERR_NOT_INITIALIZED, "forward_shade: post chain disabled in config (no colour target)");
    if (!c->have_forward) return fail(c, ASH_ERR_NOT_INITIALIZED, "forward_shade: no forward_depth for this view");
    const bool inst_color = (c->fwd_params.flags & ASH_FWD_INSTANCE_COLOR) != 0;
    if (inst_color && !c->have_instancing) return fail(c, ASH_ERR_NOT_INITIALIZED, "forward_shade: ASH_FWD_INSTANCE_COLOR without set_instances");
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, launch_forward_shade(c->fwd_params, c->fwd_keys, c->fwd_spheres, c->lights, c->n_lights, c->fwd_lights_view, c->tile_data,
                                     inst_color ? c->inst : nullptr, inst_color ? c->n_inst : 0u, c->fwd_consts, c->fwd_rays,
                                     c->fwd_rays + c->W, c->W, c->H, c->color_owned, c->L()));
    c->color = c->color_owned; c->velocity = nullptr;
    if (c->fwd_has_reproj) {          // real motion vectors for taa_resolve.comp:94-102 (previous-frame transforms, static scene)
        ASH_CUDA(c, launch_forward_velocity(c->fwd_keys, c->fwd_spheres, c->fwd_consts, c->fwd_rays, c->fwd_rays + c->W, c->fwd_reproj,
                                            c->W, c->H, c->velocity_owned, c->L()));
        c->velocity = c->velocity_owned;
    }
    c->have_color = true; c->taa_ran = false; c->bloom_valid = false;
    return ASH_OK;
}

int ashcuda_set_forward_reprojection(ashcuda_ctx* c, const float* reproj16) {
    if (!c) return ASH_ERR_INVALID;
    c->fwd_has_reproj = reproj16 != nullptr;
    if (reproj16) std::memcpy(c->fwd_reproj, reproj16, sizeof(c->fwd_reproj));
    return ASH_OK;
}

// ------------------------------------------------------------------------------ outputs
int ashcuda_hiz_level_count(ashcuda_ctx* c) { return c ? c->pyr.levels : ASH_ERR_INVALID; }

int ashcuda_mip_extent(ashcuda_ctx* c, int which, uint32_t* w, uint32_t* h) {
    if (!c || !w || !h) return ASH_ERR_INVALID;
    if (which >= 0x200 && which <= 0x200 + (int)c->bloom.mips && c->cfg.enable_post) { *w = c->bloom.w[which - 0x200]; *h = c->bloom.h[which - 0x200]; return ASH_OK; }
    if (which >= 0x100 && which < 0x100 + c->pyr.levels) { *w = c->pyr.w[which - 0x100]; *h = c->pyr.h[which - 0x100]; return ASH_OK; }
    return fail(c, ASH_ERR_NOT_FOUND, "mip_extent: unknown mip");
}

int ashcuda_get(ashcuda_ctx* c, int which, void** dev_ptr, size_t* bytes) {
    if (!c || !dev_ptr || !bytes) return fail(c, ASH_ERR_INVALID, "get: null");
    const size_t P = (size_t)c->W * c->H, tiles = (size_t)c->tiles_x * c->tiles_y;
    void* p = nullptr; size_t n = 0;
    if (which >= 0x200 && which <= 0x200 + (int)ASH_MAX_BLOOM_MIPS) {
        const int k = which - 0x200;
        if (!c->cfg.enable_post || k > (int)c->bloom.mips) return fail(c, ASH_ERR_NOT_FOUND, "get: no such bloom mip");
        p = c->bloom.d[k]; n = (size_t)c->bloom.w[k] * c->bloom.h[k] * (c->cfg.fp32_intermediates ? 16 : 8);
    } else if (which >= 0x100 && which < 0x100 + kMaxMips) {
        const int l = which - 0x100;
        if (l >= c->pyr.levels) return fail(c, ASH_ERR_NOT_FOUND, "get: no such Hi-Z mip");
        p = l == 0 ? const_cast<float*>(c->depth) : c->pyr.mips + c->pyr.off[l];
        n = (size_t)c->pyr.w[l] * c->pyr.h[l] * 4;
    } else switch (which) {
        case ASH_BUF_VISIBILITY:      p = c->visibility; n = (size_t)c->n_objs * 4; break;
        case ASH_BUF_DRAWS_OUT:       p = c->draws_out; n = (size_t)c->n_objs * sizeof(AshIndirectDrawCommand); break;
        case ASH_BUF_VISIBLE_COUNT:   p = c->visible_count; n = 4; break;
        case ASH_BUF_TILE_DATA:       p = c->tile_data; n = tiles * ASH_TILE_STRIDE * 4; break;
        case ASH_BUF_TAA_OUT:         p = c->taa[c->taa_cur]; n = P * (c->cfg.fp32_intermediates ? 16 : 8); break;
        case ASH_BUF_LDR8:            p = c->ldr8; n = P * 4; break;
        case ASH_BUF_STATS:           p = c->stats; n = sizeof(AshViewStats); break;
        case ASH_BUF_TILE_MINMAX:     p = c->tile_minmax; n = tiles * 8; break;
        case ASH_BUF_VISIBILITY_BITS: p = c->vis_bits; n = (((size_t)c->n_objs + 31) / 32) * 4; break;
        case ASH_BUF_BLOOM_THRESHOLD: p = c->bloom_t; n = P * (c->cfg.fp32_intermediates ? 16 : 8); break;
        case ASH_BUF_LOD_SELECTION:   p = c->lod_sel; n = (size_t)c->n_inst * sizeof(AshLodSelection); break;
        case ASH_BUF_INSTANCES_OUT:   p = c->inst_out; n = (size_t)c->n_inst * sizeof(AshInstanceData); break;
        case ASH_BUF_BUCKET_OFFSETS:  p = c->bucket_offsets; n = ((size_t)c->n_buckets + 1) * 4; break;
        case ASH_BUF_INSTANCE_DRAWS:  p = c->inst_draws; n = (size_t)c->n_buckets * sizeof(AshIndirectDrawCommand); break;
        case ASH_BUF_INSTANCING_STATS: p = c->inst_stats; n = sizeof(AshInstancingStats); break;
        case ASH_BUF_FORWARD_KEYS:    p = c->fwd_keys; n = P * 8; break;
        case ASH_BUF_COLOR:           p = const_cast<void*>(c->color); n = P * 8; break;
        case ASH_BUF_VELOCITY:        p = const_cast<void*>(c->velocity); n = c->velocity ? P * 4 : 0; if (!p) return fail(c, ASH_ERR_NOT_FOUND, "get: no velocity buffer"); break;
        case ASH_BUF_CULLING_STATS:   p = c->culling_stats; n = sizeof(AshCullingStats); break;
        default: return fail(c, ASH_ERR_NOT_FOUND, "get: unknown buffer id");
    }
    if (!p) return fail(c, ASH_ERR_NOT_FOUND, "get: buffer not allocated (post chain disabled, or no instances set)");
    *dev_ptr = p; *bytes = n;
    return ASH_OK;
}

int ashcuda_readback_async(ashcuda_ctx* c, int which, void* host_dst, size_t bytes) {
    void* p; size_t n;
    int rc = ashcuda_get(c, which, &p, &n);
    if (rc) return rc;
    if (!host_dst || bytes > n) return fail(c, ASH_ERR_INVALID, "readback: bad destination or size");
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, cudaMemcpyAsync(host_dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
    return ASH_OK;
}

int ashcuda_readback(ashcuda_ctx* c, int which, void* host_dst, size_t bytes) {
    int rc = ashcuda_readback_async(c, which, host_dst, bytes);
    if (rc) return rc;
    ASH_CUDA(c, cudaStreamSynchronize(c->stream));
    return ASH_OK;
}

int ashcuda_sync(ashcuda_ctx* c) {
    if (!c) return ASH_ERR_INVALID;
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, cudaStreamSynchronize(c->stream));
    return ASH_OK;
}

int ashcuda_debug_exact_taps(int kind, int n_src, int n_dst, float size_f, float inv) {
    switch (kind) {
        case 0: return axis_center_exact(n_dst, size_f) ? 1 : 0;
        case 1: return axis_taa_taps_exact(n_dst, size_f, inv) ? 1 : 0;
        case 2: return axis_down_taps_exact(n_src, n_dst) ? 1 : 0;
        case 3: return axis_up_taps_exact(n_src, n_dst) ? 1 : 0;
        case 4: { int d; float f; return axis_taa_history_regular(n_dst, size_f, inv, &d, &f) ? 1 : 0; }
        default: return 0;
    }
}

uint64_t ashcuda_kernel_launches(ashcuda_ctx* c) { return c ? c->launches : 0; }

int ashcuda_graph_stats(ashcuda_ctx* c, uint64_t* captures, uint64_t* replays) {
    if (!c || !captures || !replays) return fail(c, ASH_ERR_INVALID, "graph_stats: null");
    *captures = c->graph_captures; *replays = c->graph_replays;
    return ASH_OK;
}

int ashcuda_profile_kernels(ashcuda_ctx* c, uint32_t mask) {
    if (!c) return ASH_ERR_INVALID;
    c->prof.mask = mask & ((1u << ASH_KERNEL_COUNT) - 1u);
    return ASH_OK;
}

int ashcuda_kernel_time_ms(ashcuda_ctx* c, int k, float* ms, uint64_t* launches) {
    if (!c || !ms || !launches || k < 0 || k >= ASH_KERNEL_COUNT) return fail(c, ASH_ERR_INVALID, "kernel_time_ms: bad argument");
    ASH_CUDA(c, cudaSetDevice(c->device));
    ASH_CUDA(c, cudaStreamSynchronize(c->stream));
    c->prof.resolve();
    *ms = (float)c->prof.ms[k]; *launches = c->prof.n[k];
    c->prof.ms[k] = 0.0; c->prof.n[k] = 0;
    return ASH_OK;
}

int ashcuda_upload_view(ashcuda_ctx* c, const float* depth, const void* hdr, const void* vel) {
    if (!c) return ASH_ERR_INVALID;
    int rc = ASH_OK;
    if (depth) { rc = ashcuda_set_depth(c, depth, 0); if (rc) return rc; }
    if (hdr) { rc = ashcuda_set_color(c, hdr, vel, 0); if (rc) return rc; }
    return ASH_OK;
}

int ashcuda_download_view(ashcuda_ctx* c, uint32_t* visibility, uint32_t* tile_data, uint8_t* ldr8,
                          AshViewStats* stats, uint32_t* visible_count) {
    if (!c) return ASH_ERR_INVALID;
    ASH_CUDA(c, cudaSetDevice(c->device));
    const size_t P = (size_t)c->W * c->H, tiles = (size_t)c->tiles_x * c->tiles_y;
    if (visibility && c->n_objs)
        ASH_CUDA(c, cudaMemcpyAsync(visibility, c->visibility, (size_t)c->n_objs * 4, cudaMemcpyDeviceToHost, c->stream));
    if (tile_data)
        ASH_CUDA(c, cudaMemcpyAsync(tile_data, c->tile_data, tiles * ASH_TILE_STRIDE * 4, cudaMemcpyDeviceToHost, c->stream));
    if (ldr8) {
        if (!c->ldr8) return fail(c, ASH_ERR_NOT_FOUND, "download_view: post chain disabled");
        ASH_CUDA(c, cudaMemcpyAsync(ldr8, c->ldr8, P * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    if (stats) ASH_CUDA(c, cudaMemcpyAsync(stats, c->stats, sizeof(AshViewStats), cudaMemcpyDeviceToHost, c->stream));
    if (visible_count) ASH_CUDA(c, cudaMemcpyAsync(visible_count, c->visible_count, 4, cudaMemcpyDeviceToHost, c->stream));
    return ASH_OK;
}

int ashcuda_download_view_packed(ashcuda_ctx* c, uint32_t* visibility, uint32_t* tile_offsets, uint32_t* tile_lights,
                                 uint32_t tile_lights_capacity, uint8_t* ldr8, AshViewStats* stats, uint32_t* visible_count) {
    if (!c) return ASH_ERR_INVALID;
    ASH_CUDA(c, cudaSetDevice(c->device));
    if (tile_offsets || tile_lights) {
        if (!tile_offsets || !tile_lights || tile_lights_capacity == 0)
            return fail(c, ASH_ERR_INVALID, "download_view_packed: offsets, lights and a capacity go together");
        const uint32_t tiles = c->tiles_x * c->tiles_y;
        if (!c->tile_offsets) ASH_CUDA(c, dmalloc(&c->tile_offsets, ((size_t)tiles + 1) * 4));
        if (c->tile_packed_cap < tile_lights_capacity) {          // (stream-ordered free: earlier packs on this stream are done first)
            ASH_CUDA(c, cudaStreamSynchronize(c->stream));
            if (c->tile_packed) ASH_CUDA(c, cudaFree(c->tile_packed));
            c->tile_packed = nullptr; c->tile_packed_cap = 0;
            ASH_CUDA(c, dmalloc(&c->tile_packed, (size_t)tile_lights_capacity * 4));
            c->tile_packed_cap = tile_lights_capacity;
        }
        ASH_CUDA(c, launch_tile_pack(c->tile_data, tiles, c->tile_offsets, c->tile_packed, tile_lights_capacity, c->L()));
        ASH_CUDA(c, cudaMemcpyAsync(tile_offsets, c->tile_offsets, ((size_t)tiles + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
        ASH_CUDA(c, cudaMemcpyAsync(tile_lights, c->tile_packed, (size_t)tile_lights_capacity * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    return ashcuda_download_view(c, visibility, nullptr, ldr8, stats, visible_count);
}

}  // extern "C"
