/* ashcuda.h — C ABI of libashcuda, the sm_100a CUDA backend for Ash-Renderer's per-frame
 * data-parallel compute path (Hi-Z pyramid, frustum + Hi-Z occlusion cull, 16x16 tiled light
 * cull, TAA resolve, bloom threshold/down/up, ACES tonemap).
 *
 * Every POD struct below is byte-identical to the reference's #[repr(C)] Rust struct it names
 * (paths relative to the reference repo), so a Rust `extern "C"` block can pass `&T` straight
 * through.  Every entry point cites the reference interface it replaces.
 *
 * Conventions: return 0 (ASH_OK) on success, a negative ASH_ERR_* otherwise; never aborts, never
 * throws across the boundary.  One ctx = one CUDA device + one stream, used by one thread at a
 * time (the reference's `RenderFeature: Send`, not `Sync` — features/feature_trait.rs:25).
 * Stage calls only enqueue work on the ctx stream; ashcuda_sync / ashcuda_readback wait.
 * There is no CPU fallback: without a CUDA device ashcuda_create fails with ASH_ERR_DEVICE.
 */
#ifndef ASHCUDA_H
#define ASHCUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes (mirror `enum AshError`, src/error.rs:12-29) -------------------------- */
#define ASH_OK                   0
#define ASH_ERR_DEVICE          -1 /* VulkanError / DeviceInitFailed: a CUDA call failed      */
#define ASH_ERR_NOT_INITIALIZED -2 /* FeatureNotInitialized: a stage ran before its inputs    */
#define ASH_ERR_NOT_FOUND       -3 /* ResourceNotFound: unknown buffer id / mip               */
#define ASH_ERR_INVALID         -4 /* bad argument (null, zero extent, size mismatch)         */
#define ASH_ERR_OUT_OF_MEMORY   -5 /* cudaMalloc failed                                       */

/* ---- constants (reference defaults; all are runtime-overridable via AshCudaConfig) ------- */
#define ASH_TILE_SIZE            16u  /* features/light_culling.rs:23, light_culling.comp:9-10 */
#define ASH_MAX_LIGHTS_PER_TILE 256u  /* features/light_culling.rs:20, light_culling.comp:15   */
#define ASH_TILE_STRIDE         257u  /* light_culling.comp:119                                */
#define ASH_DEFAULT_HIZ_LEVELS   10u  /* occlusion_culling.rs:21                               */
#define ASH_DEFAULT_MAX_LIGHTS 1024u  /* features/light_culling.rs:17 (host cap, lifted here)  */
#define ASH_DEFAULT_MAX_OBJECTS 65536u/* occlusion_culling.rs:18      (host cap, lifted here)  */
#define ASH_MAX_BLOOM_MIPS        8u

/* ---- POD structs ------------------------------------------------------------------------ */

/* CullBoundingBox, occlusion_culling.rs:26-31 ; BoundingBox, occlusion_cull.comp:14-17 */
typedef struct AshCullBoundingBox {
    float center[4];
    float extents[4];
} AshCullBoundingBox;

/* CullObjectData (112 B), occlusion_culling.rs:72-91 ; ObjectData, occlusion_cull.comp:19-26.
 * model[c] is COLUMN c (the Rust fields are named model_row* but hold to_cols_array_2d()). */
typedef struct AshCullObjectData {
    AshCullBoundingBox bounds;
    float    model[4][4];
    uint32_t draw_index;
    float    lod_bias;
    uint32_t flags;      /* bit 0 = enabled */
    uint32_t _padding;
} AshCullObjectData;

/* IndirectDrawCommand (20 B), occlusion_culling.rs:114-120 ; DrawCommand, occlusion_cull.comp:28-34 */
typedef struct AshIndirectDrawCommand {
    uint32_t index_count;
    uint32_t instance_count;
    uint32_t first_index;
    int32_t  vertex_offset;
    uint32_t first_instance;
} AshIndirectDrawCommand;

/* CullingPushConstants (96 B), occlusion_culling.rs:125-136 ; occlusion_cull.comp:69-75 */
typedef struct AshCullingPushConstants {
    float    view_proj[4][4];   /* column-major: view_proj[c] is column c */
    float    screen_params[4];  /* w, h, 1/w, 1/h */
    uint32_t object_count;
    uint32_t hiz_levels;
    uint32_t _padding[2];
} AshCullingPushConstants;

/* GpuLight (64 B), features/light_culling.rs:28-37 ; Light, light_culling.comp:18-23 */
typedef struct AshGpuLight {
    float position[4];   /* xyz, radius */
    float color[4];      /* rgb, intensity */
    float direction[4];  /* xyz, type (0 point, 1 spot, 2 directional) */
    float params[4];     /* inner, outer, falloff, enabled */
} AshGpuLight;

/* LightCullingPushConstants (16 B), features/light_culling.rs:69-76 ; light_culling.comp:26-30 */
typedef struct AshLightCullingPushConstants {
    uint32_t screen_size[2];
    uint32_t light_count;
    uint32_t _padding;
} AshLightCullingPushConstants;

/* CullingCameraData (208 B), features/light_culling.rs:81-86 ; CameraData, light_culling.comp:48-53 */
typedef struct AshCullingCameraData {
    float view[4][4];
    float projection[4][4];
    float inv_projection[4][4];
    float camera_pos[4];
} AshCullingCameraData;

/* TaaPushConstants (48 B), temporal_aa.rs:86-97 ; taa_resolve.comp:22-27 */
typedef struct AshTaaPushConstants {
    float screen_params[4];  /* w, h, 1/w, 1/h */
    float params[4];         /* blend, colour clamp on, velocity rejection on, pad */
    float jitter[2];
    float prev_jitter[2];
} AshTaaPushConstants;

/* PostProcessPushConstants (16 B), fullscreen_pass.rs:152-163 ; tonemapping.frag:11-16 */
typedef struct AshPostProcessPushConstants {
    float exposure;
    float gamma;
    float bloom_intensity;
    float _padding;
} AshPostProcessPushConstants;

/* Bloom chain parameters: BloomConfig (features/bloom.rs:9-33) as the three bloom shaders see
 * it (bloom_threshold.frag:10-15, bloom_downsample.frag:10-14, bloom_upsample.frag:10-14). */
typedef struct AshBloomParams {
    float    threshold;           /* bloom.rs:27  = 0.8 */
    float    upsample_intensity;  /* bloom_upsample.frag:12, 1.0 in the pinned chain */
    uint32_t mip_count;           /* bloom.rs:29  = 5   */
    uint32_t _padding;
} AshBloomParams;

/* ---- (f-2) consumers of the cull result: LOD selection + instance batching ------------------------
 * InstanceData (96 B), src/renderer/instancing.rs:19-34; the fields named model_row* hold COLUMNS
 * (to_cols_array_2d, :38-46); position() = model_row3.xyz (:68-70). */
typedef struct AshInstanceData {
    float model_row0[4], model_row1[4], model_row2[4], model_row3[4];
    float color[4];
    float custom[4];
} AshInstanceData;

/* ---- (f-1) producers: a compute forward pass that makes the depth and HDR colour buffers on the device --
 * The uniform fields of shaders/frag.frag the compute forward pass uses (MVP :12-23, Material :25-37);
 * view and projection come from AshCullingCameraData.  Textures (bindless indices -1), the shadow map
 * (shadow = 0) and normal maps are not part of it. */
typedef struct AshForwardParams {
    float light_direction[4];   /* MVP.light_direction: world-space direction the light travels            */
    float light_color[4];       /* MVP.light_color rgb                                                     */
    float ambient_color[4];     /* MVP.ambient_color rgb                                                   */
    float base_color_factor[4]; /* Material.base_color_factor                                              */
    float emissive_factor[4];   /* Material.emissive_factor                                                */
    float parameters[4];        /* Material.parameters: x metallic, y roughness (z, w unused: no textures) */
    float background[4];        /* HDR clear colour of pixels no occluder covers                           */
    float near_plane;           /* hits nearer than this along the view axis are clipped                   */
    float radius_scale;         /* occluder sphere radius = scale * min world half-extent of the AABB      */
    uint32_t flags;             /* ASH_FWD_*                                                               */
    uint32_t _padding;
} AshForwardParams;
#define ASH_FWD_TILE_LIGHTS    1u /* add the lights of the pixel's tile list (tileData of this view's light cull) */
#define ASH_FWD_REINHARD       2u /* frag.frag:226: color / (color + 1) on shaded pixels                            */
#define ASH_FWD_INSTANCE_COLOR 4u /* base colour *= AshInstanceData.color of the same index (needs set_instances)   */
#define ASH_FWD_NO_OBJECT      0xFFFFFFFFu

#define ASH_MAX_LOD_LEVELS          8u      /* lod_system.rs:15  MAX_LOD_LEVELS          */
#define ASH_MAX_INSTANCES_PER_BATCH 65536u  /* instancing.rs:16  MAX_INSTANCES_PER_BATCH */
#define ASH_LOD_NOT_VISIBLE         0xFFFFFFFFu

/* LodLevel (24 B), lod_system.rs:61-76 */
typedef struct AshLodLevel {
    float    screen_threshold;
    float    distance_threshold;
    uint32_t triangle_count;
    uint32_t vertex_count;
    uint32_t index_offset;
    uint32_t index_count;
} AshLodLevel;

/* LodMesh (lod_system.rs:78-87) with `levels: Vec<LodLevel>` flattened to MAX_LOD_LEVELS (208 B) */
typedef struct AshLodMesh {
    AshLodLevel levels[ASH_MAX_LOD_LEVELS];
    uint32_t    level_count;
    float       bounding_radius;
    float       lod_bias;
    uint32_t    _padding;
} AshLodMesh;

/* LodConfig + LodSelectionMode (lod_system.rs:17-58), 32 B */
#define ASH_LOD_MODE_SCREEN_SIZE 0u
#define ASH_LOD_MODE_DISTANCE    1u
#define ASH_LOD_MODE_FORCE       2u
typedef struct AshLodConfig {
    uint32_t mode;               /* ASH_LOD_MODE_*                         */
    uint32_t force_level;        /* LodSelectionMode::Force(level)          */
    uint32_t smooth_transitions; /* default 1                               */
    float    transition_width;   /* default 0.1                             */
    float    global_bias;        /* default 0.0                             */
    float    cull_threshold;     /* default 0.001                           */
    uint32_t use_visibility;     /* 1: only objects with visibility[i] == 1 (the occlusion cull's output) are
                                    selected and batched; 0: every object, as the CPU managers do alone   */
    uint32_t _padding;
} AshLodConfig;

/* LodSelection (16 B), lod_system.rs:126-137; level == ASH_LOD_NOT_VISIBLE for objects the cull rejected */
typedef struct AshLodSelection {
    uint32_t level;
    float    blend;
    uint32_t culled;
    float    screen_coverage;
} AshLodSelection;

/* LodStats (lod_system.rs:146-157) + InstancingStats (instancing.rs:155-166), 72 B */
typedef struct AshInstancingStats {
    uint32_t objects_per_lod[ASH_MAX_LOD_LEVELS];
    uint32_t objects_culled;           /* LodStats: too small on screen                            */
    uint32_t total_instances;          /* InstancingStats: instances accepted into batches          */
    uint32_t batch_count;              /* non-empty (batch, LOD) buckets = draw calls               */
    float    avg_instances_per_batch;  /* total_instances / batch_count (0 when there is no batch)  */
    uint64_t triangles_before;
    uint64_t triangles_after;
    uint32_t instances_dropped;        /* beyond MAX_INSTANCES_PER_BATCH in their bucket (:223-226) */
    uint32_t _padding;
} AshInstancingStats;

/* Per-view parameters of the standalone call (view_run takes them from its cull / camera blocks) */
typedef struct AshInstancingParams {
    float    view_proj[4][4];    /* columns; LodManager::select_lod's view_proj (lod_system.rs:241) */
    float    camera_pos[4];      /* xyz                                                      */
    float    screen_size[2];     /* width, height                                            */
    uint32_t object_count;
    uint32_t _padding;
} AshInstancingParams;

/* Library configuration; zero-initialise then call ashcuda_default_config(). */
typedef struct AshCudaConfig {
    uint32_t max_objects;     /* capacity of the object/visibility/draw buffers            */
    uint32_t max_lights;      /* capacity of the light buffer                              */
    uint32_t max_draws;       /* capacity of the draw-command template buffer              */
    uint32_t hiz_levels;      /* value pushed as hizLevels is min(levels_built, this)      */
    uint32_t bloom_mips;      /* bloom chain depth allocated (<= ASH_MAX_BLOOM_MIPS)       */
    uint32_t enable_post;     /* 0: allocate no colour/TAA/bloom/LDR targets               */
    uint32_t srgb_encode;     /* reserved (swapchain.rs:81-88 sRGB target): must be 0               */
    uint32_t concurrent_frontend; /* 1 (default): view_run runs Hi-Z + cull + light cull on a second
                                     stream, concurrently with the post chain of the same view     */
    uint32_t max_batches;     /* instancing: distinct batch ids (BatchKey values), default 64, <= 128 */
    uint32_t graph_replay;    /* 1 (default): view_run captures the view into a CUDA graph once and replays it,
                                 patching the per-view kernel arguments; 0: launch every kernel each view.
                                 Same kernels, same results either way (csrc/view_graph.cu)                 */
    uint32_t precision;       /* ASH_PRECISION_EXACT (0, default): every image bit-identical to the reference semantics
                                 as the CPU oracle states them; ASH_PRECISION_FAST (1): the post-chain colour
                                 arithmetic spends the tolerance BASELINE.json's north_star grants (images within
                                 1 fp16 ulp, 8-bit output within 1 LSB): approximate reciprocals, FMA contraction.
                                 Sampling positions / weights, visibility and light lists are exact in both modes. */
    uint32_t fp32_intermediates; /* 0 (default): TAA output / history and bloom targets are RGBA16F like the reference's render
                                 targets (hdr_framebuffer.rs:34, taa_resolve.comp:19).  1: they are RGBA32F (SURVEY PIN-6's
                                 second mode: nothing rounded to fp16 between stages; ASH_BUF_TAA_OUT / bloom mips / threshold
                                 are then 16 B per texel).  The parity mode for north_star's fp32 criterion: general kernels
                                 only, no graph replay, `precision` ignored.  The HDR colour input stays RGBA16F.            */
    uint32_t _reserved[4];
} AshCudaConfig;
#define ASH_PRECISION_EXACT 0u
#define ASH_PRECISION_FAST  1u

/* Per-view statistics (OcclusionStats, occlusion_culling.rs:152-185 + tile list stats).
 * This 32-byte record is the only thing exchanged between GPUs (allgather). */
typedef struct AshViewStats {
    uint32_t view_id;
    uint32_t total_objects;
    uint32_t visible_objects;
    uint32_t tiles;
    uint32_t sum_tile_lights;
    uint32_t max_tile_lights;
    uint32_t overflow_tiles;   /* tiles whose true count exceeded 256 */
    uint32_t candidate_lights; /* lights surviving the per-view depth-slab prefilter */
} AshViewStats;

/* CullingStats (render_stats.rs:54-65) as StatsCollector::record_culling / record_lod_savings (:213-234) would
 * fill it, plus the early / late split of two-phase culling.  ASH_BUF_CULLING_STATS. */
typedef struct AshCullingStats {
    uint32_t objects_submitted;
    uint32_t after_frustum;
    uint32_t after_occlusion;
    uint32_t lights_culled;        /* lights no tile of this view can see (light_count - candidates); view_run only */
    uint64_t lod_triangles_saved;  /* LodStats triangles_before - triangles_after; view_run with ASH_STAGE_INSTANCING */
    uint32_t early_visible;        /* two-phase: visible against the previous frame's pyramid */
    uint32_t late_visible;         /* two-phase: newly visible against the pyramid of what phase 1 drew */
} AshCullingStats;

/* Everything that changes per view-frame, for ashcuda_view_run. */
typedef struct AshCudaViewParams {
    AshCullingPushConstants      cull;
    AshLightCullingPushConstants light;
    AshCullingCameraData         camera;
    AshTaaPushConstants          taa;
    AshBloomParams               bloom;
    AshPostProcessPushConstants  post;
    uint32_t view_id;
    uint32_t stages;           /* bitmask of ASH_STAGE_* */
    uint32_t _padding[2];
} AshCudaViewParams;

#define ASH_STAGE_HIZ        0x01u
#define ASH_STAGE_CULL       0x02u
#define ASH_STAGE_LIGHTS     0x04u
#define ASH_STAGE_TAA        0x08u
#define ASH_STAGE_BLOOM      0x10u
#define ASH_STAGE_TONEMAP    0x20u
#define ASH_STAGE_ALL        0x3Fu
#define ASH_STAGE_INSTANCING 0x80u /* after the cull: LOD selection + instance batching (not part of ASH_STAGE_ALL) */
#define ASH_STAGE_FORWARD_DEPTH 0x100u /* before Hi-Z: depth buffer from the cull objects' occluder spheres (f-1)       */
#define ASH_STAGE_FORWARD_SHADE 0x200u /* after the light cull: HDR colour from the tile light lists (f-1, f-2)         */
#define ASH_STAGE_TWO_PHASE 0x400u /* with FORWARD_DEPTH | HIZ | CULL (row f-3): phase 1 culls against the pyramid the previous
                                      view left, its survivors are drawn, Hi-Z is rebuilt, phase 2 tests the rest and draws the
                                      newly visible ("late") objects; visibility = early | late, final pyramid from both       */
#define ASH_STAGE_KEEP_BLOOM 0x40u /* view_run: also store the bloom buffer U_0 (the fused last upsample +
                                      tonemap otherwise never writes it; ASH_BUF_BLOOM_MIP(0) then holds D_0) */

/* Buffer ids for ashcuda_get / ashcuda_readback. */
#define ASH_BUF_HIZ_MIP(i)    (0x100 + (i)) /* fp32 W_i x H_i; mip 0 is the depth buffer      */
#define ASH_BUF_BLOOM_MIP(i)  (0x200 + (i)) /* RGBA16F; i=0 is U_0 (the bloom buffer)          */
#define ASH_BUF_VISIBILITY     1  /* u32[object_count]                    (binding 3)         */
#define ASH_BUF_DRAWS_OUT      2  /* AshIndirectDrawCommand[visible]      (binding 4)         */
#define ASH_BUF_VISIBLE_COUNT  3  /* u32                                  (binding 5)         */
#define ASH_BUF_TILE_DATA      4  /* u32[tiles*257]                       (light binding 2)   */
#define ASH_BUF_TAA_OUT        5  /* RGBA16F W x H (this frame's resolve = next history)      */
#define ASH_BUF_LDR8           6  /* RGBA8 W x H                                               */
#define ASH_BUF_STATS          7  /* AshViewStats                                              */
#define ASH_BUF_TILE_MINMAX    8  /* u32[tiles*2]: float bits of tile min/max depth            */
#define ASH_BUF_VISIBILITY_BITS 9 /* u32[ceil(n/32)] 1 bit per object                          */
#define ASH_BUF_BLOOM_THRESHOLD 10/* RGBA16F W x H: T = D_0 (staged path only)                 */
#define ASH_BUF_LOD_SELECTION  11 /* AshLodSelection[object_count]                              */
#define ASH_BUF_INSTANCES_OUT  12 /* AshInstanceData[total_instances], grouped by bucket, ascending object index inside */
#define ASH_BUF_BUCKET_OFFSETS 13 /* u32[max_batches*8 + 1]: bucket b = batch*8 + level starts at offsets[b]  */
#define ASH_BUF_INSTANCE_DRAWS 14 /* AshIndirectDrawCommand[max_batches*8]: one per bucket      */
#define ASH_BUF_INSTANCING_STATS 15 /* AshInstancingStats                                        */
#define ASH_BUF_FORWARD_KEYS   16 /* u64 W x H: (depth bits << 32) | object index, ASH_FWD_NO_OBJECT where none */
#define ASH_BUF_CULLING_STATS  18 /* AshCullingStats                                                              */
#define ASH_BUF_VELOCITY      19    /* RG16F motion vectors: set_color's, or the forward pass's (ashcuda_set_forward_reprojection) */
#define ASH_BUF_COLOR          17 /* RGBA16F W x H: the current HDR colour input (set_color / forward_shade)    */

/* Kernel classes for ashcuda_profile_kernels / ashcuda_kernel_time_ms (device-time accounting;
 * the Vulkan-timestamp TimingScopes of diagnostics/gpu_profiler.rs:25-44 play this role upstream). */
#define ASH_KERNEL_DEPTH_PYRAMID    0  /* Hi-Z mips + tile min/max, one pass over depth          */
#define ASH_KERNEL_OCCLUSION_CULL   1
#define ASH_KERNEL_CULL_COMPACT     2
#define ASH_KERNEL_LIGHT_PREPASS    3
#define ASH_KERNEL_LIGHT_TILES      4
#define ASH_KERNEL_TAA              5
#define ASH_KERNEL_BLOOM_THRESHOLD  6
#define ASH_KERNEL_BLOOM_DOWNSAMPLE 7
#define ASH_KERNEL_BLOOM_UPSAMPLE   8
#define ASH_KERNEL_TONEMAP          9
#define ASH_KERNEL_TAA_THRESHOLD   10  /* fused: TAA resolve + bloom threshold                   */
#define ASH_KERNEL_UPSAMPLE_TONEMAP 11 /* fused: last bloom upsample + tonemap + 8-bit pack      */
#define ASH_KERNEL_INSTANCING       12 /* LOD select + bucket scan + instance scatter (3 launches)  */
#define ASH_KERNEL_FORWARD_DEPTH    13 /* setup + clear + sphere splat (small, big) + depth resolve     */
#define ASH_KERNEL_FORWARD_SHADE    14 /* tiled-forward shading from the tile light lists              */
#define ASH_KERNEL_BLOOM_TAIL       15 /* bloom down / up launches whose destination has < 2^20 texels (levels >= 2 at 4K): a few MB
                                       each, bound by launch latency, not by bandwidth; kept apart from the full-size levels */
#define ASH_KERNEL_COUNT           16

typedef struct ashcuda_ctx ashcuda_ctx;

/* --- lifetime: replaces LightCullingPipeline::new / Drop (vulkan/light_culling_pipeline.rs:41-243,
 *     332-351), ComputePipeline::new (vulkan/compute_pipeline.rs:17-64), HdrFramebuffer::new
 *     (renderer/hdr_framebuffer.rs:25-110). */
void ashcuda_default_config(AshCudaConfig* cfg);
int  ashcuda_create(ashcuda_ctx** out, int device, uint32_t width, uint32_t height,
                    const AshCudaConfig* cfg_or_null);
void ashcuda_destroy(ashcuda_ctx* ctx);                       /* NULL is a no-op */
int  ashcuda_set_stream(ashcuda_ctx* ctx, void* cuda_stream); /* adopt a caller stream (NULL = own) */

/* --- inputs.  on_device=0: host pointer, copied with cudaMemcpyAsync on the ctx stream.
 *     on_device=1: device pointer, BORROWED (zero-copy) until replaced; caller keeps it alive.
 *     Replaces upload_lights / upload_camera (light_culling_pipeline.rs:248-281), the descriptor
 *     writes of :161-205, and OcclusionCulling::object_data uploads (occlusion_culling.rs:247-249). */
int ashcuda_set_depth  (ashcuda_ctx*, const float* depth_wh, int on_device);
int ashcuda_set_objects(ashcuda_ctx*, const AshCullObjectData* objs, uint32_t n,
                        const AshIndirectDrawCommand* draws_in, uint32_t n_draws, int on_device);
int ashcuda_set_lights (ashcuda_ctx*, const AshGpuLight* lights, uint32_t n, int on_device);
int ashcuda_set_camera (ashcuda_ctx*, const AshCullingCameraData* cam);
int ashcuda_set_color  (ashcuda_ctx*, const void* hdr_rgba16f, const void* velocity_rg16f_or_null,
                        int on_device);
/* TemporalAA::reset_history (temporal_aa.rs:203-207): next resolve uses current colour as history */
int ashcuda_taa_reset_history(ashcuda_ctx*);

/* --- stages: one per reference shader dispatch. */
int ashcuda_hiz_build     (ashcuda_ctx*);                                        /* hiz_generate.comp x levels */
int ashcuda_occlusion_cull(ashcuda_ctx*, const AshCullingPushConstants*);        /* occlusion_cull.comp        */
int ashcuda_light_cull    (ashcuda_ctx*, const AshLightCullingPushConstants*);   /* LightCullingPipeline::dispatch, :286-319 */
int ashcuda_taa_resolve   (ashcuda_ctx*, const AshTaaPushConstants*);            /* taa_resolve.comp           */
int ashcuda_bloom         (ashcuda_ctx*, const AshBloomParams*);                 /* bloom_{threshold,downsample,upsample}.frag */
int ashcuda_tonemap       (ashcuda_ctx*, const AshPostProcessPushConstants*);    /* tonemapping.frag           */
/* TonemappingConfig.operator / TonemappingFeature::set_operator (features/tonemapping.rs:9-20, 92-94).  The reference
 * has a shader for Aces only (tonemapping.frag:20-27); the other three are the published closed forms, pinned here
 * (PIN-13) and restated identically by the oracle, applied where tonemapping.frag:44 applies aces(), gamma after:
 *   ASH_TONEMAP_REINHARD    clamp(x / (1 + x), 0, 1)
 *   ASH_TONEMAP_UNCHARTED2  clamp(h(2 x) / h(11.2), 0, 1),  h(x) = (x (A x + C B) + D E) / (x (A x + B) + D F) - E / F
 *                           with A .15, B .50, C .10, D .20, E .02, F .30 (Hable 2010: exposure bias 2, white point 11.2)
 *   ASH_TONEMAP_NONE        clamp(x, 0, 1)
 * Applies to ashcuda_tonemap and ashcuda_view_run from the next call on; default ASH_TONEMAP_ACES. */
#define ASH_TONEMAP_ACES       0u
#define ASH_TONEMAP_REINHARD   1u
#define ASH_TONEMAP_UNCHARTED2 2u
#define ASH_TONEMAP_NONE       3u
int ashcuda_set_tonemap_operator(ashcuda_ctx*, uint32_t op);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  The ordered draw-command
 * compaction of ashcuda_occlusion_cull as a context-free call that ALLOCATES NOTHING: its scratch (visible objects per
 * CTA of 256) is a file-scope static __device__ array of ASH_STATIC_COMPACT_MAX_OBJECTS / 256 words = 16 KB, which every
 * GPU that loads the library holds until exit (csrc/cull_compact_static.cu).
 *   in  (device pointers): objs[object_count], draws_in[n_draws], visibility[object_count] (non-zero = visible)
 *   out (device pointers): draws_out[>= number of visible objects] = draws_in[objs[i].draw_index] of the visible objects in
 *        ascending object index (zeros where draw_index >= n_draws), *visible_count — the values ashcuda_occlusion_cull
 *        leaves in ASH_BUF_DRAWS_OUT / ASH_BUF_VISIBLE_COUNT for the same visibility.
 * Enqueues on `cuda_stream` (a cudaStream_t, NULL = the default stream) of `device` and returns; the caller synchronises.
 * object_count > ASH_STATIC_COMPACT_MAX_OBJECTS is rejected with ASH_ERR_INVALID.  err / err_len: optional buffer for the
 * message (there is no context to ask).  All calls on a GPU SHARE the one array and find the previous call's contents in
 * it (each call overwrites the entries it reads, nothing needs clearing); calls that OVERLAP on a GPU race: serialise. */
#define ASH_STATIC_COMPACT_MAX_OBJECTS 1048576u
int ashcuda_cull_compact_static(int device, void* cuda_stream, const AshCullObjectData* objs,
                                const AshIndirectDrawCommand* draws_in, uint32_t n_draws, const uint32_t* visibility,
                                uint32_t object_count, AshIndirectDrawCommand* draws_out, uint32_t* visible_count,
                                char* err, size_t err_len);
/* the whole view-frame in one call, fused kernels, same results as the staged calls */
int ashcuda_view_run      (ashcuda_ctx*, const AshCudaViewParams*);

/* ---- (f-2) LOD selection + instance batching on the device ------------------------------------------
 * One record per cull object (same index): its InstanceData, a dense batch id (the host's BatchKey ->
 * id map, instancing.rs:74-90) and the LodMesh it draws.  ashcuda_instancing_finalize does, for every
 * (visible) object in ascending index, what the CPU managers do per frame:
 *   sel = LodManager::select_lod(mesh, instance.position(), view_proj, camera_pos, w, h)   lod_system.rs:238-345
 *   if !sel.culled: InstancingManager::add_instance(key(batch, sel.level), instance)        instancing.rs:212-224
 * then InstancingManager::finalize (:253-263): stats, plus one indirect draw per (batch, level) bucket
 * {index_count, instance_count, first_index = level.index_offset, 0, first_instance = bucket offset}. */
int ashcuda_set_instances (ashcuda_ctx*, const AshInstanceData* instances, const uint32_t* batch_id,
                           const uint32_t* lod_mesh_id, uint32_t n, int on_device);
int ashcuda_set_lod_meshes(ashcuda_ctx*, const AshLodMesh* meshes, uint32_t n, int on_device);
int ashcuda_set_lod_config(ashcuda_ctx*, const AshLodConfig*);      /* NULL: LodConfig::default(), :47-57 */
int ashcuda_instancing_finalize(ashcuda_ctx*, const AshInstancingParams*);

/* ---- (f-1) depth / HDR producers (csrc/forward.cu) --------------------------------------------------
 * forward_depth ray-casts one occluder sphere per enabled cull object (centre = model * bounds.center,
 * radius = radius_scale * min_k(extents_k * |model column k|), inside the object's world AABB) through
 * the pixel centres, keeps per pixel the nearest hit as (ndc depth, object index) and writes the depth
 * buffer the rest of the path consumes (as set_depth would).  forward_shade evaluates frag.frag's
 * Cook-Torrance BRDF (:89-114, :167-213) at the hit for the MVP light and, with ASH_FWD_TILE_LIGHTS,
 * for every light of the pixel's 16x16 tile list (the consumer light_culling.comp was written for and
 * frag.frag never got), and writes the HDR colour buffer (as set_color would). */
int ashcuda_set_forward_params(ashcuda_ctx*, const AshForwardParams*);   /* NULL: defaults */
/* Velocity output of the forward pass (the velocity buffer binding 2 of taa_resolve.comp:14 reads, :94-102): with a
 * reprojection matrix set, ashcuda_forward_shade also writes RG16F velocity = uv - prevUV per pixel, prevUV from
 * reproj * (view-space hit, 1), reproj = prevViewProj * inverse(view) as 16 floats in columns (the caller builds it like
 * viewProj, renderer.rs:1645); the TAA of the same view then takes its velocity branch.  Static scene: CullObjectData
 * carries no previous model matrix.  Pixels without a hit get 0 (jitter-delta fallback, :97-99).  NULL switches it off. */
int ashcuda_set_forward_reprojection(ashcuda_ctx*, const float* reproj16_or_null);
int ashcuda_forward_depth(ashcuda_ctx*, const AshCullingCameraData* cam, uint32_t object_count);
int ashcuda_forward_shade(ashcuda_ctx*);   /* after forward_depth (and light_cull with ASH_FWD_TILE_LIGHTS) of the same view */

/* --- whole-view host I/O for callers whose frame data lives in host memory (the e2e path):
 *     upload = set_depth + set_color from host pointers (pinned memory makes the copies async);
 *     download = async device->host copies of the view's results on the ctx stream; any pointer may
 *     be NULL to skip that buffer.  Sizes: visibility 4*object_count, tile_data 4*257*tiles,
 *     ldr8 4*W*H, stats sizeof(AshViewStats), visible_count 4. Call ashcuda_sync before reading. */
int ashcuda_upload_view  (ashcuda_ctx*, const float* depth_wh, const void* hdr_rgba16f,
                          const void* velocity_rg16f_or_null);
int ashcuda_download_view(ashcuda_ctx*, uint32_t* visibility, uint32_t* tile_data, uint8_t* ldr8,
                          AshViewStats* stats, uint32_t* visible_count);
/* The same with the tile light lists in CSR form instead of the reference's 257-word-per-tile table
 * (light_culling.comp:43-45 / light_culling_pipeline.rs:55-56: 33 MB at 4K, nearly all unused slots):
 * tile_offsets[tiles + 1] = exclusive scan of the per-tile counts (tile_offsets[tiles] = total), tile_lights =
 * the light indices of tile 0, tile 1, ... back to back.  Exactly tile_lights_capacity words are copied into
 * tile_lights; if tile_offsets[tiles] > tile_lights_capacity the lists are truncated there and the caller
 * fetches the full table (ashcuda_download_view / ASH_BUF_TILE_DATA) instead.  tile_offsets and tile_lights
 * may both be NULL to skip the lists.  Two small kernels on the ctx stream build the packed form. */
int ashcuda_download_view_packed(ashcuda_ctx*, uint32_t* visibility, uint32_t* tile_offsets, uint32_t* tile_lights,
                                 uint32_t tile_lights_capacity, uint8_t* ldr8, AshViewStats* stats,
                                 uint32_t* visible_count);

/* --- outputs: device pointers owned by ctx, valid until the next call that rewrites them.
 *     Replaces LightCullingPipeline::tile_buffer / tile_buffer_size (:322-330). */
int ashcuda_get     (ashcuda_ctx*, int which, void** dev_ptr, size_t* bytes);
int ashcuda_readback(ashcuda_ctx*, int which, void* host_dst, size_t bytes);  /* syncs the stream */
int ashcuda_readback_async(ashcuda_ctx*, int which, void* host_dst, size_t bytes); /* no sync */
int ashcuda_sync    (ashcuda_ctx*);
int ashcuda_hiz_level_count(ashcuda_ctx*);                     /* levels_built (PIN-4) */
int ashcuda_mip_extent(ashcuda_ctx*, int which, uint32_t* w, uint32_t* h);
uint64_t ashcuda_kernel_launches(ashcuda_ctx*);                /* kernels launched so far by this ctx */
/* view_run graph replay (AshCudaConfig.graph_replay): views captured into a CUDA graph / views replayed from one */
int ashcuda_graph_stats(ashcuda_ctx*, uint64_t* captures, uint64_t* replays);
/* Device-time accounting: launches of the classes in `mask` (bit ASH_KERNEL_*) are bracketed by
 * CUDA events on the ctx stream.  ashcuda_kernel_time_ms syncs the stream, returns the time and
 * launch count accumulated for one class since the previous query of it, and resets them.
 * While any class is profiled, view_run does not fork its front end (one kernel at a time per ctx). */
int ashcuda_profile_kernels(ashcuda_ctx*, uint32_t mask);
int ashcuda_kernel_time_ms(ashcuda_ctx*, int kernel_class, float* ms, uint64_t* launches);
/* Diagnostics (host only, no device needed): does the pinned LINEAR sampler land exactly on the
 * texels/weights the fast kernels assume, for every coordinate of one axis?  kind 0: centre taps
 * (n_dst texels, size_f = extent as float); 1: TAA 3x3 taps (size_f, inv = screenParams.z|w);
 * 2: 2:1 downsample (n_src -> n_dst); 3: 1:2 upsample (n_src -> n_dst); 4: are the TAA history taps of a frame
 * without velocity buffer "regular" on this axis (one texel offset and one weight for every coordinate; inv = the
 * jitter delta in pixels) — the precondition of the shared-row history path of the FAST precision mode.  Returns 1 / 0. */
int ashcuda_debug_exact_taps(int kind, int n_src, int n_dst, float size_f, float inv);
/* shaders/compute_utils.glsl as a CUDA header (csrc/ash_compute_utils.cuh; row f-4): runs one function of
 * the library on host data on `device` (no context), for parity tests.  op:
 *   ASH_CU_MIN / MAX / SUM   in: float[n], n % 256 == 0  -> out: float[n/256]   parallelMin/Max/Sum, :22-65
 *   ASH_CU_SCAN              in: u32[n],   n % 256 == 0  -> out: u32[n]         exclusivePrefixSum per 256, :80-111
 *   ASH_CU_SPHERE_IN_FRUSTUM in: float[n][4] (centre, radius), params[0..15] = view_proj columns -> out: u32[n], :171-178
 *   ASH_CU_AABB_IN_FRUSTUM   in: float[n][6] (min, max), same params -> out: u32[n], :188-195
 *   ASH_CU_DEPTH             in: float[n][3] (u, v, depth), params = inv_proj columns, near, far
 *                            -> out: float[n][4] = linearizeDepth, uvDepthToView xyz, :205-219
 *   ASH_CU_PLANES            params = view_proj columns -> out: float[6][4] extractPlane(face 0..5), :126-160 */
#define ASH_CU_MIN 0
#define ASH_CU_MAX 1
#define ASH_CU_SUM 2
#define ASH_CU_SCAN 3
#define ASH_CU_SPHERE_IN_FRUSTUM 4
#define ASH_CU_AABB_IN_FRUSTUM 5
#define ASH_CU_DEPTH 6
#define ASH_CU_PLANES 7
int ashcuda_debug_compute_utils(int device, int op, const void* in, uint32_t n, const float* params18, void* out);
const char* ashcuda_last_error(ashcuda_ctx*);                  /* NULL ctx: last create error */
const char* ashcuda_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ASHCUDA_H */
