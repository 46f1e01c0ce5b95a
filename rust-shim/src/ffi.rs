//! extern "C" declarations of include/ashcuda.h against the crate's own #[repr(C)] types
//! (byte-identical layouts; sizes asserted below).  UNBUILT (no Rust toolchain in the image).
#![allow(non_camel_case_types)]
use crate::renderer::features::light_culling::{CullingCameraData, GpuLight, LightCullingPushConstants};
use crate::renderer::occlusion_culling::{CullObjectData, CullingPushConstants, IndirectDrawCommand};
use crate::renderer::temporal_aa::TaaPushConstants;
use crate::renderer::fullscreen_pass::PostProcessPushConstants;
use crate::renderer::instancing::InstanceData;                 // 96 B, instancing.rs:19-34 (passed by pointer unchanged)
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] pub struct ashcuda_ctx { _private: [u8; 0] }

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshBloomParams { pub threshold: f32, pub upsample_intensity: f32, pub mip_count: u32, pub _padding: u32 }

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshCudaConfig {
    pub max_objects: u32, pub max_lights: u32, pub max_draws: u32, pub hiz_levels: u32, pub bloom_mips: u32,
    pub enable_post: u32, pub srgb_encode: u32, pub concurrent_frontend: u32, pub max_batches: u32, pub graph_replay: u32, pub precision: u32, pub fp32_intermediates: u32, pub _reserved: [u32; 4],
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct AshCudaViewParams {
    pub cull: CullingPushConstants, pub light: LightCullingPushConstants, pub camera: CullingCameraData,
    pub taa: TaaPushConstants, pub bloom: AshBloomParams, pub post: PostProcessPushConstants,
    pub view_id: u32, pub stages: u32, pub _padding: [u32; 2],
}

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct AshViewStats {
    pub view_id: u32, pub total_objects: u32, pub visible_objects: u32, pub tiles: u32,
    pub sum_tile_lights: u32, pub max_tile_lights: u32, pub overflow_tiles: u32, pub candidate_lights: u32,
}

// ---- row f-2: LodManager / InstancingManager on the device.  LodMesh holds a Vec<LodLevel>, so the host
// flattens it: `AshLodMesh::from(&LodMesh)` copies up to MAX_LOD_LEVELS levels.
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshLodLevel { pub screen_threshold: f32, pub distance_threshold: f32, pub triangle_count: u32,
                         pub vertex_count: u32, pub index_offset: u32, pub index_count: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshLodMesh { pub levels: [AshLodLevel; 8], pub level_count: u32, pub bounding_radius: f32, pub lod_bias: f32, pub _padding: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshLodConfig { pub mode: u32, pub force_level: u32, pub smooth_transitions: u32, pub transition_width: f32,
                          pub global_bias: f32, pub cull_threshold: f32, pub use_visibility: u32, pub _padding: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshLodSelection { pub level: u32, pub blend: f32, pub culled: u32, pub screen_coverage: f32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshInstancingStats { pub objects_per_lod: [u32; 8], pub objects_culled: u32, pub total_instances: u32, pub batch_count: u32,
                                pub avg_instances_per_batch: f32, pub triangles_before: u64, pub triangles_after: u64,
                                pub instances_dropped: u32, pub _padding: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshInstancingParams { pub view_proj: [[f32; 4]; 4], pub camera_pos: [f32; 4], pub screen_size: [f32; 2],
                                 pub object_count: u32, pub _padding: u32 }
const _: () = assert!(std::mem::size_of::<InstanceData>() == 96 && std::mem::size_of::<AshLodMesh>() == 208);
const _: () = assert!(std::mem::size_of::<AshInstancingStats>() == 72 && std::mem::size_of::<AshInstancingParams>() == 96);
/// frag.frag MVP / Material uniform fields used by the compute forward pass (row f-1)
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshForwardParams { pub light_direction: [f32; 4], pub light_color: [f32; 4], pub ambient_color: [f32; 4],
                              pub base_color_factor: [f32; 4], pub emissive_factor: [f32; 4], pub parameters: [f32; 4],
                              pub background: [f32; 4], pub near_plane: f32, pub radius_scale: f32, pub flags: u32, pub _padding: u32 }
const _: () = assert!(std::mem::size_of::<AshForwardParams>() == 128);
pub const ASH_FWD_TILE_LIGHTS: u32 = 1;
pub const ASH_FWD_REINHARD: u32 = 2;
pub const ASH_FWD_INSTANCE_COLOR: u32 = 4;
/// CullingStats (render_stats.rs:54-65) + early / late counts of two-phase culling
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct AshCullingStats { pub objects_submitted: u32, pub after_frustum: u32, pub after_occlusion: u32, pub lights_culled: u32,
                             pub lod_triangles_saved: u64, pub early_visible: u32, pub late_visible: u32 }
const _: () = assert!(std::mem::size_of::<AshCullingStats>() == 32);
pub const ASH_STAGE_TWO_PHASE: u32 = 0x400;
pub const ASH_BUF_CULLING_STATS: c_int = 18;
pub const ASH_STAGE_FORWARD_DEPTH: u32 = 0x100;
pub const ASH_STAGE_FORWARD_SHADE: u32 = 0x200;
pub const ASH_BUF_FORWARD_KEYS: c_int = 16;
pub const ASH_BUF_COLOR: c_int = 17;
pub const ASH_STAGE_INSTANCING: u32 = 0x80;
pub const ASH_BUF_LOD_SELECTION: c_int = 11;
pub const ASH_BUF_INSTANCES_OUT: c_int = 12;
pub const ASH_BUF_BUCKET_OFFSETS: c_int = 13;
pub const ASH_BUF_INSTANCE_DRAWS: c_int = 14;
pub const ASH_BUF_INSTANCING_STATS: c_int = 15;

const _: () = assert!(std::mem::size_of::<CullObjectData>() == 112 && std::mem::size_of::<IndirectDrawCommand>() == 20);
const _: () = assert!(std::mem::size_of::<CullingPushConstants>() == 96 && std::mem::size_of::<GpuLight>() == 64);
const _: () = assert!(std::mem::size_of::<CullingCameraData>() == 208 && std::mem::size_of::<TaaPushConstants>() == 48);
const _: () = assert!(std::mem::size_of::<AshCudaViewParams>() == 416 && std::mem::size_of::<AshCudaConfig>() == 64);

pub const ASH_OK: c_int = 0;
pub const ASH_ERR_DEVICE: c_int = -1;
pub const ASH_ERR_NOT_INITIALIZED: c_int = -2;
pub const ASH_ERR_NOT_FOUND: c_int = -3;
pub const ASH_ERR_INVALID: c_int = -4;
pub const ASH_ERR_OUT_OF_MEMORY: c_int = -5;
pub const ASH_STAGE_ALL: u32 = 0x3F;
pub const ASH_BUF_VISIBILITY: c_int = 1;
pub const ASH_BUF_DRAWS_OUT: c_int = 2;
pub const ASH_BUF_VISIBLE_COUNT: c_int = 3;
pub const ASH_BUF_TILE_DATA: c_int = 4;
pub const ASH_BUF_TAA_OUT: c_int = 5;
pub const ASH_BUF_LDR8: c_int = 6;
pub const ASH_BUF_STATS: c_int = 7;

extern "C" {
    pub fn ashcuda_default_config(cfg: *mut AshCudaConfig);
    pub fn ashcuda_create(out: *mut *mut ashcuda_ctx, device: c_int, width: u32, height: u32, cfg: *const AshCudaConfig) -> c_int;
    pub fn ashcuda_destroy(ctx: *mut ashcuda_ctx);
    pub fn ashcuda_set_stream(ctx: *mut ashcuda_ctx, cuda_stream: *mut c_void) -> c_int;
    pub fn ashcuda_set_depth(ctx: *mut ashcuda_ctx, depth: *const f32, on_device: c_int) -> c_int;
    pub fn ashcuda_set_objects(ctx: *mut ashcuda_ctx, objs: *const CullObjectData, n: u32, draws_in: *const IndirectDrawCommand, n_draws: u32, on_device: c_int) -> c_int;
    pub fn ashcuda_set_lights(ctx: *mut ashcuda_ctx, lights: *const GpuLight, n: u32, on_device: c_int) -> c_int;
    pub fn ashcuda_set_camera(ctx: *mut ashcuda_ctx, cam: *const CullingCameraData) -> c_int;
    pub fn ashcuda_set_color(ctx: *mut ashcuda_ctx, hdr_rgba16f: *const c_void, velocity_rg16f: *const c_void, on_device: c_int) -> c_int;
    pub fn ashcuda_taa_reset_history(ctx: *mut ashcuda_ctx) -> c_int;
    pub fn ashcuda_hiz_build(ctx: *mut ashcuda_ctx) -> c_int;
    pub fn ashcuda_occlusion_cull(ctx: *mut ashcuda_ctx, pc: *const CullingPushConstants) -> c_int;
    pub fn ashcuda_light_cull(ctx: *mut ashcuda_ctx, pc: *const LightCullingPushConstants) -> c_int;
    pub fn ashcuda_taa_resolve(ctx: *mut ashcuda_ctx, pc: *const TaaPushConstants) -> c_int;
    pub fn ashcuda_bloom(ctx: *mut ashcuda_ctx, bp: *const AshBloomParams) -> c_int;
    pub fn ashcuda_tonemap(ctx: *mut ashcuda_ctx, pc: *const PostProcessPushConstants) -> c_int;
    pub fn ashcuda_set_tonemap_operator(ctx: *mut ashcuda_ctx, op: u32) -> c_int;
    pub fn ashcuda_set_forward_reprojection(ctx: *mut ashcuda_ctx, reproj16_or_null: *const f32) -> c_int;
    /// no counterpart in the reference: allocation-free ordered compaction, static 16 KB scratch per GPU
    pub fn ashcuda_cull_compact_static(device: c_int, cuda_stream: *mut c_void, objs: *const CullObjectData, draws_in: *const IndirectDrawCommand,
                                       n_draws: u32, visibility: *const u32, object_count: u32, draws_out: *mut IndirectDrawCommand,
                                       visible_count: *mut u32, err: *mut c_char, err_len: usize) -> c_int;
    pub fn ashcuda_view_run(ctx: *mut ashcuda_ctx, vp: *const AshCudaViewParams) -> c_int;
    pub fn ashcuda_set_instances(ctx: *mut ashcuda_ctx, instances: *const InstanceData, batch_id: *const u32, lod_mesh_id: *const u32, n: u32, on_device: c_int) -> c_int;
    pub fn ashcuda_set_lod_meshes(ctx: *mut ashcuda_ctx, meshes: *const AshLodMesh, n: u32, on_device: c_int) -> c_int;
    pub fn ashcuda_set_lod_config(ctx: *mut ashcuda_ctx, cfg: *const AshLodConfig) -> c_int;
    pub fn ashcuda_instancing_finalize(ctx: *mut ashcuda_ctx, ip: *const AshInstancingParams) -> c_int;
    pub fn ashcuda_set_forward_params(ctx: *mut ashcuda_ctx, params: *const AshForwardParams) -> c_int;
    pub fn ashcuda_forward_depth(ctx: *mut ashcuda_ctx, cam: *const CullingCameraData, object_count: u32) -> c_int;
    pub fn ashcuda_forward_shade(ctx: *mut ashcuda_ctx) -> c_int;
    pub fn ashcuda_debug_compute_utils(device: c_int, op: c_int, input: *const c_void, n: u32, params18: *const f32, out: *mut c_void) -> c_int;
    pub fn ashcuda_upload_view(ctx: *mut ashcuda_ctx, depth: *const f32, hdr: *const c_void, velocity: *const c_void) -> c_int;
    pub fn ashcuda_download_view(ctx: *mut ashcuda_ctx, visibility: *mut u32, tile_data: *mut u32, ldr8: *mut u8, stats: *mut AshViewStats, visible_count: *mut u32) -> c_int;
    pub fn ashcuda_download_view_packed(ctx: *mut ashcuda_ctx, visibility: *mut u32, tile_offsets: *mut u32, tile_lights: *mut u32, tile_lights_capacity: u32, ldr8: *mut u8, stats: *mut AshViewStats, visible_count: *mut u32) -> c_int;
    pub fn ashcuda_get(ctx: *mut ashcuda_ctx, which: c_int, dev_ptr: *mut *mut c_void, bytes: *mut usize) -> c_int;
    pub fn ashcuda_readback(ctx: *mut ashcuda_ctx, which: c_int, host_dst: *mut c_void, bytes: usize) -> c_int;
    pub fn ashcuda_readback_async(ctx: *mut ashcuda_ctx, which: c_int, host_dst: *mut c_void, bytes: usize) -> c_int;
    pub fn ashcuda_sync(ctx: *mut ashcuda_ctx) -> c_int;
    pub fn ashcuda_hiz_level_count(ctx: *mut ashcuda_ctx) -> c_int;
    pub fn ashcuda_mip_extent(ctx: *mut ashcuda_ctx, which: c_int, w: *mut u32, h: *mut u32) -> c_int;
    pub fn ashcuda_kernel_launches(ctx: *mut ashcuda_ctx) -> u64;
    pub fn ashcuda_graph_stats(ctx: *mut ashcuda_ctx, captures: *mut u64, replays: *mut u64) -> c_int;
    pub fn ashcuda_profile_kernels(ctx: *mut ashcuda_ctx, mask: u32) -> c_int;
    pub fn ashcuda_kernel_time_ms(ctx: *mut ashcuda_ctx, kernel_class: c_int, ms: *mut f32, launches: *mut u64) -> c_int;
    pub fn ashcuda_debug_exact_taps(kind: c_int, n_src: c_int, n_dst: c_int, size_f: f32, inv: f32) -> c_int;
    pub fn ashcuda_last_error(ctx: *mut ashcuda_ctx) -> *const c_char;
    pub fn ashcuda_version() -> *const c_char;
}
