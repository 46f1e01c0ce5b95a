//! CudaComputeFeature: RenderFeature — the drop-in plugin (features/feature_trait.rs:25-34).
//! UNBUILT (no Rust toolchain in the image); mirrors include/ashcuda_host.hpp line for line.
use super::ffi::*;
use crate::error::{AshError, Result};
use crate::renderer::features::bloom::BloomConfig;
use crate::renderer::features::light_culling::{CullingCameraData, LightCullingPass};
use crate::renderer::features::tonemapping::TonemappingConfig;
use crate::renderer::features::{FeatureFrameContext, FeatureRenderContext, RenderFeature};
use crate::renderer::occlusion_culling::{IndirectDrawCommand, OcclusionCulling};
use crate::renderer::temporal_aa::TemporalAA;
use glam::{Mat4, Vec3};
use std::ffi::CStr;

/// What the reference's contexts do not carry (feature_trait.rs:10-23): set before render_frame.
pub struct FrameInputs<'a> {
    pub view: Mat4, pub projection: Mat4, pub camera_pos: Vec3,
    pub depth: &'a [f32], pub hdr_rgba16f: Option<&'a [u16]>, pub view_id: u32,
}

pub struct CudaComputeFeature {
    pub culling: OcclusionCulling, pub lights: LightCullingPass, pub taa: TemporalAA,
    pub bloom: BloomConfig, pub tonemapping: TonemappingConfig, pub draw_templates: Vec<IndirectDrawCommand>,
    ctx: *mut ashcuda_ctx, width: u32, height: u32, cfg: AshCudaConfig, scene_dirty: bool, params: Option<AshCudaViewParams>,
}
unsafe impl Send for CudaComputeFeature {}   // one thread at a time, like every RenderFeature

fn check(ctx: *mut ashcuda_ctx, rc: i32) -> Result<()> {
    if rc == ASH_OK { return Ok(()); }
    let msg = unsafe { CStr::from_ptr(ashcuda_last_error(ctx)) }.to_string_lossy().into_owned();
    Err(match rc {
        ASH_ERR_NOT_INITIALIZED => AshError::FeatureNotInitialized(msg),
        ASH_ERR_NOT_FOUND => AshError::ResourceNotFound(msg),
        ASH_ERR_DEVICE | ASH_ERR_OUT_OF_MEMORY => AshError::DeviceInitFailed(msg),
        _ => AshError::VulkanError(msg),
    })
}

impl CudaComputeFeature {
    pub fn new(width: u32, height: u32, max_objects: u32, max_lights: u32, hiz_levels: u32) -> Self {
        let mut cfg = AshCudaConfig::default();
        unsafe { ashcuda_default_config(&mut cfg) };
        cfg.max_objects = max_objects; cfg.max_lights = max_lights; cfg.hiz_levels = hiz_levels;
        Self { culling: OcclusionCulling::new(), lights: LightCullingPass::new(), taa: TemporalAA::new(),
               bloom: BloomConfig::default(), tonemapping: TonemappingConfig::default(),
               draw_templates: vec![IndirectDrawCommand::default()], ctx: std::ptr::null_mut(), width, height, cfg,
               scene_dirty: true, params: None }
    }
    /// Additive API: hand the frame's camera + buffers over before `Renderer::render_frame`.
    pub fn set_frame_inputs(&mut self, fi: &FrameInputs) -> Result<()> {
        if self.ctx.is_null() { return Err(AshError::FeatureNotInitialized("CudaComputeFeature before on_added".into())); }
        check(self.ctx, unsafe { ashcuda_set_depth(self.ctx, fi.depth.as_ptr(), 0) })?;
        if let Some(hdr) = fi.hdr_rgba16f {
            check(self.ctx, unsafe { ashcuda_set_color(self.ctx, hdr.as_ptr().cast(), std::ptr::null(), 0) })?;
        }
        let view_proj = fi.projection * fi.view;                                     // renderer.rs:1645
        let mut cull = self.culling.push_constants(view_proj, self.width, self.height);
        cull.hiz_levels = cull.hiz_levels.min(unsafe { ashcuda_hiz_level_count(self.ctx) } as u32);
        let camera = CullingCameraData {                                             // light_culling_integration.rs:112-123
            view: fi.view.to_cols_array_2d(), projection: fi.projection.to_cols_array_2d(),
            inv_projection: fi.projection.inverse().to_cols_array_2d(),
            camera_pos: [fi.camera_pos.x, fi.camera_pos.y, fi.camera_pos.z, 1.0] };
        let mut stages = 0x01 | if self.culling.is_enabled() { 0x02 } else { 0 } | if self.lights.is_enabled() { 0x04 } else { 0 };
        if fi.hdr_rgba16f.is_some() {
            stages |= 0x20 | if self.taa.is_enabled() { 0x08 } else { 0 } | if self.bloom.enabled { 0x10 } else { 0 };
        }
        self.params = Some(AshCudaViewParams {
            cull, light: self.lights.get_push_constants(self.width, self.height), camera,
            taa: self.taa.push_constants(self.width, self.height),
            bloom: AshBloomParams { threshold: self.bloom.threshold, upsample_intensity: 1.0, mip_count: self.bloom.mip_count, _padding: 0 },
            post: crate::renderer::fullscreen_pass::PostProcessPushConstants {
                exposure: self.tonemapping.exposure, gamma: self.tonemapping.gamma,
                bloom_intensity: if self.bloom.enabled { self.bloom.intensity } else { 0.0 }, _padding: 0.0 },
            view_id: fi.view_id, stages, _padding: [0; 2] });
        Ok(())
    }
    pub fn visible_count(&mut self) -> Result<u32> {
        let mut n = 0u32;
        check(self.ctx, unsafe { ashcuda_readback(self.ctx, ASH_BUF_VISIBLE_COUNT, (&mut n as *mut u32).cast(), 4) })?;
        self.culling.update_stats(n);
        Ok(n)
    }
}

impl RenderFeature for CudaComputeFeature {
    fn name(&self) -> &'static str { "CudaCompute" }
    fn on_added(&mut self, _device: &ash::Device) {
        self.cfg.bloom_mips = self.bloom.mip_count;
        let rc = unsafe { ashcuda_create(&mut self.ctx, 0, self.width, self.height, &self.cfg) };
        if rc != ASH_OK { log::error!("ashcuda_create failed ({rc}): no CUDA device; there is no CPU fallback"); self.ctx = std::ptr::null_mut(); return; }
        self.lights.calculate_tiles(self.width, self.height);
    }
    fn before_frame(&mut self, _ctx: &mut FeatureFrameContext<'_>) {
        if self.ctx.is_null() { return; }
        self.taa.begin_frame();
        if self.scene_dirty {
            let o = self.culling.object_data();
            let l = self.lights.get_light_buffer_data();
            unsafe {
                ashcuda_set_objects(self.ctx, o.as_ptr(), o.len() as u32, self.draw_templates.as_ptr(), self.draw_templates.len() as u32, 0);
                ashcuda_set_lights(self.ctx, l.as_ptr(), l.len() as u32, 0);
            }
            self.scene_dirty = false;
        }
    }
    unsafe fn render(&self, _ctx: &FeatureRenderContext<'_>) {
        if let (false, Some(vp)) = (self.ctx.is_null(), self.params.as_ref()) {
            ashcuda_set_tonemap_operator(self.ctx, self.tonemapping.operator as u32);   // Aces, Reinhard, Uncharted2, None = 0..3
            if let Err(e) = check(self.ctx, ashcuda_view_run(self.ctx, vp)) { log::error!("ashcuda_view_run: {e}"); }
        }
    }
    fn on_removed(&mut self, _device: &ash::Device) { unsafe { ashcuda_destroy(self.ctx) }; self.ctx = std::ptr::null_mut(); }
}
