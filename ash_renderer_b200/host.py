"""Host-side mirror of the reference's CPU state for the path (same names, argument meaning and
defaults), feeding the C ABI through `backend.Context`.

  OcclusionCulling / CullBoundingBox / OcclusionStats   src/renderer/occlusion_culling.rs:26-272
  GpuLight helpers / LightCullingPass                   src/renderer/features/light_culling.rs:28-243
  LightCullingIntegration.create_camera_data            src/renderer/light_culling_integration.rs:112-123
  HaltonSequence / TaaConfig / TemporalAA               src/renderer/temporal_aa.rs:14-207
  BloomConfig / TonemappingConfig                       features/bloom.rs:9-33, tonemapping.rs:22-44
  RenderFeature / FeatureManager                        features/feature_trait.rs:25-98
  CudaComputeFeature                                    the plugin a maintainer registers

The reference's host caps (MAX_LIGHTS 1024, MAX_CULLABLE_OBJECTS 65536, HIZ_LEVELS 10) are the
defaults and can be raised per instance: the shaders have no such caps.
"""
import logging

import numpy as np

from . import types as T

log = logging.getLogger("ash_renderer_b200")

MAX_CULLABLE_OBJECTS = 65536     # occlusion_culling.rs:18
HIZ_LEVELS = 10                  # occlusion_culling.rs:21
MAX_LIGHTS = 1024                # features/light_culling.rs:17
MAX_LIGHTS_PER_TILE = 256        # features/light_culling.rs:20
TILE_SIZE = 16                   # features/light_culling.rs:23
F32_MAX = float(np.finfo(np.float32).max)


def _f32(x):
    return np.float32(x)


# ------------------------------------------------------------------ occlusion culling
def cull_bounding_box_from_min_max(mn, mx):
    """CullBoundingBox::from_min_max (occlusion_culling.rs:34-42): center=(min+max)*0.5, extents=(max-min)*0.5."""
    mn, mx = np.asarray(mn, np.float32), np.asarray(mx, np.float32)
    c, e = (mn + mx) * _f32(0.5), (mx - mn) * _f32(0.5)
    return np.array([*c, 0.0], np.float32), np.array([*e, 0.0], np.float32)


def cull_bounding_box_from_center_extents(center, extents):
    return (np.array([*np.asarray(center, np.float32), 0.0], np.float32),
            np.array([*np.asarray(extents, np.float32), 0.0], np.float32))


class OcclusionStats:
    """occlusion_culling.rs:152-185"""

    def __init__(self):
        self.total_objects = 0
        self.after_frustum = 0
        self.after_occlusion = 0
        self.draws_culled = 0
        self.triangles_culled = 0

    def cull_rate(self):
        if self.total_objects == 0:
            return 0.0
        return 1.0 - self.after_occlusion / self.total_objects

    def format(self):
        return (f"Occlusion: {self.after_occlusion}/{self.total_objects} visible "
                f"({self.cull_rate() * 100.0:.1f}% culled), {self.draws_culled} draws saved")


class OcclusionCulling:
    """CPU-side state of the cull pass (occlusion_culling.rs:188-272)."""

    def __init__(self, max_objects=MAX_CULLABLE_OBJECTS, hiz_levels=HIZ_LEVELS):
        self.enabled = True
        self.frustum_only = False
        self.max_objects = max_objects
        self.hiz_levels = hiz_levels
        self._objects = np.zeros(1024, T.CULL_OBJECT)
        self._n = 0
        self._stats = OcclusionStats()

    def set_enabled(self, enabled):
        self.enabled = bool(enabled)

    def is_enabled(self):
        return self.enabled

    def set_frustum_only(self, frustum_only):
        self.frustum_only = bool(frustum_only)

    def begin_frame(self):
        self._n = 0
        self._stats = OcclusionStats()

    def add_object(self, bounds, model_cols, draw_index):
        """bounds = (center4, extents4); model_cols[c] = column c (Mat4::to_cols_array_2d)."""
        if self._n >= self.max_objects:
            log.warning("Occlusion culling: max object limit reached")
            return
        if self._n == len(self._objects):
            self._objects = np.concatenate([self._objects, np.zeros_like(self._objects)])
        o = self._objects[self._n]
        o["center"], o["extents"] = bounds
        o["model"] = np.asarray(model_cols, np.float32)
        o["draw_index"], o["lod_bias"], o["flags"], o["_padding"] = draw_index, 0.0, 1, 0
        self._n += 1

    def object_data(self):
        return self._objects[:self._n]

    def object_count(self):
        return self._n

    def push_constants(self, view_proj_cols, width, height):
        pc = np.zeros(1, T.CULLING_PC)
        pc["view_proj"][0] = np.asarray(view_proj_cols, np.float32)
        w, h = _f32(width), _f32(height)
        pc["screen_params"][0] = [w, h, _f32(1.0) / w, _f32(1.0) / h]
        pc["object_count"] = self._n
        pc["hiz_levels"] = self.hiz_levels
        return pc

    def update_stats(self, visible_count):
        self._stats.total_objects = self._n
        self._stats.after_occlusion = int(visible_count)
        self._stats.draws_culled = max(0, self._n - int(visible_count))

    def stats(self):
        return self._stats


# ------------------------------------------------------------------ light culling
class PointLight:
    """features/lighting.rs:22-39"""

    def __init__(self, position=(0.0, 0.0, 0.0), color=(1.0, 1.0, 1.0), intensity=1.0, radius=10.0):
        self.position, self.color, self.intensity, self.radius = position, color, intensity, radius


class DirectionalLight:
    """features/lighting.rs:6-20"""

    def __init__(self, direction=(0.0, -1.0, 0.0), color=(1.0, 1.0, 1.0), intensity=1.0):
        self.direction, self.color, self.intensity = direction, color, intensity


def gpu_light_from_point_light(l):
    g = np.zeros((), T.GPU_LIGHT)
    g["position"] = [*l.position, l.radius]
    g["color"] = [*l.color, l.intensity]
    g["direction"] = [0.0, 0.0, 0.0, 0.0]
    g["params"] = [0.0, 0.0, 1.0, 1.0]
    return g


def gpu_light_from_directional_light(l):
    g = np.zeros((), T.GPU_LIGHT)
    g["position"] = [0.0, 0.0, 0.0, F32_MAX]
    g["color"] = [*l.color, l.intensity]
    g["direction"] = [*l.direction, 2.0]
    g["params"] = [0.0, 0.0, 1.0, 1.0]
    return g


class LightCullingPass:
    """features/light_culling.rs:121-243"""

    def __init__(self, max_lights=MAX_LIGHTS):
        self.enabled = True
        self.debug_tiles = False
        self.max_lights = max_lights
        self._lights = np.zeros(0, T.GPU_LIGHT)
        self.tiles_x = self.tiles_y = 0
        self._last = (0, 0)

    def update_lights(self, point_lights, directional_lights=()):
        out = []
        for l in point_lights:
            if len(out) >= self.max_lights:
                log.warning("Light culling: exceeded max lights (%d)", self.max_lights)
                break
            out.append(gpu_light_from_point_light(l))
        for l in directional_lights:
            if len(out) >= self.max_lights:
                break
            out.append(gpu_light_from_directional_light(l))
        self._lights = np.array(out, T.GPU_LIGHT) if out else np.zeros(0, T.GPU_LIGHT)

    def set_light_buffer_data(self, lights):
        """Bulk path for already-packed GpuLight arrays (truncated to max_lights like update_lights)."""
        self._lights = np.ascontiguousarray(lights[:self.max_lights], T.GPU_LIGHT)

    def calculate_tiles(self, width, height):
        if (width, height) != self._last:
            self.tiles_x = -(-width // TILE_SIZE)
            self.tiles_y = -(-height // TILE_SIZE)
            self._last = (width, height)

    def get_dispatch_dimensions(self):
        return (self.tiles_x, self.tiles_y, 1)

    def get_push_constants(self, screen_width, screen_height):
        pc = np.zeros(1, T.LIGHT_PC)
        pc["screen_size"][0] = [screen_width, screen_height]
        pc["light_count"] = len(self._lights)
        return pc

    def get_light_buffer_data(self):
        return self._lights

    def light_count(self):
        return len(self._lights)

    def get_tile_buffer_size(self):
        return self.tiles_x * self.tiles_y * (MAX_LIGHTS_PER_TILE + 1) * 4

    def is_enabled(self):
        return self.enabled and len(self._lights) > 0


def create_camera_data(view_cols, projection_cols, camera_pos):
    """LightCullingIntegration::create_camera_data: inv_projection = projection.inverse()."""
    cd = np.zeros(1, T.CAMERA_DATA)
    cd["view"][0] = np.asarray(view_cols, np.float32)
    cd["projection"][0] = np.asarray(projection_cols, np.float32)
    p64 = np.asarray(projection_cols, np.float64)
    cd["inv_projection"][0] = np.linalg.inv(p64.T).T.astype(np.float32)
    cd["camera_pos"][0] = [*np.asarray(camera_pos, np.float32)[:3], 1.0]
    return cd


# ------------------------------------------------------------------ TAA
class TaaConfig:
    def __init__(self, enabled=True, blend_factor=0.9, color_clamp=True, velocity_rejection=True,
                 jitter_scale=1.0):
        self.enabled, self.blend_factor, self.color_clamp = enabled, blend_factor, color_clamp
        self.velocity_rejection, self.jitter_scale = velocity_rejection, jitter_scale


class HaltonSequence:
    """temporal_aa.rs:42-83: Halton(2,3) - 0.5, period 16, all arithmetic in f32."""

    def __init__(self):
        self.index = 0

    @staticmethod
    def halton(index, base):
        f, r = _f32(1.0), _f32(0.0)
        while index > 0:
            f = f / _f32(base)
            r = r + f * _f32(index % base)
            index //= base
        return r

    def next_jitter(self):
        j = (self.halton(self.index + 1, 2) - _f32(0.5), self.halton(self.index + 1, 3) - _f32(0.5))
        self.index = (self.index + 1) % 16
        return j

    def reset(self):
        self.index = 0


class TemporalAA:
    """temporal_aa.rs:111-207"""

    def __init__(self, config=None):
        self._config = config or TaaConfig()
        self.halton = HaltonSequence()
        self._cur = (_f32(0.0), _f32(0.0))
        self._prev = (_f32(0.0), _f32(0.0))
        self.frame_index = 0

    def begin_frame(self):
        self._prev = self._cur
        j = self.halton.next_jitter()
        s = _f32(self._config.jitter_scale)
        self._cur = (j[0] * s, j[1] * s)
        self.frame_index += 1

    def jitter_projection(self, projection_cols, width, height):
        p = np.array(projection_cols, np.float32)
        if not self._config.enabled:
            return p
        p[3][0] += self._cur[0] * _f32(2.0) / _f32(width)
        p[3][1] += self._cur[1] * _f32(2.0) / _f32(height)
        return p

    def push_constants(self, width, height):
        pc = np.zeros(1, T.TAA_PC)
        w, h = _f32(width), _f32(height)
        pc["screen_params"][0] = [w, h, _f32(1.0) / w, _f32(1.0) / h]
        c = self._config
        pc["params"][0] = [c.blend_factor, 1.0 if c.color_clamp else 0.0, 1.0 if c.velocity_rejection else 0.0, 0.0]
        pc["jitter"][0] = self._cur
        pc["prev_jitter"][0] = self._prev
        return pc

    def current_jitter(self):
        return self._cur

    def is_enabled(self):
        return self._config.enabled

    def config(self):
        return self._config

    def reset_history(self):
        self.halton.reset()
        self._cur = (_f32(0.0), _f32(0.0))
        self._prev = (_f32(0.0), _f32(0.0))


# ------------------------------------------------------------------ bloom / tonemap configs
class BloomConfig:
    """features/bloom.rs:9-33 (soft_knee is carried but the shader hard-codes 0.5, PIN-10)."""

    def __init__(self, threshold=0.8, intensity=0.5, mip_count=5, soft_knee=0.5, enabled=True):
        self.threshold, self.intensity, self.mip_count = threshold, intensity, mip_count
        self.soft_knee, self.enabled = soft_knee, enabled

    def params(self):
        bp = np.zeros(1, T.BLOOM_PARAMS)
        bp["threshold"], bp["upsample_intensity"], bp["mip_count"] = self.threshold, 1.0, self.mip_count
        return bp


class TonemappingConfig:
    """features/tonemapping.rs:22-44 (only the ACES operator has a shader)."""

    def __init__(self, exposure=1.0, gamma=2.2):
        self.operator, self.exposure, self.gamma = "Aces", exposure, gamma

    def push_constants(self, bloom_intensity):
        pc = np.zeros(1, T.POST_PC)
        pc["exposure"], pc["gamma"], pc["bloom_intensity"] = self.exposure, self.gamma, bloom_intensity
        return pc


# ------------------------------------------------------------------ plugin API
class RenderFeature:
    """features/feature_trait.rs:25-34"""

    def name(self):
        raise NotImplementedError

    def on_added(self, device):
        pass

    def before_frame(self, ctx):
        pass

    def render(self, ctx):
        pass

    def on_removed(self, device):
        pass


class FeatureManager:
    """features/feature_trait.rs:36-98: one instance per type, insertion order, on_added if a device is set."""

    def __init__(self):
        self._features, self._device = [], None

    def set_device(self, device):
        self._device = device

    def add_feature(self, feature):
        if any(type(f) is type(feature) for f in self._features):
            return
        if self._device is not None:
            feature.on_added(self._device)
        self._features.append(feature)

    def before_frame(self, ctx):
        for f in self._features:
            f.before_frame(ctx)

    def render(self, ctx):
        for f in self._features:
            f.render(ctx)

    def cleanup(self):
        if self._device is not None:
            for f in self._features:
                f.on_removed(self._device)
        self._features = []


class FrameInputs:
    """What the additive setter API hands the feature each frame (the reference's contexts carry no
    camera / depth / colour, feature_trait.rs:10-23)."""

    def __init__(self, view, projection, camera_pos, depth, hdr, velocity=None, view_id=0):
        self.view, self.projection, self.camera_pos = view, projection, camera_pos
        self.depth, self.hdr, self.velocity, self.view_id = depth, hdr, velocity, view_id


class CudaComputeFeature(RenderFeature):
    """The drop-in feature: owns a backend.Context plus the reference's CPU state objects and runs
    the whole per-frame compute path in `render`."""

    def __init__(self, width, height, max_objects=MAX_CULLABLE_OBJECTS, max_lights=MAX_LIGHTS,
                 hiz_levels=HIZ_LEVELS, bloom=None, tonemapping=None, taa=None, device=0):
        self.width, self.height, self._device_index = width, height, device
        self.culling = OcclusionCulling(max_objects, hiz_levels)
        self.lights = LightCullingPass(max_lights)
        self.taa = TemporalAA(taa)
        self.bloom = bloom or BloomConfig()
        self.tonemapping = tonemapping or TonemappingConfig()
        self.draw_templates = np.zeros(1, T.DRAW_COMMAND)
        self._ctx_args = dict(max_objects=max_objects, max_lights=max_lights, hiz_levels=hiz_levels,
                              bloom_mips=self.bloom.mip_count)
        self.ctx = None
        self.inputs = None
        self._scene_dirty = True

    def name(self):
        return "CudaCompute"

    def on_added(self, device):
        from . import backend
        self.ctx = backend.Context(self.width, self.height, self._device_index, **self._ctx_args)
        self.lights.calculate_tiles(self.width, self.height)

    def on_removed(self, device):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None

    def set_frame_inputs(self, inputs):
        self.inputs = inputs

    def mark_scene_dirty(self):
        self._scene_dirty = True

    def view_params(self):
        fi = self.inputs
        vp = np.zeros(1, T.VIEW_PARAMS)
        view_proj = (np.asarray(fi.projection, np.float64).T @ np.asarray(fi.view, np.float64).T).T
        vp["cull"] = self.culling.push_constants(view_proj, self.width, self.height)
        vp["cull"]["hiz_levels"] = min(self.culling.hiz_levels, self.ctx.hiz_level_count())
        vp["light"] = self.lights.get_push_constants(self.width, self.height)
        vp["camera"] = create_camera_data(fi.view, fi.projection, fi.camera_pos)
        vp["taa"] = self.taa.push_constants(self.width, self.height)
        vp["bloom"] = self.bloom.params()
        vp["post"] = self.tonemapping.push_constants(self.bloom.intensity if self.bloom.enabled else 0.0)
        vp["view_id"] = fi.view_id
        st = T.STAGE_HIZ
        if self.culling.is_enabled():
            st |= T.STAGE_CULL
        if self.lights.is_enabled():
            st |= T.STAGE_LIGHTS
        if fi.hdr is not None:
            if self.taa.is_enabled():
                st |= T.STAGE_TAA
            if self.bloom.enabled:
                st |= T.STAGE_BLOOM
            st |= T.STAGE_TONEMAP
        vp["stages"] = st
        return vp

    def before_frame(self, ctx=None):
        if self.ctx is None:
            from .backend import AshError, ASH_ERR_NOT_INITIALIZED
            raise AshError(ASH_ERR_NOT_INITIALIZED, "CudaComputeFeature used before on_added")
        self.taa.begin_frame()
        if self._scene_dirty:
            self.ctx.set_objects(self.culling.object_data(), self.draw_templates)
            self.ctx.upload_lights(self.lights.get_light_buffer_data())
            self._scene_dirty = False

    def render(self, ctx=None):
        fi = self.inputs
        self.ctx.set_depth(fi.depth)
        if fi.hdr is not None:
            self.ctx.set_color(fi.hdr, fi.velocity)
        self.ctx.view_run(self.view_params())
