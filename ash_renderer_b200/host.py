"""Host-side mirror of the reference's CPU state for the path (same names, argument meaning and
defaults), feeding the C ABI through `backend.Context`.

  OcclusionCulling / CullBoundingBox / OcclusionStats   src/renderer/occlusion_culling.rs:26-272
  GpuLight helpers / LightCullingPass                   src/renderer/features/light_culling.rs:28-243
  LightCullingIntegration.create_camera_data            src/renderer/light_culling_integration.rs:112-123
  HaltonSequence / TaaConfig / TemporalAA               src/renderer/temporal_aa.rs:14-207
  BloomConfig / TonemappingConfig                       features/bloom.rs:9-33, tonemapping.rs:22-44
  RenderFeature / FeatureManager                        features/feature_trait.rs:25-98
  InstanceData / BatchKey / InstancingManager           src/renderer/instancing.rs:19-277   (row f-2)
  LodLevel / LodMesh / LodConfig / LodManager           src/renderer/lod_system.rs:15-394    (row f-2)
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


class CullingStats:
    """render_stats.rs:54-65"""

    def __init__(self):
        self.objects_submitted = 0
        self.after_frustum = 0
        self.after_occlusion = 0
        self.lights_culled = 0
        self.lod_triangles_saved = 0


class StatsCollector:
    """The culling part of render_stats.rs:132-240 (record_culling :213-217, record_lod_savings :232-234)."""

    def __init__(self):
        self.culling = CullingStats()

    def record_culling(self, submitted, after_frustum, after_occlusion):
        self.culling.objects_submitted = int(submitted)
        self.culling.after_frustum = int(after_frustum)
        self.culling.after_occlusion = int(after_occlusion)

    def record_lod_savings(self, triangles_saved):
        self.culling.lod_triangles_saved = int(triangles_saved)

    def record_device(self, cs):
        """Feed one AshCullingStats record (Context.culling_stats()) through the reference's recorders."""
        self.record_culling(cs["objects_submitted"], cs["after_frustum"], cs["after_occlusion"])
        self.record_lod_savings(cs["lod_triangles_saved"])
        self.culling.lights_culled = int(cs["lights_culled"])
        return self.culling


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
    """features/tonemapping.rs:9-44.  `operator` is one of TonemapOperator's names; only Aces has a shader in the
    reference, the others follow PIN-13 (include/ashcuda.h)."""
    OPERATORS = {"Aces": T.TONEMAP_ACES, "Reinhard": T.TONEMAP_REINHARD, "Uncharted2": T.TONEMAP_UNCHARTED2, "None": T.TONEMAP_NONE}

    def __init__(self, exposure=1.0, gamma=2.2, operator="Aces", enabled=True):
        if operator not in self.OPERATORS:
            raise ValueError(f"unknown TonemapOperator {operator!r}")
        self.operator, self.exposure, self.gamma, self.enabled = operator, exposure, gamma, enabled

    def operator_id(self):
        return self.OPERATORS[self.operator]

    def push_constants(self, bloom_intensity):
        pc = np.zeros(1, T.POST_PC)
        pc["exposure"], pc["gamma"], pc["bloom_intensity"] = self.exposure, self.gamma, bloom_intensity
        return pc


# ------------------------------------------------------------------ row f-2: instancing + LOD
MAX_INSTANCES_PER_BATCH = 65536  # instancing.rs:16
MAX_LOD_LEVELS = 8               # lod_system.rs:15


def instance_data_from_matrix(model_cols, color=(1.0, 1.0, 1.0, 1.0), custom=(0.0, 0.0, 0.0, 0.0)):
    """InstanceData::from_matrix / ::new / with_custom (instancing.rs:37-66): model_row* hold COLUMNS."""
    d = np.zeros((), T.INSTANCE_DATA)
    m = np.asarray(model_cols, np.float32).reshape(4, 4)
    d["model_row0"], d["model_row1"], d["model_row2"], d["model_row3"] = m[0], m[1], m[2], m[3]
    d["color"], d["custom"] = color, custom
    return d


def instance_position(instance):
    """InstanceData::position (instancing.rs:68-70)."""
    return tuple(float(v) for v in np.asarray(instance["model_row3"]).reshape(-1)[:3])


class BatchKey:
    """instancing.rs:74-90"""

    def __init__(self, mesh_id, material_id):
        self.mesh_id, self.material_id = int(mesh_id), int(material_id)

    def __eq__(self, o):
        return (self.mesh_id, self.material_id) == (o.mesh_id, o.material_id)

    def __hash__(self):
        return hash((self.mesh_id, self.material_id))


class InstancingStats:
    """instancing.rs:155-186"""

    def __init__(self):
        self.total_instances = 0
        self.batch_count = 0
        self.instances_culled = 0
        self.avg_instances_per_batch = 0.0

    def efficiency(self):
        return 0.0 if self.batch_count == 0 else self.avg_instances_per_batch / MAX_INSTANCES_PER_BATCH

    def format(self):
        return (f"Instancing: {self.total_instances} instances in {self.batch_count} batches "
                f"(avg {self.avg_instances_per_batch:.1f}), {self.instances_culled} culled")


class InstancingManager:
    """InstancingManager (instancing.rs:188-277).  Instances are kept in submission order together with
    the dense id of their BatchKey (first-seen order), which is what the device stage consumes."""

    def __init__(self):
        self._keys = {}                # BatchKey -> dense batch id
        self._counts = []              # instances per batch this frame
        self._instances, self._batch_id = [], []
        self._stats = InstancingStats()
        self.frustum_cull = True

    def begin_frame(self):
        self._counts = [0] * len(self._counts)
        self._instances, self._batch_id = [], []
        self._stats = InstancingStats()

    def add_instance(self, key, instance):
        b = self._keys.setdefault(key, len(self._keys))
        if b == len(self._counts):
            self._counts.append(0)
        if self._counts[b] < MAX_INSTANCES_PER_BATCH:          # :219-223
            self._counts[b] += 1
            self._instances.append(instance)
            self._batch_id.append(b)
            self._stats.total_instances += 1

    def add_instances(self, key, instances):
        for i in instances:
            self.add_instance(key, i)

    def finalize(self):
        self._stats.batch_count = sum(1 for c in self._counts if c)          # empty batches are dropped, :255
        if self._stats.batch_count:
            self._stats.avg_instances_per_batch = float(np.float32(self._stats.total_instances) / np.float32(self._stats.batch_count))

    def stats(self):
        return self._stats

    def set_frustum_cull(self, enabled):
        self.frustum_cull = bool(enabled)

    def batch_count_of(self, key):
        return self._counts[self._keys[key]] if key in self._keys else 0

    def batch_index(self, key):
        return self._keys.get(key)

    def instance_data(self):
        return np.array(self._instances, T.INSTANCE_DATA).reshape(-1) if self._instances else np.zeros(0, T.INSTANCE_DATA)

    def batch_ids(self):
        return np.asarray(self._batch_id, np.uint32)


class LodLevel:
    """lod_system.rs:61-76"""

    def __init__(self, screen_threshold=0.0, distance_threshold=0.0, triangle_count=0, vertex_count=0,
                 index_offset=0, index_count=0):
        self.screen_threshold, self.distance_threshold = screen_threshold, distance_threshold
        self.triangle_count, self.vertex_count = triangle_count, vertex_count
        self.index_offset, self.index_count = index_offset, index_count


class LodMesh:
    """lod_system.rs:78-124"""

    def __init__(self, levels, bounding_radius):
        self.levels, self.bounding_radius, self.lod_bias = list(levels), bounding_radius, 0.0

    @staticmethod
    def single(triangle_count, bounding_radius):
        return LodMesh([LodLevel(0.0, F32_MAX, triangle_count, triangle_count, 0, triangle_count * 3)], bounding_radius)

    def with_level(self, level):
        self.levels.append(level)
        return self

    def level_count(self):
        return len(self.levels)

    def record(self):
        """-> one AshLodMesh (levels flattened to MAX_LOD_LEVELS)."""
        if len(self.levels) > MAX_LOD_LEVELS:
            raise ValueError("more than MAX_LOD_LEVELS levels")
        r = np.zeros((), T.LOD_MESH)
        for i, l in enumerate(self.levels):
            r["levels"][i] = (l.screen_threshold, l.distance_threshold, l.triangle_count, l.vertex_count,
                              l.index_offset, l.index_count)
        r["level_count"], r["bounding_radius"], r["lod_bias"] = len(self.levels), self.bounding_radius, self.lod_bias
        return r


class LodConfig:
    """lod_system.rs:35-58; mode is "ScreenSize", "Distance" or ("Force", level) (:17-26)."""

    def __init__(self, mode="ScreenSize", smooth_transitions=True, transition_width=0.1, global_bias=0.0,
                 cull_threshold=0.001):
        self.mode, self.smooth_transitions, self.transition_width = mode, smooth_transitions, transition_width
        self.global_bias, self.cull_threshold = global_bias, cull_threshold

    def record(self, use_visibility=True):
        r = np.zeros(1, T.LOD_CONFIG)
        if isinstance(self.mode, tuple):
            r["mode"], r["force_level"] = T.LOD_MODE_FORCE, self.mode[1]
        else:
            r["mode"] = {"ScreenSize": T.LOD_MODE_SCREEN_SIZE, "Distance": T.LOD_MODE_DISTANCE}[self.mode]
        r["smooth_transitions"], r["transition_width"] = int(self.smooth_transitions), self.transition_width
        r["global_bias"], r["cull_threshold"], r["use_visibility"] = self.global_bias, self.cull_threshold, int(use_visibility)
        return r


class LodSelection:
    """lod_system.rs:126-137"""

    def __init__(self, level, blend, culled, screen_coverage):
        self.level, self.blend, self.culled, self.screen_coverage = level, blend, culled, screen_coverage


class LodStats:
    """lod_system.rs:146-183"""

    def __init__(self):
        self.objects_per_lod = [0] * MAX_LOD_LEVELS
        self.objects_culled = 0
        self.triangles_before = 0
        self.triangles_after = 0

    def reduction_rate(self):
        return 0.0 if self.triangles_before == 0 else 1.0 - self.triangles_after / self.triangles_before

    def format(self):
        return (f"LOD: {self.reduction_rate() * 100.0:.1f}% reduction ({self.triangles_before / 1e6:.1f}M → "
                f"{self.triangles_after / 1e6:.1f}M tris), {self.objects_culled} culled")

    def reset(self):
        self.__init__()


def _rust_as_usize(x):
    x = np.float32(x)
    if np.isnan(x) or x <= 0:
        return 0
    return int(min(float(x), 4294967295.0))


class LodManager:
    """LodManager (lod_system.rs:139-394): the scalar host version, one object per call, in fp32 with the
    reference's evaluation order.  The device stage (ashcuda_instancing_finalize) does the same per frame
    for every visible object; tests compare the two through the oracle."""

    def __init__(self, config=None):
        self.config = config or LodConfig()
        self._stats = LodStats()

    def begin_frame(self):
        self._stats.reset()

    @staticmethod
    def calculate_screen_coverage(position, radius, view_proj_cols, screen_width, screen_height):   # :207-235
        f = np.float32
        m = np.asarray(view_proj_cols, np.float32).reshape(4, 4)
        p = [f(v) for v in position]
        with np.errstate(all="ignore"):
            clip_w = f(f(f(m[0][3] * p[0]) + f(m[1][3] * p[1])) + f(m[2][3] * p[2])) + f(m[3][3] * f(1.0))
            if clip_w <= 0:
                return f(0.0)
            proj_radius = f(radius) / clip_w
            sw, sh = f(screen_width), f(screen_height)
            mx = sh if (np.isnan(sw) or sw < sh) and not np.isnan(sh) else sw
            pixel_radius = f(f(proj_radius * mx) * f(0.5))
            area = f(f(f(np.pi) * pixel_radius) * pixel_radius)
            cov = f(area / f(sw * sh))
            return f(1.0) if np.isnan(cov) or cov > 1 else cov

    def select_lod(self, mesh, world_position, view_proj_cols, camera_position, screen_width, screen_height):   # :238-345
        f = np.float32
        st, cfg, n = self._stats, self.config, len(mesh.levels)
        last = max(n - 1, 0)
        if n:
            st.triangles_before += mesh.levels[0].triangle_count
        if isinstance(cfg.mode, tuple):
            level = min(cfg.mode[1], last)
            if n:
                st.triangles_after += mesh.levels[level].triangle_count
                st.objects_per_lod[level] += 1
            return LodSelection(level, 0.0, False, 1.0)
        cov = self.calculate_screen_coverage(world_position, mesh.bounding_radius, view_proj_cols, screen_width, screen_height)
        if cov < f(cfg.cull_threshold):
            st.objects_culled += 1
            return LodSelection(last, 0.0, True, float(cov))
        level, blend = last, f(0.0)
        with np.errstate(all="ignore"):
            if cfg.mode == "ScreenSize":
                for i, l in enumerate(mesh.levels):
                    th = f(l.screen_threshold)
                    if cov >= th:
                        if cfg.smooth_transitions and i + 1 < n:
                            rng = f(th - f(mesh.levels[i + 1].screen_threshold))
                            if rng > 0:
                                b = f(f(f(th - cov) / rng) / f(cfg.transition_width))
                                blend = b if np.isnan(b) else min(max(b, f(0.0)), f(1.0))
                        level = i
                        break
            else:
                d = [f(f(a) - f(b)) for a, b in zip(world_position, camera_position)]
                dist = np.sqrt(f(f(f(d[0] * d[0]) + f(d[1] * d[1])) + f(d[2] * d[2])))
                for i, l in enumerate(mesh.levels):
                    if dist <= f(l.distance_threshold):
                        level = i
                        break
            x = f(f(f(level) + f(cfg.global_bias)) + f(mesh.lod_bias))
            if x < 0:
                x = f(0.0)
            if x > f(last):
                x = f(last)
        biased = _rust_as_usize(x)
        if biased < n:
            st.triangles_after += mesh.levels[biased].triangle_count
            if biased < MAX_LOD_LEVELS:
                st.objects_per_lod[biased] += 1
        return LodSelection(biased, float(blend), False, float(cov))

    def stats(self):
        return self._stats

    def config_mut(self):
        return self.config


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
            if self.tonemapping.enabled:             # TonemappingFeature::render returns early when disabled (:127-129)
                st |= T.STAGE_TONEMAP
        vp["stages"] = st
        return vp

    def before_frame(self, ctx=None):
        if self.ctx is None:
            from .backend import AshError, ASH_ERR_NOT_INITIALIZED
            raise AshError(ASH_ERR_NOT_INITIALIZED, "CudaComputeFeature used before on_added")
        self.taa.begin_frame()
        self.ctx.set_tonemap_operator(self.tonemapping.operator_id())
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
