"""POD layouts of the reference's #[repr(C)] structs as numpy dtypes (host-side mirror).

Sizes are the reference's: CullObjectData 112 B (occlusion_culling.rs:72-91, test :311-314),
IndirectDrawCommand 20 B (:114-120), CullingPushConstants 96 B (:125-136), GpuLight 64 B
(features/light_culling.rs:28-37), LightCullingPushConstants 16 B (:69-76), CullingCameraData
208 B (:81-86), TaaPushConstants 48 B (temporal_aa.rs:86-97), PostProcessPushConstants 16 B
(fullscreen_pass.rs:152-163); row f-2: InstanceData 96 B (instancing.rs:19-34), LodLevel 24 B
(lod_system.rs:61-76).  include/ashcuda.h declares the same layouts for C.
"""
import numpy as np

CULL_OBJECT = np.dtype([("center", "<f4", 4), ("extents", "<f4", 4), ("model", "<f4", (4, 4)),
                        ("draw_index", "<u4"), ("lod_bias", "<f4"), ("flags", "<u4"),
                        ("_padding", "<u4")])
DRAW_COMMAND = np.dtype([("index_count", "<u4"), ("instance_count", "<u4"), ("first_index", "<u4"),
                         ("vertex_offset", "<i4"), ("first_instance", "<u4")])
GPU_LIGHT = np.dtype([("position", "<f4", 4), ("color", "<f4", 4), ("direction", "<f4", 4),
                      ("params", "<f4", 4)])
CULLING_PC = np.dtype([("view_proj", "<f4", (4, 4)), ("screen_params", "<f4", 4),
                       ("object_count", "<u4"), ("hiz_levels", "<u4"), ("_padding", "<u4", 2)])
LIGHT_PC = np.dtype([("screen_size", "<u4", 2), ("light_count", "<u4"), ("_padding", "<u4")])
CAMERA_DATA = np.dtype([("view", "<f4", (4, 4)), ("projection", "<f4", (4, 4)),
                        ("inv_projection", "<f4", (4, 4)), ("camera_pos", "<f4", 4)])
TAA_PC = np.dtype([("screen_params", "<f4", 4), ("params", "<f4", 4), ("jitter", "<f4", 2),
                   ("prev_jitter", "<f4", 2)])
POST_PC = np.dtype([("exposure", "<f4"), ("gamma", "<f4"), ("bloom_intensity", "<f4"),
                    ("_padding", "<f4")])
BLOOM_PARAMS = np.dtype([("threshold", "<f4"), ("upsample_intensity", "<f4"), ("mip_count", "<u4"),
                         ("_padding", "<u4")])
CONFIG = np.dtype([("max_objects", "<u4"), ("max_lights", "<u4"), ("max_draws", "<u4"),
                   ("hiz_levels", "<u4"), ("bloom_mips", "<u4"), ("enable_post", "<u4"),
                   ("srgb_encode", "<u4"), ("concurrent_frontend", "<u4"), ("max_batches", "<u4"),
                   ("graph_replay", "<u4"), ("precision", "<u4"), ("fp32_intermediates", "<u4"), ("_reserved", "<u4", 4)])
PRECISION_EXACT, PRECISION_FAST = 0, 1
VIEW_STATS = np.dtype([("view_id", "<u4"), ("total_objects", "<u4"), ("visible_objects", "<u4"),
                       ("tiles", "<u4"), ("sum_tile_lights", "<u4"), ("max_tile_lights", "<u4"),
                       ("overflow_tiles", "<u4"), ("candidate_lights", "<u4")])
VIEW_PARAMS = np.dtype([("cull", CULLING_PC), ("light", LIGHT_PC), ("camera", CAMERA_DATA),
                        ("taa", TAA_PC), ("bloom", BLOOM_PARAMS), ("post", POST_PC),
                        ("view_id", "<u4"), ("stages", "<u4"), ("_padding", "<u4", 2)])

# ---- row f-2: LOD selection + instance batching
MAX_LOD_LEVELS, MAX_INSTANCES_PER_BATCH, LOD_NOT_VISIBLE = 8, 65536, 0xFFFFFFFF
LOD_MODE_SCREEN_SIZE, LOD_MODE_DISTANCE, LOD_MODE_FORCE = 0, 1, 2
INSTANCE_DATA = np.dtype([("model_row0", "<f4", 4), ("model_row1", "<f4", 4), ("model_row2", "<f4", 4),
                          ("model_row3", "<f4", 4), ("color", "<f4", 4), ("custom", "<f4", 4)])
LOD_LEVEL = np.dtype([("screen_threshold", "<f4"), ("distance_threshold", "<f4"), ("triangle_count", "<u4"),
                      ("vertex_count", "<u4"), ("index_offset", "<u4"), ("index_count", "<u4")])
LOD_MESH = np.dtype([("levels", LOD_LEVEL, MAX_LOD_LEVELS), ("level_count", "<u4"), ("bounding_radius", "<f4"),
                     ("lod_bias", "<f4"), ("_padding", "<u4")])
LOD_CONFIG = np.dtype([("mode", "<u4"), ("force_level", "<u4"), ("smooth_transitions", "<u4"),
                       ("transition_width", "<f4"), ("global_bias", "<f4"), ("cull_threshold", "<f4"),
                       ("use_visibility", "<u4"), ("_padding", "<u4")])
LOD_SELECTION = np.dtype([("level", "<u4"), ("blend", "<f4"), ("culled", "<u4"), ("screen_coverage", "<f4")])
INSTANCING_STATS = np.dtype([("objects_per_lod", "<u4", MAX_LOD_LEVELS), ("objects_culled", "<u4"),
                             ("total_instances", "<u4"), ("batch_count", "<u4"), ("avg_instances_per_batch", "<f4"),
                             ("triangles_before", "<u8"), ("triangles_after", "<u8"), ("instances_dropped", "<u4"),
                             ("_padding", "<u4")])
INSTANCING_PARAMS = np.dtype([("view_proj", "<f4", (4, 4)), ("camera_pos", "<f4", 4), ("screen_size", "<f4", 2),
                              ("object_count", "<u4"), ("_padding", "<u4")])
assert INSTANCE_DATA.itemsize == 96 and LOD_LEVEL.itemsize == 24 and LOD_MESH.itemsize == 208
assert LOD_CONFIG.itemsize == 32 and LOD_SELECTION.itemsize == 16 and INSTANCING_STATS.itemsize == 72
assert INSTANCING_PARAMS.itemsize == 96
# (f-1) compute forward pass: the frag.frag uniform fields it uses (AshForwardParams)
FORWARD_PARAMS = np.dtype([("light_direction", "<f4", 4), ("light_color", "<f4", 4), ("ambient_color", "<f4", 4),
                           ("base_color_factor", "<f4", 4), ("emissive_factor", "<f4", 4), ("parameters", "<f4", 4),
                           ("background", "<f4", 4), ("near_plane", "<f4"), ("radius_scale", "<f4"), ("flags", "<u4"),
                           ("_padding", "<u4")])
assert FORWARD_PARAMS.itemsize == 128
FWD_TILE_LIGHTS, FWD_REINHARD, FWD_INSTANCE_COLOR = 1, 2, 4
FWD_NO_OBJECT = 0xFFFFFFFF
# CullingStats (render_stats.rs:54-65) + the early / late split of two-phase culling (AshCullingStats)
CULLING_STATS = np.dtype([("objects_submitted", "<u4"), ("after_frustum", "<u4"), ("after_occlusion", "<u4"),
                          ("lights_culled", "<u4"), ("lod_triangles_saved", "<u8"), ("early_visible", "<u4"),
                          ("late_visible", "<u4")])
assert CULLING_STATS.itemsize == 32

assert CULL_OBJECT.itemsize == 112 and DRAW_COMMAND.itemsize == 20 and CULLING_PC.itemsize == 96
assert GPU_LIGHT.itemsize == 64 and LIGHT_PC.itemsize == 16 and CAMERA_DATA.itemsize == 208
assert TAA_PC.itemsize == 48 and POST_PC.itemsize == 16 and BLOOM_PARAMS.itemsize == 16
assert CONFIG.itemsize == 64 and VIEW_STATS.itemsize == 32
assert VIEW_PARAMS.itemsize == 96 + 16 + 208 + 48 + 16 + 16 + 16

TONEMAP_ACES, TONEMAP_REINHARD, TONEMAP_UNCHARTED2, TONEMAP_NONE = 0, 1, 2, 3   # TonemapOperator, tonemapping.rs:9-20

TILE_SIZE = 16
MAX_LIGHTS_PER_TILE = 256
TILE_STRIDE = 257

STAGE_HIZ, STAGE_CULL, STAGE_LIGHTS, STAGE_TAA, STAGE_BLOOM, STAGE_TONEMAP = 1, 2, 4, 8, 16, 32
STAGE_ALL = 0x3F
STAGE_KEEP_BLOOM = 0x40
STAGE_INSTANCING = 0x80
STAGE_FORWARD_DEPTH, STAGE_FORWARD_SHADE = 0x100, 0x200
STAGE_TWO_PHASE = 0x400

BUF_VISIBILITY, BUF_DRAWS_OUT, BUF_VISIBLE_COUNT, BUF_TILE_DATA = 1, 2, 3, 4
BUF_TAA_OUT, BUF_LDR8, BUF_STATS, BUF_TILE_MINMAX, BUF_VISIBILITY_BITS = 5, 6, 7, 8, 9
BUF_BLOOM_THRESHOLD = 10
BUF_LOD_SELECTION, BUF_INSTANCES_OUT, BUF_BUCKET_OFFSETS, BUF_INSTANCE_DRAWS, BUF_INSTANCING_STATS = 11, 12, 13, 14, 15
BUF_FORWARD_KEYS, BUF_COLOR, BUF_CULLING_STATS, BUF_VELOCITY = 16, 17, 18, 19


def BUF_HIZ_MIP(i):
    return 0x100 + i


def BUF_BLOOM_MIP(i):
    return 0x200 + i

# kernel classes (ashcuda_profile_kernels / ashcuda_kernel_time_ms)
(KERNEL_DEPTH_PYRAMID, KERNEL_OCCLUSION_CULL, KERNEL_CULL_COMPACT, KERNEL_LIGHT_PREPASS, KERNEL_LIGHT_TILES,
 KERNEL_TAA, KERNEL_BLOOM_THRESHOLD, KERNEL_BLOOM_DOWNSAMPLE, KERNEL_BLOOM_UPSAMPLE, KERNEL_TONEMAP,
 KERNEL_TAA_THRESHOLD, KERNEL_UPSAMPLE_TONEMAP, KERNEL_INSTANCING, KERNEL_FORWARD_DEPTH, KERNEL_FORWARD_SHADE,
 KERNEL_BLOOM_TAIL) = range(16)
KERNEL_COUNT = 16
