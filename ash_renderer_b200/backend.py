"""ctypes binding of libashcuda.so (include/ashcuda.h) — the only compute path of this package.

There is NO CPU fallback: importing works without a GPU (so the symbol table can be checked), but
`Context(...)` raises `AshError` when the library is missing or no CUDA device is usable.  The
method names follow the reference's per-stage wrapper `LightCullingPipeline`
(vulkan/light_culling_pipeline.rs:41-351: new / upload_lights / upload_camera / dispatch /
tile_buffer / Drop) generalised to every stage of the path.
"""
import ctypes as C
import os

import numpy as np

from . import types as T

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libashcuda.so")

ASH_OK, ASH_ERR_DEVICE, ASH_ERR_NOT_INITIALIZED, ASH_ERR_NOT_FOUND, ASH_ERR_INVALID, ASH_ERR_OOM = 0, -1, -2, -3, -4, -5
_ERR_NAMES = {ASH_ERR_DEVICE: "DeviceInitFailed", ASH_ERR_NOT_INITIALIZED: "FeatureNotInitialized",
              ASH_ERR_NOT_FOUND: "ResourceNotFound", ASH_ERR_INVALID: "InvalidArgument",
              ASH_ERR_OOM: "OutOfMemory"}

# every symbol include/ashcuda.h declares (tests/test_abi.py checks header <-> this list <-> .so)
EXPORTS = [
    "ashcuda_default_config", "ashcuda_create", "ashcuda_destroy", "ashcuda_set_stream",
    "ashcuda_set_depth", "ashcuda_set_objects", "ashcuda_set_lights", "ashcuda_set_camera",
    "ashcuda_set_color", "ashcuda_taa_reset_history", "ashcuda_hiz_build", "ashcuda_occlusion_cull",
    "ashcuda_light_cull", "ashcuda_taa_resolve", "ashcuda_bloom", "ashcuda_tonemap", "ashcuda_set_tonemap_operator", "ashcuda_cull_compact_static",
    "ashcuda_view_run", "ashcuda_get", "ashcuda_readback", "ashcuda_readback_async", "ashcuda_sync",
    "ashcuda_hiz_level_count", "ashcuda_mip_extent", "ashcuda_kernel_launches", "ashcuda_graph_stats", "ashcuda_last_error",
    "ashcuda_version", "ashcuda_profile_kernels", "ashcuda_kernel_time_ms", "ashcuda_upload_view", "ashcuda_download_view", "ashcuda_download_view_packed",
    "ashcuda_debug_exact_taps",
    "ashcuda_set_instances", "ashcuda_set_lod_meshes", "ashcuda_set_lod_config", "ashcuda_instancing_finalize",
    "ashcuda_set_forward_params", "ashcuda_set_forward_reprojection", "ashcuda_forward_depth", "ashcuda_forward_shade", "ashcuda_debug_compute_utils",
]


class AshError(RuntimeError):
    """Mirror of `enum AshError` (src/error.rs:12-29): code + message from ashcuda_last_error."""

    def __init__(self, code, msg):
        super().__init__(f"{_ERR_NAMES.get(code, 'AshError')}({code}): {msg}")
        self.code = code


_lib = None


def load_library():
    """dlopen libashcuda.so; raises AshError(DeviceInitFailed) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AshError(ASH_ERR_DEVICE, f"{LIB_PATH} not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                                       "there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp, u32, i32 = C.c_void_p, C.c_uint32, C.c_int
    sig = {
        "ashcuda_default_config": (None, [vp]),
        "ashcuda_create": (i32, [C.POINTER(vp), i32, u32, u32, vp]),
        "ashcuda_destroy": (None, [vp]),
        "ashcuda_set_stream": (i32, [vp, vp]),
        "ashcuda_set_depth": (i32, [vp, vp, i32]),
        "ashcuda_set_objects": (i32, [vp, vp, u32, vp, u32, i32]),
        "ashcuda_set_lights": (i32, [vp, vp, u32, i32]),
        "ashcuda_set_camera": (i32, [vp, vp]),
        "ashcuda_set_color": (i32, [vp, vp, vp, i32]),
        "ashcuda_taa_reset_history": (i32, [vp]),
        "ashcuda_hiz_build": (i32, [vp]),
        "ashcuda_occlusion_cull": (i32, [vp, vp]),
        "ashcuda_light_cull": (i32, [vp, vp]),
        "ashcuda_taa_resolve": (i32, [vp, vp]),
        "ashcuda_bloom": (i32, [vp, vp]),
        "ashcuda_tonemap": (i32, [vp, vp]),
        "ashcuda_set_tonemap_operator": (i32, [vp, u32]),
        "ashcuda_cull_compact_static": (i32, [i32, vp, vp, vp, u32, vp, u32, vp, vp, C.c_char_p, C.c_size_t]),
        "ashcuda_view_run": (i32, [vp, vp]),
        "ashcuda_get": (i32, [vp, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "ashcuda_readback": (i32, [vp, i32, vp, C.c_size_t]),
        "ashcuda_readback_async": (i32, [vp, i32, vp, C.c_size_t]),
        "ashcuda_sync": (i32, [vp]),
        "ashcuda_hiz_level_count": (i32, [vp]),
        "ashcuda_mip_extent": (i32, [vp, i32, C.POINTER(u32), C.POINTER(u32)]),
        "ashcuda_kernel_launches": (C.c_uint64, [vp]),
        "ashcuda_graph_stats": (C.c_int, [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
        "ashcuda_last_error": (C.c_char_p, [vp]),
        "ashcuda_version": (C.c_char_p, []),
        "ashcuda_profile_kernels": (i32, [vp, u32]),
        "ashcuda_kernel_time_ms": (i32, [vp, i32, C.POINTER(C.c_float), C.POINTER(C.c_uint64)]),
        "ashcuda_debug_exact_taps": (i32, [i32, i32, i32, C.c_float, C.c_float]),
        "ashcuda_upload_view": (i32, [vp, vp, vp, vp]),
        "ashcuda_download_view": (i32, [vp, vp, vp, vp, vp, vp]),
        "ashcuda_download_view_packed": (i32, [vp, vp, vp, vp, u32, vp, vp, vp]),
        "ashcuda_set_instances": (i32, [vp, vp, vp, vp, u32, i32]),
        "ashcuda_set_lod_meshes": (i32, [vp, vp, u32, i32]),
        "ashcuda_set_lod_config": (i32, [vp, vp]),
        "ashcuda_instancing_finalize": (i32, [vp, vp]),
        "ashcuda_set_forward_params": (i32, [vp, vp]),
        "ashcuda_set_forward_reprojection": (i32, [vp, vp]),
        "ashcuda_forward_depth": (i32, [vp, vp, u32]),
        "ashcuda_forward_shade": (i32, [vp]),
        "ashcuda_debug_compute_utils": (i32, [i32, i32, vp, u32, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def _ptr(a):
    """numpy array -> host pointer; torch CUDA tensor / int -> device pointer. Returns (ptr, on_device)."""
    if a is None:
        return None, 0
    if isinstance(a, int):
        return C.c_void_p(a), 1
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise AshError(ASH_ERR_INVALID, "host arrays must be C-contiguous")
        return C.c_void_p(a.ctypes.data), 0
    if hasattr(a, "data_ptr"):          # torch tensor
        if not a.is_contiguous():
            raise AshError(ASH_ERR_INVALID, "tensors must be contiguous")
        return C.c_void_p(a.data_ptr()), (1 if a.is_cuda else 0)
    raise AshError(ASH_ERR_INVALID, f"unsupported buffer type {type(a)}")


def default_config(**over):
    cfg = np.zeros(1, T.CONFIG)
    load_library().ashcuda_default_config(cfg.ctypes.data)
    for k, v in over.items():
        cfg[k] = v
    return cfg


class Context:
    """One CUDA device + one stream + every buffer of the path (ashcuda_ctx)."""

    def __init__(self, width, height, device=0, **config):
        self._lib = load_library()
        self._h = C.c_void_p()
        self.W, self.H = int(width), int(height)
        self.cfg = default_config(**config)
        rc = self._lib.ashcuda_create(C.byref(self._h), int(device), self.W, self.H, self.cfg.ctypes.data)
        if rc != ASH_OK:
            msg = self._lib.ashcuda_last_error(None)
            self._h = C.c_void_p()
            raise AshError(rc, (msg or b"").decode())
        self._keep = {}

    # -- lifetime (Drop, light_culling_pipeline.rs:332-351)
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.ashcuda_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != ASH_OK:
            raise AshError(rc, (self._lib.ashcuda_last_error(self._h) or b"").decode())

    def set_stream(self, cuda_stream_ptr):
        self._ck(self._lib.ashcuda_set_stream(self._h, C.c_void_p(cuda_stream_ptr) if cuda_stream_ptr else None))

    # -- inputs
    def set_depth(self, depth):
        p, dev = _ptr(depth)
        self._keep["depth"] = depth
        self._ck(self._lib.ashcuda_set_depth(self._h, p, dev))

    def set_objects(self, objs, draws_in):
        po, dev = _ptr(objs)
        pd, dev2 = _ptr(draws_in)
        n = objs.shape[0] if isinstance(objs, np.ndarray) else objs.numel() * objs.element_size() // T.CULL_OBJECT.itemsize
        nd = draws_in.shape[0] if isinstance(draws_in, np.ndarray) else draws_in.numel() * draws_in.element_size() // T.DRAW_COMMAND.itemsize
        self._keep["objs"] = (objs, draws_in)
        self._ck(self._lib.ashcuda_set_objects(self._h, po, n, pd, nd, dev))

    def upload_lights(self, lights):          # LightCullingPipeline::upload_lights :248-263
        p, dev = _ptr(lights)
        n = lights.shape[0] if isinstance(lights, np.ndarray) else lights.numel() * lights.element_size() // T.GPU_LIGHT.itemsize
        self._keep["lights"] = lights
        self._ck(self._lib.ashcuda_set_lights(self._h, p, n, dev))

    def upload_camera(self, camera_data):     # LightCullingPipeline::upload_camera :268-281
        cd = np.ascontiguousarray(camera_data, T.CAMERA_DATA)
        self._ck(self._lib.ashcuda_set_camera(self._h, cd.ctypes.data))

    def set_color(self, hdr, velocity=None):
        p, dev = _ptr(hdr)
        pv, _ = _ptr(velocity)
        self._keep["color"] = (hdr, velocity)
        self._ck(self._lib.ashcuda_set_color(self._h, p, pv, dev))

    def taa_reset_history(self):
        self._ck(self._lib.ashcuda_taa_reset_history(self._h))

    # -- stages
    def hiz_build(self):
        self._ck(self._lib.ashcuda_hiz_build(self._h))

    def occlusion_cull(self, pc):
        pc = np.ascontiguousarray(pc, T.CULLING_PC)
        self._ck(self._lib.ashcuda_occlusion_cull(self._h, pc.ctypes.data))

    def light_cull(self, pc):                 # LightCullingPipeline::dispatch :286-319
        pc = np.ascontiguousarray(pc, T.LIGHT_PC)
        self._ck(self._lib.ashcuda_light_cull(self._h, pc.ctypes.data))

    def taa_resolve(self, pc):
        pc = np.ascontiguousarray(pc, T.TAA_PC)
        self._ck(self._lib.ashcuda_taa_resolve(self._h, pc.ctypes.data))

    def bloom(self, params):
        bp = np.ascontiguousarray(params, T.BLOOM_PARAMS)
        self._ck(self._lib.ashcuda_bloom(self._h, bp.ctypes.data))

    def set_tonemap_operator(self, op):          # TonemappingFeature::set_operator, tonemapping.rs:92-94
        self._ck(self._lib.ashcuda_set_tonemap_operator(self._h, int(op)))

    def tonemap(self, pc):
        pc = np.ascontiguousarray(pc, T.POST_PC)
        self._ck(self._lib.ashcuda_tonemap(self._h, pc.ctypes.data))

    def view_run(self, view_params):
        vp = np.ascontiguousarray(view_params, T.VIEW_PARAMS)
        self._ck(self._lib.ashcuda_view_run(self._h, vp.ctypes.data))

    def upload_view(self, depth_host, hdr_host, velocity_host=None):
        """Host -> device copies of one view's per-frame inputs (pinned or pageable host memory)."""
        pd, _ = _ptr(depth_host)
        ph, _ = _ptr(hdr_host)
        pv, _ = _ptr(velocity_host)
        self._ck(self._lib.ashcuda_upload_view(self._h, pd, ph, pv))

    def download_view(self, visibility=None, tile_data=None, ldr8=None, stats=None, visible_count=None):
        """Device -> host copies of one view's results (async on the ctx stream; call sync())."""
        ptrs = [_ptr(a)[0] for a in (visibility, tile_data, ldr8, stats, visible_count)]
        self._ck(self._lib.ashcuda_download_view(self._h, *ptrs))

    def download_view_packed(self, visibility=None, tile_offsets=None, tile_lights=None, ldr8=None, stats=None,
                             visible_count=None):
        """download_view with the tile light lists in CSR form: tile_offsets (tiles + 1 words) and tile_lights
        (its length is the capacity).  After sync(), tile_offsets[-1] > len(tile_lights) means truncated."""
        cap = 0 if tile_lights is None else int(tile_lights.numel() if hasattr(tile_lights, "numel") else tile_lights.size)
        ptrs = [_ptr(a)[0] for a in (visibility, tile_offsets, tile_lights)]
        tail = [_ptr(a)[0] for a in (ldr8, stats, visible_count)]
        self._ck(self._lib.ashcuda_download_view_packed(self._h, *ptrs, cap, *tail))

    def sync(self):
        self._ck(self._lib.ashcuda_sync(self._h))

    # -- outputs
    def get(self, which):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._lib.ashcuda_get(self._h, int(which), C.byref(p), C.byref(n)))
        return p.value, n.value

    def readback(self, which, dtype=np.uint8, count=None):
        _, nbytes = self.get(which)
        dt = np.dtype(dtype)
        if count is not None:
            nbytes = min(nbytes, count * dt.itemsize)
        out = np.empty(nbytes // dt.itemsize, dt)
        if nbytes:
            self._ck(self._lib.ashcuda_readback(self._h, int(which), out.ctypes.data, nbytes))
        else:
            self.sync()
        return out

    def hiz_level_count(self):
        return int(self._lib.ashcuda_hiz_level_count(self._h))

    def mip_extent(self, which):
        w, h = C.c_uint32(), C.c_uint32()
        self._ck(self._lib.ashcuda_mip_extent(self._h, int(which), C.byref(w), C.byref(h)))
        return w.value, h.value

    def hiz_mip(self, level):
        w, h = self.mip_extent(T.BUF_HIZ_MIP(level))
        return self.readback(T.BUF_HIZ_MIP(level), np.float32).reshape(h, w)

    def bloom_mip(self, k):
        w, h = self.mip_extent(T.BUF_BLOOM_MIP(k))
        return self.readback(T.BUF_BLOOM_MIP(k), self._image_dtype()).reshape(h, w, 4)

    def visibility(self):
        return self.readback(T.BUF_VISIBILITY, np.uint32)

    def visible_count(self):
        return int(self.readback(T.BUF_VISIBLE_COUNT, np.uint32)[0])

    def draws_out(self, count):
        return self.readback(T.BUF_DRAWS_OUT, T.DRAW_COMMAND, count)

    def set_forward_reprojection(self, reproj_cols=None):
        """reproj = prevViewProj * inverse(view) as a (4,4) array of COLUMNS (None: no velocity output)."""
        if reproj_cols is None:
            self._ck(self._lib.ashcuda_set_forward_reprojection(self._h, None))
        else:
            m = np.ascontiguousarray(reproj_cols, np.float32).reshape(16)
            self._ck(self._lib.ashcuda_set_forward_reprojection(self._h, m.ctypes.data))

    def velocity(self):
        return self.readback(T.BUF_VELOCITY, np.float16).reshape(self.H, self.W, 2)

    def tile_data(self):
        return self.readback(T.BUF_TILE_DATA, np.uint32).reshape(-1, T.TILE_STRIDE)

    def _image_dtype(self):
        return np.float32 if int(self.cfg["fp32_intermediates"][0]) else np.float16

    def taa_out(self):
        return self.readback(T.BUF_TAA_OUT, self._image_dtype()).reshape(self.H, self.W, 4)

    def ldr8(self):
        return self.readback(T.BUF_LDR8, np.uint8).reshape(self.H, self.W, 4)

    def stats(self):
        return self.readback(T.BUF_STATS, T.VIEW_STATS)[0]

    # ---- row f-2: LOD selection + instance batching (lod_system.rs / instancing.rs on the device)
    def set_instances(self, instances, batch_id, lod_mesh_id):
        pi, d0 = _ptr(instances)
        pb, d1 = _ptr(batch_id)
        pm, d2 = _ptr(lod_mesh_id)
        if not (d0 == d1 == d2):
            raise AshError(ASH_ERR_INVALID, "set_instances: all three arrays must live on the same side")
        n = batch_id.shape[0] if isinstance(batch_id, np.ndarray) else batch_id.numel()
        self._keep["instances"] = (instances, batch_id, lod_mesh_id)
        self._ck(self._lib.ashcuda_set_instances(self._h, pi, pb, pm, n, d0))

    def set_lod_meshes(self, meshes):
        p, dev = _ptr(meshes)
        n = meshes.shape[0] if isinstance(meshes, np.ndarray) else meshes.numel() * meshes.element_size() // T.LOD_MESH.itemsize
        self._keep["lod_meshes"] = meshes
        self._ck(self._lib.ashcuda_set_lod_meshes(self._h, p, n, dev))

    def set_lod_config(self, cfg=None):
        self._ck(self._lib.ashcuda_set_lod_config(self._h, None if cfg is None else np.ascontiguousarray(cfg, T.LOD_CONFIG).ctypes.data))

    def instancing_finalize(self, params):
        ip = np.ascontiguousarray(params, T.INSTANCING_PARAMS)
        self._ck(self._lib.ashcuda_instancing_finalize(self._h, ip.ctypes.data))

    def lod_selection(self):
        return self.readback(T.BUF_LOD_SELECTION, T.LOD_SELECTION)

    def instancing_stats(self):
        return self.readback(T.BUF_INSTANCING_STATS, T.INSTANCING_STATS)[0]

    def bucket_offsets(self):
        return self.readback(T.BUF_BUCKET_OFFSETS, np.uint32)

    def instance_draws(self):
        return self.readback(T.BUF_INSTANCE_DRAWS, T.DRAW_COMMAND)

    def instances_out(self, count):
        return self.readback(T.BUF_INSTANCES_OUT, T.INSTANCE_DATA, count)

    # ---- row f-1: depth / HDR producers (compute forward pass)
    def set_forward_params(self, params=None):
        self._ck(self._lib.ashcuda_set_forward_params(
            self._h, None if params is None else np.ascontiguousarray(params, T.FORWARD_PARAMS).ctypes.data))

    def forward_depth(self, camera, object_count):
        cam = np.ascontiguousarray(camera, T.CAMERA_DATA)
        self._ck(self._lib.ashcuda_forward_depth(self._h, cam.ctypes.data, int(object_count)))

    def forward_shade(self):
        self._ck(self._lib.ashcuda_forward_shade(self._h))

    def culling_stats(self):
        """CullingStats of the last cull (render_stats.rs:54-65) + early / late counts of two-phase culling."""
        return self.readback(T.BUF_CULLING_STATS, T.CULLING_STATS)[0]

    def forward_keys(self):
        return self.readback(T.BUF_FORWARD_KEYS, np.uint64).reshape(self.H, self.W)

    def depth(self):
        return self.readback(T.BUF_HIZ_MIP(0), np.float32).reshape(self.H, self.W)

    def color(self):
        return self.readback(T.BUF_COLOR, np.float16).reshape(self.H, self.W, 4)

    def kernel_launches(self):
        return int(self._lib.ashcuda_kernel_launches(self._h))

    def graph_stats(self):
        """(views captured into a CUDA graph, views replayed from one) by view_run."""
        cap, rep = C.c_uint64(0), C.c_uint64(0)
        self._ck(self._lib.ashcuda_graph_stats(self._h, C.byref(cap), C.byref(rep)))
        return int(cap.value), int(rep.value)

    def profile_kernels(self, mask):
        self._ck(self._lib.ashcuda_profile_kernels(self._h, int(mask)))

    def kernel_time_ms(self, kernel_id):
        """(accumulated ms, launches) of one kernel class since the last call; needs profiling on."""
        ms, n = C.c_float(), C.c_uint64()
        self._ck(self._lib.ashcuda_kernel_time_ms(self._h, int(kernel_id), C.byref(ms), C.byref(n)))
        return ms.value, n.value


STATIC_COMPACT_MAX_OBJECTS = 1 << 20          # ASH_STATIC_COMPACT_MAX_OBJECTS


def cull_compact_static(objs, draws_in, visibility, object_count, draws_out, visible_count, device=0, stream=None):
    """ashcuda_cull_compact_static (no counterpart in the reference; operators' request): the ordered draw-command
    compaction without any allocation — scratch in a static 16 KB device array shared by every call on the GPU (calls
    that overlap on a GPU race).  All arrays are device tensors the CALLER owns; enqueues on `stream` and returns."""
    po, _ = _ptr(objs)
    pd, _ = _ptr(draws_in)
    pv, _ = _ptr(visibility)
    pout, _ = _ptr(draws_out)
    pcnt, _ = _ptr(visible_count)
    n_draws = draws_in.numel() * draws_in.element_size() // T.DRAW_COMMAND.itemsize
    err = C.create_string_buffer(256)
    rc = load_library().ashcuda_cull_compact_static(int(device), C.c_void_p(stream) if stream else None, po, pd, n_draws, pv,
                                                    int(object_count), pout, pcnt, err, len(err))
    if rc != ASH_OK:
        raise AshError(rc, err.value.decode())


def version():
    return load_library().ashcuda_version().decode()


CU_MIN, CU_MAX, CU_SUM, CU_SCAN, CU_SPHERE_IN_FRUSTUM, CU_AABB_IN_FRUSTUM, CU_DEPTH, CU_PLANES = range(8)


def compute_utils(op, data=None, params=None, device=0):
    """shaders/compute_utils.glsl on the device (csrc/ash_compute_utils.cuh), one function per call; see ashcuda.h."""
    lib = load_library()
    p = np.zeros(18, np.float32)
    if params is not None:
        q = np.asarray(params, np.float32).reshape(-1)
        p[:len(q)] = q
    if op == CU_PLANES:
        out = np.zeros((6, 4), np.float32)
        rc = lib.ashcuda_debug_compute_utils(device, op, None, 0, p.ctypes.data, out.ctypes.data)
    else:
        a = np.ascontiguousarray(data, np.uint32 if op == CU_SCAN else np.float32)
        n = a.shape[0]
        out = (np.zeros(n // 256, np.float32) if op <= CU_SUM else np.zeros(n, np.uint32) if op in (CU_SCAN, CU_SPHERE_IN_FRUSTUM, CU_AABB_IN_FRUSTUM)
               else np.zeros((n, 4), np.float32))
        rc = lib.ashcuda_debug_compute_utils(device, op, a.ctypes.data, n, p.ctypes.data, out.ctypes.data)
    if rc != ASH_OK:
        raise AshError(rc, "ashcuda_debug_compute_utils")
    return out

