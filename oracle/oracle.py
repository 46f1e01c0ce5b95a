"""ctypes/numpy front-end of the CPU oracle (oracle/ash_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (ash_renderer_b200/) never imports it.
Checked bit for bit against the reference's own GLSL compiled as C++ (oracle/ref.py, tests/test_oracle_ref.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libashoracle.so")

TILE_STRIDE = 257

OBJECT_DTYPE = np.dtype([("center", "<f4", 4), ("extents", "<f4", 4), ("model", "<f4", (4, 4)),
                         ("draw_index", "<u4"), ("lod_bias", "<f4"), ("flags", "<u4"), ("_padding", "<u4")])
DRAW_DTYPE = np.dtype([("index_count", "<u4"), ("instance_count", "<u4"), ("first_index", "<u4"),
                       ("vertex_offset", "<i4"), ("first_instance", "<u4")])
LIGHT_DTYPE = np.dtype([("position", "<f4", 4), ("color", "<f4", 4), ("direction", "<f4", 4),
                        ("params", "<f4", 4)])
CULL_PC_DTYPE = np.dtype([("view_proj", "<f4", (4, 4)), ("screen_params", "<f4", 4),
                          ("object_count", "<u4"), ("hiz_levels", "<u4"), ("_padding", "<u4", 2)])
LIGHT_PC_DTYPE = np.dtype([("screen_size", "<u4", 2), ("light_count", "<u4"), ("_padding", "<u4")])
CAMERA_DTYPE = np.dtype([("view", "<f4", (4, 4)), ("projection", "<f4", (4, 4)),
                         ("inv_projection", "<f4", (4, 4)), ("camera_pos", "<f4", 4)])
TAA_PC_DTYPE = np.dtype([("screen_params", "<f4", 4), ("params", "<f4", 4), ("jitter", "<f4", 2),
                         ("prev_jitter", "<f4", 2)])
POST_PC_DTYPE = np.dtype([("exposure", "<f4"), ("gamma", "<f4"), ("bloom_intensity", "<f4"),
                          ("_padding", "<f4")])


def build(force=False):
    """Compile libashoracle.so with the committed Makefile (g++ -O2 -fopenmp -ffp-contract=off)."""
    srcs = [os.path.join(_HERE, f) for f in ("ash_oracle.cpp", "ash_oracle_instancing.cpp", "ash_oracle_forward.cpp", "ash_oracle_utils.cpp")]
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(f) for f in srcs)):
        return _LIB_PATH
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, env=env,
                   stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.ash_oracle_hiz_build.restype = C.c_size_t
        _lib.ash_oracle_occlusion_cull.restype = C.c_uint32
        _lib.ash_oracle_halton.restype = C.c_float
        _lib.ash_oracle_halton.argtypes = [C.c_uint32, C.c_uint32]
        _lib.ash_oracle_h2f.restype = C.c_float
        _lib.ash_oracle_f2h.restype = C.c_uint16
        _lib.ash_oracle_f2h.argtypes = [C.c_float]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def num_threads():
    return int(lib().ash_oracle_num_threads())


WEIGHTS_Q8, WEIGHTS_EXACT = 0, 1


def set_weight_mode(mode):
    """PIN-6 LINEAR-sampler weights: WEIGHTS_Q8 (8 fractional bits, the library's model, default) or
    WEIGHTS_EXACT (fp32 fraction, SURVEY.md PIN-6 as first written)."""
    lib().ash_oracle_set_weight_mode(int(mode))


# ---------------------------------------------------------------- Hi-Z
def mip_extents(W, H):
    """[(w,h)] for mips 0..levels-1, PIN-3/4: floor halving, floor(log2(max))+1 levels."""
    n = int(lib().ash_oracle_hiz_level_count(int(W), int(H)))
    out, w, h = [(W, H)], W, H
    for _ in range(1, n):
        w, h = max(1, w // 2), max(1, h // 2)
        out.append((w, h))
    return out


def hiz_build(depth):
    """depth: (H,W) f32 -> list of mips [mip1, mip2, ...] (each (h,w) f32) and the flat buffer."""
    depth = _c(depth, np.float32)
    H, W = depth.shape
    ext = mip_extents(W, H)
    total = sum(w * h for (w, h) in ext[1:])
    flat = np.empty(max(total, 1), np.float32)
    n = lib().ash_oracle_hiz_build(_p(depth), W, H, _p(flat))
    assert n == total
    mips, off = [], 0
    for (w, h) in ext[1:]:
        mips.append(flat[off:off + w * h].reshape(h, w))
        off += w * h
    return mips, flat


# ---------------------------------------------------------------- occlusion cull
def occlusion_cull(objs, draws_in, pc, depth, mips_flat=None):
    objs = _c(objs, OBJECT_DTYPE)
    draws_in = _c(draws_in, DRAW_DTYPE)
    pc = _c(pc, CULL_PC_DTYPE).reshape(1)
    depth = _c(depth, np.float32)
    H, W = depth.shape
    if mips_flat is None:
        _, mips_flat = hiz_build(depth)
    n = int(pc["object_count"][0])
    vis = np.zeros(max(n, 1), np.uint32)
    draws_out = np.zeros(max(n, 1), DRAW_DTYPE)
    lib().ash_oracle_occlusion_cull_n.restype = C.c_uint32
    cnt = lib().ash_oracle_occlusion_cull_n(_p(objs), _p(draws_in), C.c_uint32(len(draws_in.reshape(-1))), _p(pc), _p(depth),
                                            _p(mips_flat), W, H, _p(vis), _p(draws_out))
    return vis[:n], draws_out[:cnt], int(cnt)


# ---------------------------------------------------------------- light cull
def light_cull(lights, pc, cam, depth):
    lights = _c(lights, LIGHT_DTYPE)
    pc = _c(pc, LIGHT_PC_DTYPE).reshape(1)
    cam = _c(cam, CAMERA_DTYPE).reshape(1)
    depth = _c(depth, np.float32)
    W, H = int(pc["screen_size"][0][0]), int(pc["screen_size"][0][1])
    assert depth.shape == (H, W)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    tile_data = np.full(tiles * TILE_STRIDE, 0xFFFFFFFF, np.uint32)
    true_counts = np.zeros(tiles, np.uint32)
    minmax = np.zeros(tiles * 2, np.uint32)
    lib().ash_oracle_light_cull(_p(lights), _p(pc), _p(cam), _p(depth), _p(tile_data), _p(true_counts),
                                _p(minmax))
    return tile_data.reshape(tiles, TILE_STRIDE), true_counts, minmax.reshape(tiles, 2)


# ---------------------------------------------------------------- post chain
# Images are (H, W, 4) float16 — RGBA16F, the reference's render-target format — or float32: the fp32_intermediates mode of
# PIN-6, where TAA output / history and the bloom targets are stored unrounded.  `out32` selects the output's storage.
def _img(a):
    a = np.asarray(a)
    return np.ascontiguousarray(a, np.float32 if a.dtype == np.float32 else np.float16)


def _is32(a):
    return 1 if a.dtype == np.float32 else 0


def taa_resolve(current, history, velocity, pc, out32=False):
    """current/history: (H,W,4) float16 or float32; velocity: (H,W,2) float16 or None."""
    current, history = _img(current), _img(history)
    velocity = None if velocity is None else _c(velocity, np.float16)
    pc = _c(pc, TAA_PC_DTYPE).reshape(1)
    out = np.empty(current.shape, np.float32 if out32 else np.float16)
    lib().ash_oracle_taa_resolve_fmt(_p(current), _p(history), _p(velocity), _p(pc), _p(out),
                                     _is32(current) | (_is32(history) << 1) | (int(bool(out32)) << 2))
    return out


def halton(index, base):
    return float(lib().ash_oracle_halton(int(index), int(base)))


def bloom_threshold(hdr, threshold, out32=False):
    hdr = _img(hdr)
    H, W, _ = hdr.shape
    out = np.empty(hdr.shape, np.float32 if out32 else np.float16)
    lib().ash_oracle_bloom_threshold_fmt(_p(hdr), W, H, C.c_float(threshold), _p(out), _is32(hdr) | (int(bool(out32)) << 2))
    return out


def bloom_downsample(src):
    src = _img(src)
    sh, sw, _ = src.shape
    dw, dh = max(1, sw // 2), max(1, sh // 2)
    dst = np.empty((dh, dw, 4), src.dtype)
    lib().ash_oracle_bloom_downsample_fmt(_p(src), sw, sh, _p(dst), dw, dh, _is32(src))
    return dst


def bloom_upsample_add(src, dst, intensity=1.0):
    """returns RN(dst + upsample(src)) in the storage type of src; dst is not modified."""
    src = _img(src)
    out = np.array(dst, dtype=src.dtype, order="C", copy=True)
    sh, sw, _ = src.shape
    dh, dw, _ = out.shape
    lib().ash_oracle_bloom_upsample_add_fmt(_p(src), sw, sh, _p(out), dw, dh, C.c_float(intensity), _is32(src))
    return out


def bloom_chain(hdr, threshold=0.8, mip_count=5, intensity=1.0, out32=False):
    """PIN-9 topology. Returns (U list [U_0..U_mip_count], D list [D_0..D_mip_count]); out32: fp32 bloom targets."""
    D = [bloom_threshold(hdr, threshold, out32=out32)]
    for _ in range(mip_count):
        D.append(bloom_downsample(D[-1]))
    U = [None] * (mip_count + 1)
    U[mip_count] = D[mip_count]
    for k in range(mip_count - 1, -1, -1):
        U[k] = bloom_upsample_add(U[k + 1], D[k], intensity)
    return U, D


TONEMAP_ACES, TONEMAP_REINHARD, TONEMAP_UNCHARTED2, TONEMAP_NONE = range(4)


def tonemap(hdr, bloom, pc, op=TONEMAP_ACES):
    hdr = _img(hdr)
    bloom = None if bloom is None else _img(bloom)
    pc = _c(pc, POST_PC_DTYPE).reshape(1)
    H, W, _ = hdr.shape
    out = np.empty((H, W, 4), np.uint8)
    fmt = _is32(hdr) | ((0 if bloom is None else _is32(bloom)) << 1)
    lib().ash_oracle_tonemap_fmt(_p(hdr), _p(bloom), W, H, _p(pc), C.c_uint32(op), _p(out), fmt)
    return out


# ---- row f-2 (oracle/ash_oracle_instancing.cpp): lod_system.rs + instancing.rs, serially
INSTANCE_DTYPE = np.dtype([("model_row0", "<f4", 4), ("model_row1", "<f4", 4), ("model_row2", "<f4", 4),
                           ("model_row3", "<f4", 4), ("color", "<f4", 4), ("custom", "<f4", 4)])
LOD_LEVEL_DTYPE = np.dtype([("screen_threshold", "<f4"), ("distance_threshold", "<f4"), ("triangle_count", "<u4"),
                            ("vertex_count", "<u4"), ("index_offset", "<u4"), ("index_count", "<u4")])
LOD_MESH_DTYPE = np.dtype([("levels", LOD_LEVEL_DTYPE, 8), ("level_count", "<u4"), ("bounding_radius", "<f4"),
                           ("lod_bias", "<f4"), ("_padding", "<u4")])
LOD_CONFIG_DTYPE = np.dtype([("mode", "<u4"), ("force_level", "<u4"), ("smooth_transitions", "<u4"),
                             ("transition_width", "<f4"), ("global_bias", "<f4"), ("cull_threshold", "<f4"),
                             ("use_visibility", "<u4"), ("_padding", "<u4")])
LOD_SELECTION_DTYPE = np.dtype([("level", "<u4"), ("blend", "<f4"), ("culled", "<u4"), ("screen_coverage", "<f4")])
INSTANCING_STATS_DTYPE = np.dtype([("objects_per_lod", "<u4", 8), ("objects_culled", "<u4"), ("total_instances", "<u4"),
                                   ("batch_count", "<u4"), ("avg_instances_per_batch", "<f4"),
                                   ("triangles_before", "<u8"), ("triangles_after", "<u8"),
                                   ("instances_dropped", "<u4"), ("_padding", "<u4")])
INSTANCING_PARAMS_DTYPE = np.dtype([("view_proj", "<f4", (4, 4)), ("camera_pos", "<f4", 4), ("screen_size", "<f4", 2),
                                    ("object_count", "<u4"), ("_padding", "<u4")])


def screen_coverage(view_proj_cols, position, radius, width, height):
    """LodManager::calculate_screen_coverage (lod_system.rs:207-235); view_proj as 4 columns."""
    f = lib().ash_oracle_screen_coverage
    f.restype = C.c_float
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float]
    vp = _c(view_proj_cols, np.float32)
    pos = _c(position, np.float32)
    return float(f(_p(vp), _p(pos), radius, width, height))


def instancing(instances, batch_id, mesh_id, meshes, params, cfg, n_batches, visibility=None):
    """Whole frame of select_lod + add_instance + finalize; returns (sel, out, offsets, draws, stats)."""
    inst, bid, mid = _c(instances, INSTANCE_DTYPE), _c(batch_id, np.uint32), _c(mesh_id, np.uint32)
    ms, ip, k = _c(meshes, LOD_MESH_DTYPE), _c(params, INSTANCING_PARAMS_DTYPE), _c(cfg, LOD_CONFIG_DTYPE)
    n = int(ip.reshape(-1)[0]["object_count"])
    nb = int(n_batches) * 8
    vis = None if visibility is None else _c(visibility, np.uint32)
    sel = np.zeros(n, LOD_SELECTION_DTYPE)
    out = np.zeros(max(n, 1), INSTANCE_DTYPE)
    offsets = np.zeros(nb + 1, np.uint32)
    draws = np.zeros(nb, DRAW_DTYPE)
    stats = np.zeros(1, INSTANCING_STATS_DTYPE)
    f = lib().ash_oracle_instancing
    f.restype = None
    f.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32] + [C.c_void_p] * 5
    f(_p(inst), _p(bid), _p(mid), _p(ms), len(ms.reshape(-1)), None if vis is None else _p(vis), _p(ip), _p(k), int(n_batches),
      _p(sel), _p(out), _p(offsets), _p(draws), _p(stats))
    return sel, out[:int(offsets[-1])], offsets, draws, stats[0]


# ---- row f-1: compute forward pass (oracle/ash_oracle_forward.cpp)
FORWARD_PARAMS_DTYPE = np.dtype([("light_direction", "<f4", 4), ("light_color", "<f4", 4), ("ambient_color", "<f4", 4),
                                 ("base_color_factor", "<f4", 4), ("emissive_factor", "<f4", 4), ("parameters", "<f4", 4),
                                 ("background", "<f4", 4), ("near_plane", "<f4"), ("radius_scale", "<f4"), ("flags", "<u4"),
                                 ("_padding", "<u4")])


def forward_depth(objs, cam, fp, W, H, brute=False, mask=None, keys_in=None):
    """Nearest occluder-sphere hit per pixel: (keys u64 HxW, depth f32 HxW).  brute=True tests every pixel
    against every sphere (no silhouette rectangle): the check that the rectangle is conservative.
    mask: draw only these objects; keys_in: draw on top of an earlier pass of the same view (two-phase culling)."""
    o, c, p = _c(objs, OBJECT_DTYPE), _c(cam, CAMERA_DTYPE), _c(fp, FORWARD_PARAMS_DTYPE)
    keys = np.zeros((H, W), np.uint64) if keys_in is None else np.ascontiguousarray(keys_in, np.uint64).copy()
    depth = np.zeros((H, W), np.float32)
    m = None if mask is None else _c(mask, np.uint32)
    f = lib().ash_oracle_forward_depth
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_int,
                  C.c_void_p, C.c_void_p]
    f(_p(o), len(o.reshape(-1)), _p(c), _p(p), W, H, int(bool(brute)), None if m is None else _p(m), int(keys_in is not None),
      _p(keys), _p(depth))
    return keys, depth


def two_phase_cull(objs, draws, cull_pc, cam, fp, W, H, prev_depth=None):
    """Row f-3, the procedure of ashcuda_view_run with ASH_STAGE_TWO_PHASE, from the single-phase oracle pieces:
    phase 1 = occlusion_cull against the previous frame's depth pyramid (a first frame has none: a far-plane
    depth buffer occludes nothing, leaving the frustum test); draw its survivors; phase 2 = occlusion_cull of
    the rest against the pyramid of that depth; draw the late objects.  Returns a dict."""
    n = int(np.asarray(cull_pc).reshape(-1)[0]["object_count"])
    d0 = np.ones((H, W), np.float32) if prev_depth is None else prev_depth
    vis1, _, early = occlusion_cull(objs, draws, cull_pc, d0)
    keys, depth1 = forward_depth(objs[:n], cam, fp, W, H, mask=vis1)
    vis_a, _, _ = occlusion_cull(objs, draws, cull_pc, depth1)
    late = (vis_a != 0) & (vis1 == 0)
    vis = ((vis1 != 0) | late).astype(np.uint32)
    keys, depth = forward_depth(objs[:n], cam, fp, W, H, mask=late.astype(np.uint32), keys_in=keys)
    o = _c(objs, OBJECT_DTYPE).reshape(-1)[:n]
    draws_out = _c(draws, DRAW_DTYPE).reshape(-1)[o["draw_index"][vis != 0]]
    return dict(visibility=vis, late=late.astype(np.uint32), early_count=int(early), late_count=int(late.sum()),
                count=int(vis.sum()), draws_out=draws_out, keys=keys, depth=depth, depth_phase1=depth1)


def forward_velocity(keys, objs, cam, fp, reproj_cols):
    """RG16F velocity (H, W, 2) of the forward pass: uv - prevUV, prevUV from reproj * (view-space hit, 1)."""
    H, W = keys.shape
    k, o, c, p = _c(keys, np.uint64), _c(objs, OBJECT_DTYPE), _c(cam, CAMERA_DTYPE), _c(fp, FORWARD_PARAMS_DTYPE)
    m = _c(reproj_cols, np.float32).reshape(4, 4)
    vel = np.zeros((H, W, 2), np.float16)
    f = lib().ash_oracle_forward_velocity
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    f(_p(k), _p(o), len(o.reshape(-1)), _p(c), _p(p), _p(m), W, H, _p(vel))
    return vel


def forward_shade(keys, objs, lights, tile_data, cam, fp, instances=None):
    """frag.frag's BRDF at every hit for the MVP light + the tile light lists: RGBA16F HxWx4."""
    H, W = keys.shape
    k, o, l = _c(keys, np.uint64), _c(objs, OBJECT_DTYPE), _c(lights, LIGHT_DTYPE)
    td = _c(tile_data, np.uint32)
    c, p = _c(cam, CAMERA_DTYPE), _c(fp, FORWARD_PARAMS_DTYPE)
    inst = None if instances is None else _c(instances, INSTANCE_DTYPE)
    hdr = np.zeros((H, W, 4), np.float16)
    f = lib().ash_oracle_forward_shade
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32,
                  C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    f(_p(k), _p(o), len(o.reshape(-1)), _p(l), len(l.reshape(-1)), _p(td), None if inst is None else _p(inst),
      0 if inst is None else len(inst.reshape(-1)), _p(c), _p(p), W, H, _p(hdr))
    return hdr


# ---- row f-4: shaders/compute_utils.glsl (oracle/ash_oracle_utils.cpp); op numbering = ASH_CU_* of ashcuda.h
CU_MIN, CU_MAX, CU_SUM, CU_SCAN, CU_SPHERE_IN_FRUSTUM, CU_AABB_IN_FRUSTUM, CU_DEPTH, CU_PLANES = range(8)


def compute_utils(op, data=None, params=None):
    p = np.zeros(18, np.float32)
    if params is not None:
        q = np.asarray(params, np.float32).reshape(-1)
        p[:len(q)] = q
    f = lib().ash_oracle_compute_utils
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    if op == CU_PLANES:
        out = np.zeros((6, 4), np.float32)
        assert f(op, None, 0, _p(p), _p(out)) == 0
        return out
    a = np.ascontiguousarray(data, np.uint32 if op == CU_SCAN else np.float32)
    n = a.shape[0]
    out = (np.zeros(n // 256, np.float32) if op <= CU_SUM else np.zeros(n, np.uint32) if op in (CU_SCAN, CU_SPHERE_IN_FRUSTUM, CU_AABB_IN_FRUSTUM)
           else np.zeros((n, 4), np.float32))
    assert f(op, _p(a), n, _p(p), _p(out)) == 0
    return out

