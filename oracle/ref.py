"""ctypes/numpy front-end of oracle/_ref/libashref.so: the REFERENCE'S OWN GLSL compiled as C++ (oracle/ref/).

TEST INFRASTRUCTURE ONLY.  Same call signatures as oracle/oracle.py for the eight shaders on the path, so the
restatement (oracle/ash_oracle.cpp) can be checked against the reference's text bit for bit
(tests/test_oracle_ref.py) and the golden vectors generated from it (tests/golden/make_golden.py).
The library is built here from /root/reference/shaders (oracle/ref/Makefile); on a machine without the
reference tree the prebuilt, git-ignored oracle/_ref/libashref.so is used, and `available()` says whether it exists.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import oracle as _o

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libashref.so")
REFERENCE_SHADERS = "/root/reference/shaders"

_lib = None


def build():
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.run(["make", "-C", os.path.join(_HERE, "ref"), "-s"], check=True, env=env, stdout=subprocess.DEVNULL)
    return _LIB_PATH


def available():
    if os.path.isdir(REFERENCE_SHADERS):
        build()
    return os.path.exists(_LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libashref.so is absent and /root/reference is not here to build it")
        _lib = C.CDLL(_LIB_PATH)
        _lib.ash_ref_hiz_build.restype = C.c_size_t
        _lib.ash_ref_occlusion_cull.restype = C.c_uint32
        _lib.ash_ref_h2f.restype = C.c_float
        _lib.ash_ref_f2h.restype = C.c_uint16
        _lib.ash_ref_f2h.argtypes = [C.c_float]
    return _lib


_p, _c = _o._p, _o._c
WEIGHTS_Q8, WEIGHTS_EXACT = 0, 1


def set_weight_mode(mode):
    """PIN-6 bilinear weights of the compat sampler: WEIGHTS_EXACT or WEIGHTS_Q8 (default)."""
    lib().ash_ref_set_weight_mode(int(mode))


def hiz_build(depth):
    depth = _c(depth, np.float32)
    H, W = depth.shape
    ext = _o.mip_extents(W, H)
    total = sum(w * h for (w, h) in ext[1:])
    flat = np.empty(max(total, 1), np.float32)
    n = lib().ash_ref_hiz_build(_p(depth), W, H, _p(flat))
    assert n == total
    mips, off = [], 0
    for (w, h) in ext[1:]:
        mips.append(flat[off:off + w * h].reshape(h, w))
        off += w * h
    return mips, flat


def occlusion_cull(objs, draws_in, pc, depth, mips_flat=None):
    objs = _c(objs, _o.OBJECT_DTYPE)
    draws_in = _c(draws_in, _o.DRAW_DTYPE)
    pc = _c(pc, _o.CULL_PC_DTYPE).reshape(1)
    depth = _c(depth, np.float32)
    H, W = depth.shape
    if mips_flat is None:
        _, mips_flat = hiz_build(depth)
    n = int(pc["object_count"][0])
    vis = np.zeros(max(n, 1), np.uint32)
    draws_out = np.zeros(max(n, 1), _o.DRAW_DTYPE)
    cnt = lib().ash_ref_occlusion_cull(_p(objs), _p(draws_in), _p(pc), _p(depth), _p(mips_flat), W, H, _p(vis),
                                       _p(draws_out))
    return vis[:n], draws_out[:cnt], int(cnt)


def light_cull(lights, pc, cam, depth):
    lights = _c(lights, _o.LIGHT_DTYPE)
    pc = _c(pc, _o.LIGHT_PC_DTYPE).reshape(1)
    cam = _c(cam, _o.CAMERA_DTYPE).reshape(1)
    depth = _c(depth, np.float32)
    W, H = int(pc["screen_size"][0][0]), int(pc["screen_size"][0][1])
    assert depth.shape == (H, W)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    tile_data = np.full(tiles * _o.TILE_STRIDE, 0xFFFFFFFF, np.uint32)
    true_counts = np.zeros(tiles, np.uint32)
    lib().ash_ref_light_cull(_p(lights), _p(pc), _p(cam), _p(depth), _p(tile_data), _p(true_counts))
    return tile_data.reshape(tiles, _o.TILE_STRIDE), true_counts


def taa_resolve(current, history, velocity, pc, out32=False):
    current, history = _o._img(current), _o._img(history)
    velocity = None if velocity is None else _c(velocity, np.float16)
    pc = _c(pc, _o.TAA_PC_DTYPE).reshape(1)
    out = np.zeros(current.shape, np.float32 if out32 else np.float16)
    lib().ash_ref_taa_resolve_fmt(_p(current), _p(history), _p(velocity), _p(pc), _p(out),
                                  _o._is32(current) | (_o._is32(history) << 1) | (int(bool(out32)) << 2))
    return out


def bloom_threshold(hdr, threshold, out32=False):
    hdr = _o._img(hdr)
    H, W, _ = hdr.shape
    out = np.empty(hdr.shape, np.float32 if out32 else np.float16)
    lib().ash_ref_bloom_threshold_fmt(_p(hdr), W, H, C.c_float(threshold), _p(out), _o._is32(hdr) | (int(bool(out32)) << 2))
    return out


def bloom_downsample(src):
    src = _o._img(src)
    sh, sw, _ = src.shape
    dw, dh = max(1, sw // 2), max(1, sh // 2)
    dst = np.empty((dh, dw, 4), src.dtype)
    lib().ash_ref_bloom_downsample_fmt(_p(src), sw, sh, _p(dst), dw, dh, _o._is32(src))
    return dst


def bloom_upsample_add(src, dst, intensity=1.0):
    src = _o._img(src)
    out = np.array(dst, dtype=src.dtype, order="C", copy=True)
    sh, sw, _ = src.shape
    dh, dw, _ = out.shape
    lib().ash_ref_bloom_upsample_add_fmt(_p(src), sw, sh, _p(out), dw, dh, C.c_float(intensity), _o._is32(src))
    return out


def bloom_chain(hdr, threshold=0.8, mip_count=5, intensity=1.0, out32=False):
    D = [bloom_threshold(hdr, threshold, out32=out32)]
    for _ in range(mip_count):
        D.append(bloom_downsample(D[-1]))
    U = [None] * (mip_count + 1)
    U[mip_count] = D[mip_count]
    for k in range(mip_count - 1, -1, -1):
        U[k] = bloom_upsample_add(U[k + 1], D[k], intensity)
    return U, D


def tonemap(hdr, bloom, pc):
    hdr = _o._img(hdr)
    bloom = None if bloom is None else _o._img(bloom)
    pc = _c(pc, _o.POST_PC_DTYPE).reshape(1)
    H, W, _ = hdr.shape
    out = np.empty((H, W, 4), np.uint8)
    lib().ash_ref_tonemap_fmt(_p(hdr), _p(bloom), W, H, _p(pc), _p(out), _o._is32(hdr) | ((0 if bloom is None else _o._is32(bloom)) << 1))
    return out
