"""Pins for the CPU oracle (oracle/ash_oracle.cpp).  The reference holds no golden vectors for
the shaders' arithmetic, so the oracle is pinned by (a) the reference's own
known answers at this boundary, (b) independent numpy/float64 re-evaluations of each shader on
small inputs, (c) property tests, and (d) committed regression vectors (tests/golden)."""
import math
import os

import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_small.npz")


# ------------------------------------------------------------------ fp16 helpers
def test_fp16_conversions_match_numpy(oracle):
    lib = oracle.lib()
    allh = np.arange(65536, dtype=np.uint16)
    out = np.empty(65536, np.float32)
    lib.ash_oracle_h2f_array(allh.ctypes.data, out.ctypes.data, 65536)
    ref = allh.view(np.float16).astype(np.float32)
    assert np.array_equal(out.view(np.uint32)[~np.isnan(ref)], ref.view(np.uint32)[~np.isnan(ref)])
    rng = np.random.default_rng(0)
    f = np.concatenate([rng.standard_normal(200000).astype(np.float32) * 10.0 ** rng.integers(-9, 6, 200000),
                        np.array([0, -0.0, 65504, 65519.99, 65520, 1e-8, 5.96e-8, 2.98e-8, 2.9803e-8, 6.1e-5,
                                  np.inf, -np.inf], np.float32),
                        ref[~np.isnan(ref)]]).astype(np.float32)
    # ties: exactly halfway between two halfs
    h = rng.integers(0, 0x7BFF, 20000).astype(np.uint16)
    mid = ((h.view(np.float16).astype(np.float64) + (h + 1).astype(np.uint16).view(np.float16).astype(np.float64)) / 2)
    f = np.concatenate([f, mid.astype(np.float32)])
    got = np.empty(len(f), np.uint16)
    lib.ash_oracle_f2h_array(f.ctypes.data, got.ctypes.data, len(f))
    with np.errstate(over="ignore"):
        assert np.array_equal(got, f.astype(np.float16).view(np.uint16))


# ------------------------------------------------------------------ Hi-Z
def test_hiz_levels_and_extents(oracle):
    assert len(oracle.mip_extents(1920, 1080)) == 11          # PIN-4
    assert len(oracle.mip_extents(3840, 2160)) == 12
    e = oracle.mip_extents(1920, 1080)
    assert e[1:] == [(960, 540), (480, 270), (240, 135), (120, 67), (60, 33), (30, 16), (15, 8), (7, 4), (3, 2), (1, 1)]
    assert sum(w * h for w, h in e[1:]) == 691055
    assert sum(w * h for w, h in oracle.mip_extents(3840, 2160)[1:]) == 2764655


@pytest.mark.parametrize("W,H", [(64, 64), (67, 45), (1, 1), (2, 1), (1, 7), (130, 70), (255, 257)])
def test_hiz_matches_numpy_2x2_max(oracle, W, H):
    rng = np.random.default_rng(W * 1000 + H)
    d = rng.random((H, W), dtype=np.float32)
    mips, _ = oracle.hiz_build(d)
    src = d
    for m in mips:
        h, w = m.shape
        assert (w, h) == (max(1, src.shape[1] // 2), max(1, src.shape[0] // 2))
        ys = np.minimum(np.arange(h)[:, None] * 2 + np.array([0, 1])[None, :], src.shape[0] - 1)
        xs = np.minimum(np.arange(w)[:, None] * 2 + np.array([0, 1])[None, :], src.shape[1] - 1)
        ref = src[ys[:, None, :, None], xs[None, :, None, :]].max(axis=(2, 3))
        assert np.array_equal(m, ref)
        src = m
    assert len(mips) == len(oracle.mip_extents(W, H)) - 1


def test_hiz_texel_bounds_the_covered_depth(oracle):
    d = synth.random_depth(256, 128, seed=3)
    mips, _ = oracle.hiz_build(d)
    for lvl, m in enumerate(mips, start=1):
        s = 1 << lvl
        for (y, x) in [(0, 0), (m.shape[0] - 1, m.shape[1] - 1), (m.shape[0] // 2, m.shape[1] // 3)]:
            assert m[y, x] == d[y * s:(y + 1) * s, x * s:(x + 1) * s].max()


# ------------------------------------------------------------------ occlusion cull
def _np_cull(objs, pc, depth, mips):
    """float64-free numpy restatement of occlusion_cull.comp, fp32 ops in the shader's order."""
    f = np.float32
    vp = pc["view_proj"][0].astype(f)            # vp[c] = column c
    W, H = int(pc["screen_params"][0][0]), int(pc["screen_params"][0][1])
    levels = int(pc["hiz_levels"][0])
    planes = []
    for i in range(3):
        for sgn in (+1, -1):
            p = vp[3] + vp[i] if sgn > 0 else vp[3] - vp[i]
            ln = np.sqrt(f(f(p[0] * p[0] + p[1] * p[1]) + p[2] * p[2]))
            planes.append(p / ln)

    def mulv(m, v):
        return f(f(f(m[0] * v[0] + m[1] * v[1]) + m[2] * v[2]) + m[3] * v[3])

    pyr = [depth] + list(mips)
    out = np.zeros(len(objs), np.uint32)
    for k, o in enumerate(objs):
        if not (o["flags"] & 1):
            continue
        m = o["model"].astype(f)
        c = o["center"][:3].astype(f)
        wc = np.array([f(f(f(m[0][r] * c[0] + m[1][r] * c[1]) + m[2][r] * c[2]) + m[3][r] * f(1)) for r in range(3)], f)
        ln = [np.sqrt(f(f(m[j][0] * m[j][0] + m[j][1] * m[j][1]) + m[j][2] * m[j][2])) for j in range(3)]
        we = o["extents"][:3].astype(f) * np.array(ln, f)
        culled = False
        for p in planes:
            pv = wc + we * np.sign(p[:3]).astype(f)
            if f(f(f(p[0] * pv[0] + p[1] * pv[1]) + p[2] * pv[2]) + p[3]) < 0:
                culled = True
                break
        if culled:
            continue
        sx = [-1, 1, 1, -1, -1, 1, 1, -1]
        sy = [-1, -1, 1, 1, -1, -1, 1, 1]
        sz = [-1, -1, -1, -1, 1, 1, 1, 1]
        mn, mx, mnz, behind = np.array([1, 1], f), np.array([0, 0], f), f(1), False
        for j in range(8):
            v = np.array([wc[0] + f(sx[j]) * we[0], wc[1] + f(sy[j]) * we[1], wc[2] + f(sz[j]) * we[2], 1], f)
            clip = np.array([mulv(vp[:, r], v) for r in range(4)], f)
            if clip[3] <= 0:
                behind = True
                break
            ndc = clip[:3] / clip[3]
            s = ndc[:2] * f(0.5) + f(0.5)
            mn, mx, mnz = np.minimum(mn, s), np.maximum(mx, s), min(mnz, ndc[2])
        if behind:
            out[k] = 1
            continue
        mn, mx = np.clip(mn, 0, 1), np.clip(mx, 0, 1)
        ss = (mx - mn) * np.array([W, H], f)
        smax = float(max(ss[0], ss[1]))
        lvl = 0 if smax <= 0 else int(min(max(math.ceil(math.log2(smax)), 0), levels - 1))
        uv = (mn + mx) * f(0.5)
        img = pyr[lvl]
        ty = int(min(max(math.floor(float(f(uv[1] * f(img.shape[0])))), 0), img.shape[0] - 1))
        tx = int(min(max(math.floor(float(f(uv[0] * f(img.shape[1])))), 0), img.shape[1] - 1))
        out[k] = 0 if mnz > img[ty, tx] else 1
    return out


def _cull_case(W, H, grid, seed, eye=(0.0, 2.0, 22.0), disabled=0.1):
    cam = synth.Camera(eye, (0, 0, 0), W, H)
    objs, draws = synth.make_objects(grid, seed=seed, jitter=0.3, disabled_frac=disabled)
    depth = synth.splat_depth(objs, cam)
    return cam, objs, draws, depth


def test_cull_matches_independent_numpy_restatement(oracle):
    W, H = 320, 180
    cam, objs, draws, depth = _cull_case(W, H, (6, 6, 6), seed=11)
    mips, flat = oracle.hiz_build(depth)
    pc = cam.cull_pc(len(objs), min(10, len(mips) + 1))
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth, flat)
    with np.errstate(all="ignore"):
        ref = _np_cull(objs, pc, depth, mips)
    assert np.array_equal(vis, ref)
    assert 0 < cnt < len(objs) and cnt == int(vis.sum())
    # PIN-7: draw commands in ascending object order
    assert np.array_equal(dout, draws[objs["draw_index"][vis == 1]])


def test_cull_semantics_edge_cases(oracle):
    W, H = 128, 96
    cam = synth.Camera((0, 0, 10), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((1, 1, 1), seed=1)
    objs = np.repeat(objs, 6)
    objs["model"][1][3][:3] = [0, 0, 50]       # behind the camera: not inside the (column-built) frustum
    objs["model"][2][3][:3] = [0, 0, 9.9]      # straddles the camera plane: clip.w <= 0 -> visible
    objs["flags"][3] = 0                       # disabled
    objs["flags"][4] = 2                       # bit 0 clear
    objs["extents"][5] = 0                     # zero-size AABB: screenSize 0 -> mip 0
    depth = np.full((H, W), 1.0, np.float32)
    pc = cam.cull_pc(len(objs), 8)
    vis, _, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    assert vis[0] == 1 and vis[3] == 0 and vis[4] == 0 and vis[5] == 1
    # an occluder plane in front of everything hides the unit cube at the origin
    depth[:] = 0.0
    vis2, _, _ = oracle.occlusion_cull(objs, draws, pc, depth)
    assert vis2[0] == 0
    # n = 0
    vis0, d0, c0 = oracle.occlusion_cull(objs[:0], draws, cam.cull_pc(0, 8), depth)
    assert c0 == 0 and len(vis0) == 0 and len(d0) == 0


def test_exact_ceil_log2_mip_selection(oracle):
    # A screen rect of exactly 2^k pixels selects mip k; 2^k + 2^-10 px selects k+1
    W = H = 1024
    depth = np.full((H, W), 1.0, np.float32)
    mips, flat = oracle.hiz_build(depth)
    # make every level distinguishable: level l holds 0.5 + l/100 so minZ>hiz flips at a known level
    off = 0
    for l, m in enumerate(mips, start=1):
        flat[off:off + m.size] = 0.5 + l / 100.0
        off += m.size
    depth[:] = 0.5
    ident = np.eye(4, dtype=np.float32)
    for k in (1, 3, 6):
        for bump, expect_level in ((0, k), (1, k + 1)):
            half_ndc = np.float32((2.0 ** k) / W)           # rect of 2^k pixels in ndc units (x0.5 later)
            # the bump must survive 0.5 +- ext/2 in fp32 (ulp 2^-24 there): 2^-20 does, 1 ulp of ext would not
            ext = np.float32(half_ndc + np.float32(2.0 ** -20)) if bump else half_ndc
            objs = np.zeros(1, T.CULL_OBJECT)
            objs["model"][0] = ident
            objs["extents"][0][:3] = [ext, ext / 4, 0.0]
            objs["center"][0][:3] = [0.0, 0.0, 0.5 + expect_level / 100.0 - 0.001]
            objs["flags"] = 1
            pc = np.zeros(1, T.CULLING_PC)
            pc["view_proj"][0] = ident
            pc["screen_params"][0] = [W, H, 1 / W, 1 / H]
            pc["object_count"], pc["hiz_levels"] = 1, 11
            draws = np.zeros(1, T.DRAW_COMMAND)
            vis, _, _ = oracle.occlusion_cull(objs, draws, pc, depth, flat)
            assert vis[0] == 1        # minZ just below the selected level's value
            objs["center"][0][2] = 0.5 + expect_level / 100.0 + 0.001
            vis, _, _ = oracle.occlusion_cull(objs, draws, pc, depth, flat)
            assert vis[0] == 0        # minZ just above: occluded at exactly that level


# ------------------------------------------------------------------ light cull
def _light_case(W, H, n_lights, seed, eye=(0.0, 2.0, 22.0), extent=10.0):
    cam = synth.Camera(eye, (0, 0, 0), W, H)
    objs, _ = synth.make_objects((5, 5, 5), seed=seed)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(n_lights, extent=extent, seed=seed, center=eye)
    return cam, depth, lights


def test_tile_counts_known_answer():
    from ash_renderer_b200.host import LightCullingPass
    p = LightCullingPass()
    p.calculate_tiles(1920, 1080)                 # features/light_culling.rs:270-276
    assert (p.tiles_x, p.tiles_y) == (120, 68)
    assert p.get_tile_buffer_size() == 8388480


def test_light_lists_against_float64_reevaluation(oracle):
    W, H = 200, 120           # ragged: 13 x 8 tiles, last column 8 px wide, last row 8 px high
    cam, depth, lights = _light_case(W, H, 300, seed=5)
    td, true_counts, minmax = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    tiles_x, tiles_y = (W + 15) // 16, (H + 15) // 16
    assert td.shape == (tiles_x * tiles_y, 257)
    cd = cam.camera_data()
    view = cd["view"][0].astype(np.float64)
    ip = cd["inv_projection"][0].astype(np.float64)
    pos = lights["position"].astype(np.float64)
    lv = pos[:, :3] @ view[:3, :3] + view[3, :3]
    knife = 0
    for t in range(tiles_x * tiles_y):
        tx, ty = t % tiles_x, t // tiles_x
        blk = depth[ty * 16:min(H, ty * 16 + 16), tx * 16:min(W, tx * 16 + 16)]
        assert minmax[t, 0] == blk.min().view(np.uint32) and minmax[t, 1] == blk.max().view(np.uint32)
        ndc = lambda i, n: (i * 16.0) / n * 2.0 - 1.0
        cs = []
        for (cx, cy) in ((ndc(tx, W), ndc(ty, H)), (ndc(tx + 1, W), ndc(ty, H)), (ndc(tx + 1, W), ndc(ty + 1, H)), (ndc(tx, W), ndc(ty + 1, H))):
            q = np.array([cx, cy, 0.0, 1.0]) @ ip
            cs.append(q[:3] / q[3])
        ns = [np.cross(cs[3], cs[0]), np.cross(cs[1], cs[2]), np.cross(cs[0], cs[1]), np.cross(cs[2], cs[3])]
        ns = [n / np.linalg.norm(n) for n in ns]
        dist = np.stack([lv @ n for n in ns] + [-lv[:, 2] - float(blk.min()), lv[:, 2] + float(blk.max())], 1)
        margin = (dist + pos[:, 3:4]).min(1)
        enabled = lights["params"][:, 3] >= 0.5
        sure_in = set(np.nonzero(enabled & (margin > 1e-4))[0])
        maybe = set(np.nonzero(enabled & (np.abs(margin) <= 1e-4))[0])
        got = td[t, 1:1 + td[t, 0]]
        assert np.all(np.diff(got.astype(np.int64)) > 0)          # PIN-7 ascending
        assert sure_in <= set(got) <= (sure_in | maybe)
        knife += len(maybe)
        assert true_counts[t] == td[t, 0]
    assert knife < 50
    assert td[:, 0].max() > 0


def test_light_overflow_keeps_first_256_and_directionals_always_pass(oracle):
    W, H = 64, 48
    cam = synth.Camera((0, 0, 5), (0, 0, 0), W, H)
    depth = np.full((H, W), 1.0, np.float32)
    lights = synth.make_lights(700, extent=0.2, rmin=50, rmax=60, seed=9, n_directional=3, enabled_p=0.9,
                               center=(0, 0, 5))
    td, true_counts, _ = oracle.light_cull(lights, cam.light_pc(700), cam.camera_data(), depth)
    enabled = np.nonzero(lights["params"][:, 3] >= 0.5)[0]
    assert np.all(true_counts == len(enabled)) and len(enabled) > 256
    assert np.all(td[:, 0] == 256)
    assert np.array_equal(td[0, 1:257], enabled[:256])
    # zero lights: count 0 written, rest untouched (PIN-8)
    td0, _, _ = oracle.light_cull(lights[:0], cam.light_pc(0), cam.camera_data(), depth)
    assert np.all(td0[:, 0] == 0) and np.all(td0[:, 1:] == 0xFFFFFFFF)


# ------------------------------------------------------------------ TAA
def _taa_pc(W, H, blend=0.9, clamp=1.0, vel=1.0, jitter=(0.25, -0.1666667), prev=(0.0, 0.0)):
    pc = np.zeros(1, T.TAA_PC)
    w, h = np.float32(W), np.float32(H)
    pc["screen_params"][0] = [w, h, np.float32(1) / w, np.float32(1) / h]
    pc["params"][0] = [blend, clamp, vel, 0.0]
    pc["jitter"][0], pc["prev_jitter"][0] = jitter, prev
    return pc


def test_halton_known_answers(oracle):
    # temporal_aa.rs:52-58 / SURVEY §4: (0.5,1/3),(0.25,2/3),(0.75,1/9) ...
    assert oracle.halton(1, 2) == 0.5 and oracle.halton(2, 2) == 0.25 and oracle.halton(3, 2) == 0.75
    assert oracle.halton(1, 3) == np.float32(1) / np.float32(3)
    from ash_renderer_b200.host import HaltonSequence
    hs = HaltonSequence()
    seq = [hs.next_jitter() for _ in range(32)]
    assert seq[:16] == seq[16:]                                   # period 16
    assert all(-0.5 <= j[0] <= 0.5 and -0.5 <= j[1] <= 0.5 for j in seq)
    assert seq[0] != seq[1]                                       # temporal_aa.rs:221-229
    for i in range(1, 17):
        assert float(HaltonSequence.halton(i, 2)) == oracle.halton(i, 2)
        assert float(HaltonSequence.halton(i, 3)) == oracle.halton(i, 3)


def test_taa_properties(oracle):
    W, H = 96, 64
    cur = synth.random_hdr(W, H, seed=1)
    hist = synth.random_hdr(W, H, seed=2)
    # blend = 0 returns current (alpha forced to 1)
    out = oracle.taa_resolve(cur, hist, None, _taa_pc(W, H, blend=0.0))
    assert np.array_equal(out[..., :3], cur[..., :3]) and np.all(out[..., 3] == 1.0)
    # constant images stay constant (within the fp16 rounding of the tonemap round trip)
    const = np.full((H, W, 4), 0.75, np.float16)
    out = oracle.taa_resolve(const, const, None, _taa_pc(W, H))
    assert np.max(np.abs(out[..., :3].astype(np.float32) - 0.75)) <= 2e-3
    # clamp off + zero jitter: exact mix(current, history, blend) at texel centres
    out = oracle.taa_resolve(cur, hist, None, _taa_pc(W, H, clamp=0.0, vel=0.0, jitter=(0, 0)))
    c, h = cur[..., :3].astype(np.float32), hist[..., :3].astype(np.float32)
    ref = (c * (np.float32(1) - np.float32(0.9)) + h * np.float32(0.9)).astype(np.float16)
    assert np.array_equal(out[..., :3], ref)
    # neighbourhood clamp: where the 3x3 neighbourhood is dim, a wildly bright history is pulled back
    hot = hist.copy()
    hot[..., :3] = 500.0
    out = oracle.taa_resolve(cur, hot, None, _taa_pc(W, H)).astype(np.float32)
    pad = np.pad(cur[..., :3].astype(np.float32), ((1, 1), (1, 1), (0, 0)), mode="edge")
    nb_max = np.max([pad[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=0).max(-1)
    dim = nb_max < 2.0                      # tonemapped < 2/3: the 1.25x AABB stays well below 1
    assert dim.sum() > 100 and np.all(out[..., :3].max(-1)[dim] < 20.0)
    # velocity rejection: 10+ px of motion forces blend to 0
    vel = np.zeros((H, W, 2), np.float16)
    vel[..., 0] = 12.0 / W
    out = oracle.taa_resolve(cur, hist, vel, _taa_pc(W, H))
    assert np.array_equal(out[..., :3], cur[..., :3])


def test_taa_matches_numpy_restatement_on_texel_centres(oracle):
    """Independent numpy fp32 restatement (zero jitter delta => every tap is a texel centre)."""
    W, H = 40, 24
    f = np.float32
    cur = synth.random_hdr(W, H, seed=7)
    hist = synth.random_hdr(W, H, seed=8)
    out = oracle.taa_resolve(cur, hist, None, _taa_pc(W, H, jitter=(0.25, 0.25), prev=(0.25, 0.25)))
    c = cur[..., :3].astype(f)
    h = hist[..., :3].astype(f)

    def ycocg(x):
        r, g, b = x[..., 0], x[..., 1], x[..., 2]
        return np.stack([f(0.25) * r + f(0.5) * g + f(0.25) * b, f(0.5) * r - f(0.5) * b,
                         f(-0.25) * r + f(0.5) * g - f(0.25) * b], -1)

    tm = lambda x: x / (f(1) + x)
    pad = np.pad(c, ((1, 1), (1, 1), (0, 0)), mode="edge")
    nb = np.stack([ycocg(tm(pad[dy:dy + H, dx:dx + W])) for dy in range(3) for dx in range(3)])
    mn, mx = nb.min(0), nb.max(0)
    ctr, ext = (mn + mx) * f(0.5), (mx - mn) * f(0.5)
    mn, mx = ctr - ext * f(1.25), ctr + ext * f(1.25)
    hy = ycocg(tm(h))
    center = f(0.5) * (mx + mn)
    half = f(0.5) * (mx - mn) + f(0.001)
    delta = hy - center
    v = np.abs(delta / half).max(-1, keepdims=True)
    clipped = np.where(v > 1, center + delta / v, hy)
    y, co, cg = clipped[..., 0], clipped[..., 1], clipped[..., 2]
    rgb = np.stack([y + co - cg, y + cg, y - co - cg], -1)
    hh = rgb / (f(1) - rgb + f(0.0001))
    ref = (c * (f(1) - f(0.9)) + hh * f(0.9)).astype(np.float16)
    assert np.array_equal(out[..., :3], ref)


# ------------------------------------------------------------------ bloom + tonemap
def test_bloom_properties(oracle):
    W, H = 64, 40
    dark = np.full((H, W, 4), 0.2, np.float16)
    U, D = oracle.bloom_chain(dark, 0.8, 3)
    assert all(np.all(u[..., :3] == 0) for u in U)              # below threshold-knee => exactly 0
    bright = np.full((H, W, 4), 4.0, np.float16)
    t = oracle.bloom_threshold(bright, 0.8)
    assert np.allclose(t[..., :3].astype(np.float32), 4.0 * (4.0 - 0.8) / 4.0, rtol=1e-3)
    d1 = oracle.bloom_downsample(t)
    assert d1.shape == (20, 32, 4) and np.allclose(d1[..., :3].astype(np.float32), 3.2, rtol=2e-3)
    # upsample of a constant is that constant; additive blend doubles it
    up = oracle.bloom_upsample_add(d1, t, 1.0)
    assert np.allclose(up[..., :3].astype(np.float32), 6.4, rtol=2e-3)
    # soft knee: brightness inside [th-knee, th+knee] gets the quadratic ramp
    mid = np.full((2, 2, 4), 0.8, np.float16)
    tm = oracle.bloom_threshold(mid, 0.8).astype(np.float32)
    br = np.float32(np.float16(0.8))
    soft = (br - np.float32(0.8) + np.float32(0.5)) ** 2 / (4 * 0.5 + 1e-5)
    assert np.allclose(tm[..., :3], br * soft / br, rtol=2e-3)
    # chain extents follow floor halving (PIN-9)
    U, D = oracle.bloom_chain(synth.random_hdr(67, 45, seed=3), 0.8, 5)
    assert [d.shape[:2] for d in D] == [(45, 67), (22, 33), (11, 16), (5, 8), (2, 4), (1, 2)]


def test_downsample_matches_numpy_for_even_extent(oracle):
    """For even sizes every tap of the 2:1 downsample sits midway between 4 texels (weights 1/2)."""
    W, H = 32, 20
    f = np.float32
    src = synth.random_hdr(W, H, seed=5)
    got = oracle.bloom_downsample(src)
    s = np.pad(src[..., :3].astype(f), ((2, 2), (2, 2), (0, 0)), mode="edge")

    def tap(ox, oy):     # tap centred between texels (2x+ox, 2x+ox+1) x (2y+oy, 2y+oy+1)
        a = s[2 + oy:2 + oy + H:2, 2 + ox:2 + ox + W:2]
        b = s[2 + oy:2 + oy + H:2, 3 + ox:3 + ox + W:2]
        c = s[3 + oy:3 + oy + H:2, 2 + ox:2 + ox + W:2]
        d = s[3 + oy:3 + oy + H:2, 3 + ox:3 + ox + W:2]
        top = a * f(0.5) + b * f(0.5)
        bot = c * f(0.5) + d * f(0.5)
        return top * f(0.5) + bot * f(0.5)

    order = [(0, 0, 0.25), (-1, -1, 0.0625), (1, -1, 0.0625), (-1, 1, 0.0625), (1, 1, 0.0625),
             (-1, 0, 0.125), (1, 0, 0.125), (0, -1, 0.125), (0, 1, 0.125)]
    res, sm = np.zeros((H // 2, W // 2, 3), f), np.zeros((H // 2, W // 2, 1), f)
    for ox, oy, wt in order:
        t = tap(ox, oy)
        luma = (t[..., 0:1] * f(0.2126) + t[..., 1:2] * f(0.7152)) + t[..., 2:3] * f(0.0722)
        w = f(1) / (f(1) + luma)
        res = res + (t * w) * f(wt)
        sm = sm + w * f(wt)
    ref = (res / sm).astype(np.float16)
    assert np.array_equal(got[..., :3], ref)


def test_tonemap_known_values(oracle):
    pc = np.zeros(1, T.POST_PC)
    pc["exposure"], pc["gamma"], pc["bloom_intensity"] = 1.0, 2.2, 0.5
    vals = np.array([0.0, 0.18, 1.0, 4.0, 60000.0], np.float16)
    hdr = np.zeros((1, len(vals), 4), np.float16)
    hdr[0, :, :3] = vals[:, None]
    out = oracle.tonemap(hdr, None, pc)
    x = vals.astype(np.float64)
    aces = np.clip((x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14), 0, 1)
    assert abs(aces[2] - 0.8038) < 1e-3
    ref = np.floor(aces ** (1 / 2.2) * 255 + 0.5)
    assert np.all(np.abs(out[0, :, 0].astype(np.int64) - ref) <= 1) and out[0, 0, 0] == 0
    assert np.all(out[..., 3] == 255)
    # bloom adds before exposure
    bloom = np.full_like(hdr, 2.0)
    out2 = oracle.tonemap(hdr, bloom, pc)
    x2 = x + 2.0 * 0.5
    ref2 = np.floor(np.clip((x2 * (2.51 * x2 + 0.03)) / (x2 * (2.43 * x2 + 0.59) + 0.14), 0, 1) ** (1 / 2.2) * 255 + 0.5)
    assert np.all(np.abs(out2[0, :, 1].astype(np.int64) - ref2) <= 1)


def test_tonemap_operators_match_the_published_closed_forms(oracle):
    """TonemapOperator::{Reinhard, Uncharted2, None} (features/tonemapping.rs:9-20) have no shader in the reference:
    PIN-13 pins the published closed forms; the oracle is checked against a float64 restatement of each."""
    pc = np.zeros(1, T.POST_PC)
    pc["exposure"], pc["gamma"], pc["bloom_intensity"] = 1.5, 2.2, 0.0
    vals = np.concatenate([np.linspace(0, 4, 200), [8.0, 11.2, 50.0, 60000.0]]).astype(np.float16)
    hdr = np.zeros((1, len(vals), 4), np.float16)
    hdr[0, :, :3] = vals[:, None]
    x = vals.astype(np.float64) * 1.5

    def hable(v):
        A, B, C, D, E, F = 0.15, 0.50, 0.10, 0.20, 0.02, 0.30
        return ((v * (A * v + C * B) + D * E) / (v * (A * v + B) + D * F)) - E / F
    curves = {oracle.TONEMAP_REINHARD: x / (1 + x), oracle.TONEMAP_UNCHARTED2: hable(2 * x) / hable(11.2), oracle.TONEMAP_NONE: x,
              oracle.TONEMAP_ACES: (x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14)}
    for op, curve in curves.items():
        out = oracle.tonemap(hdr, None, pc, op)
        ref = np.floor(np.clip(curve, 0, 1) ** (1 / 2.2) * 255 + 0.5)
        assert np.all(np.abs(out[0, :, 0].astype(np.int64) - ref) <= 1), op
        assert np.all(out[..., 3] == 255) and np.all(np.diff(out[0, :, 1].astype(int)) >= 0)     # monotone curves
    assert abs(hable(2 * 11.2 / 2) / hable(11.2) - 1.0) < 1e-12      # the white point maps to 1
    assert np.array_equal(oracle.tonemap(hdr, None, pc), oracle.tonemap(hdr, None, pc, oracle.TONEMAP_ACES))


# ------------------------------------------------------------------ committed regression vectors
def test_golden_vectors(oracle):
    """The committed vectors are outputs of the reference's own GLSL (oracle/_ref, tests/golden/make_golden.py)."""
    from tests.golden.make_golden import compute, META
    g = np.load(GOLDEN)
    assert "reference GLSL" in str(g["generator"])
    now = compute(oracle)
    assert sorted(k for k in g.files if k not in META) == sorted(now.keys())
    for k in now:
        assert np.array_equal(g[k], now[k]), k
