"""Row f-1 (depth / HDR producers) on the CPU: properties of the oracle's compute forward pass
(oracle/ash_oracle_forward.cpp) and an independent float64 restatement of frag.frag's BRDF."""
import math

import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T
from oracle import oracle

NO_OBJ = 0xFFFFFFFF


def _scene(W, H, eye=(0.0, 2.0, 22.0), target=(0, 0, 0), grid=(8, 6, 6), seed=3, near=0.5, **kw):
    cam = synth.Camera(eye, target, W, H, near=near)
    objs, draws = synth.make_objects(grid, spacing=2.0, seed=seed, jitter=0.3, disabled_frac=0.1)
    return cam, objs, draws, synth.forward_params(near=near, **kw)


@pytest.mark.parametrize("eye,target", [((0.0, 2.0, 22.0), (0, 0, 0)), ((3.0, 1.0, 9.0), (0, 0, 0)),
                                        ((0.3, 0.2, 0.1), (5, 1, 4)),            # inside the grid: spheres cross the near plane
                                        ((-14.0, 9.0, 3.0), (0, 0, 0)),
                                        ((1.1, -0.9, 2.7), (-6, 2, -3)), ((-2.9, 3.1, -1.2), (4, -3, 5)),       # more eyes inside the grid
                                        ((6.9, 0.1, 0.4), (-7, 0.5, 0.2)), ((0.0, 4.95, 0.0), (0.2, -5, 0.1))])
def test_silhouette_rectangle_is_conservative(eye, target):
    W, H = 160, 90
    cam, objs, _, fp = _scene(W, H, eye, target)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    keys_b, depth_b = oracle.forward_depth(objs, cam.camera_data(), fp, W, H, brute=True)
    assert np.array_equal(keys, keys_b) and np.array_equal(depth.view(np.uint32), depth_b.view(np.uint32))
    ids = (keys & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    assert (ids != NO_OBJ).any()
    assert np.array_equal(depth.view(np.uint32), (keys >> np.uint64(32)).astype(np.uint32))
    assert (depth[ids == NO_OBJ] == 1.0).all() and (depth >= 0).all() and (depth <= 1).all()
    assert not (objs["flags"][ids[ids != NO_OBJ]] & 1 == 0).any()          # disabled objects never appear


def test_depth_matches_analytic_sphere():
    """One sphere on the view axis: the centre pixel's depth is the projected near pole."""
    W, H = 65, 65
    cam = synth.Camera((0.0, 0.0, 10.0), (0, 0, 0), W, H, near=0.5, far=100.0)
    objs, _ = synth.make_objects((1, 1, 1), seed=1)
    objs["model"][0] = np.eye(4, dtype=np.float32) * [2.0, 2.0, 2.0, 1.0]      # scale 2 -> radius 1
    fp = synth.forward_params(near=0.5)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    t = 10.0 - 1.0
    P = cam.proj
    expect = (P[3][2] - P[2][2] * t) / t
    assert abs(float(depth[32, 32]) - expect) < 1e-6
    assert int(keys[32, 32]) & 0xFFFFFFFF == 0 and int(keys[0, 0]) & 0xFFFFFFFF == NO_OBJ
    # silhouette radius in pixels: tan(asin(r/d)) * P11 * H/2
    rad_px = math.tan(math.asin(1.0 / 10.0)) * abs(P[1][1]) * H / 2
    covered = ((keys & np.uint64(0xFFFFFFFF)) == 0).sum()
    assert abs(covered - math.pi * rad_px ** 2) < 0.15 * math.pi * rad_px ** 2


def test_forward_depth_feeds_the_occlusion_cull():
    """Ties row f-1 to the hot path.  One layer of separated objects: every sphere is fully visible and sits
    inside its own AABB, so occlusion_cull.comp (one Hi-Z texel at the centre of the projected box) must keep
    every object that reached the depth buffer.  With more layers behind, the same spheres occlude."""
    W, H = 256, 128
    cam = synth.Camera((0.0, 0.0, 20.0), (0, 0, 0), W, H, near=0.5)
    fp = synth.forward_params(near=0.5)
    objs, draws = synth.make_objects((6, 4, 1), spacing=3.0, seed=11)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    vis, _, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), 7), depth)
    ids = np.unique((keys & np.uint64(0xFFFFFFFF)).astype(np.uint32))
    ids = ids[ids != NO_OBJ]
    assert len(ids) == len(objs) and vis[ids].all()
    objs, draws = synth.make_objects((6, 4, 6), spacing=3.0, seed=11)
    objs["model"][:, 0, 0] = objs["model"][:, 1, 1] = objs["model"][:, 2, 2] = 2.4      # big spheres: the front layer hides the rest
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    vis, _, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), 7), depth)
    assert 24 <= cnt < len(objs)


def _brdf64(N, V, L, base, metallic, roughness, light):
    """frag.frag:89-114, 167-213 in float64 (independent restatement, numpy)."""
    N, V, L, base, light = (np.asarray(a, np.float64) for a in (N, V, L, base, light))
    NdotL = max(N @ L, 0.0)
    F0 = 0.04 * (1 - metallic) + base * metallic
    Hh = (V + L) / np.linalg.norm(V + L)
    NdotV, NdotH, VdotH = max(N @ V, 0.001), max(N @ Hh, 0.0), max(V @ Hh, 0.0)
    a2 = roughness ** 4
    den = NdotH * NdotH * (a2 - 1.0) + 1.0
    D = a2 / (3.14159265359 * den * den)
    k = (roughness + 1.0) ** 2 * 0.125
    G = NdotV / (NdotV * (1 - k) + k) * (NdotL / (NdotL * (1 - k) + k))
    t5 = min(max(1.0 - VdotH, 0.0), 1.0) ** 5
    F = F0 + (1.0 - F0) * t5
    spec = np.minimum(D * G * F / (4.0 * NdotV * NdotL + 0.001), 10.0 / np.maximum(0.04, F0))
    diffuse = (1.0 - F) * (1.0 - metallic) * base / 3.14159265359
    return (diffuse + spec) * light * NdotL


@pytest.mark.parametrize("metallic,roughness,reinhard", [(0.0, 0.5, False), (0.7, 0.2, False), (0.3, 0.9, True)])
def test_shading_matches_float64_brdf(metallic, roughness, reinhard):
    W, H = 33, 33
    cam = synth.Camera((0.0, 0.0, 6.0), (0, 0, 0), W, H, near=0.5)
    objs, _ = synth.make_objects((1, 1, 1), seed=1)
    objs["model"][0] = np.eye(4, dtype=np.float32) * [2.0, 2.0, 2.0, 1.0]
    lights = np.zeros(1, T.GPU_LIGHT)                                         # one point light up and to the right
    lights["position"][0] = [2.0, 2.0, 3.0, 8.0]
    lights["color"][0] = [1.0, 0.5, 0.25, 3.0]
    lights["params"][0] = [0, 0, 1, 1]
    sun_dir = np.array([-0.3, -1.0, -0.2])
    fp = synth.forward_params(light_direction=sun_dir, base_color=(0.8, 0.6, 0.4), metallic=metallic, roughness=roughness,
                              emissive=(0.01, 0.0, 0.02), flags=T.FWD_TILE_LIGHTS | (T.FWD_REINHARD if reinhard else 0))
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    td = np.zeros((tiles, 257), np.uint32)
    td[:, 0] = 1                                                              # every tile lists light 0
    hdr = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fp).astype(np.float64)
    base = np.array([0.8, 0.6, 0.4])
    view = cam.view.T                                                         # cam.view[c] = column c
    checked = 0
    for (x, y) in [(16, 16), (12, 14), (20, 19), (16, 10)]:
        assert int(keys[y, x]) & 0xFFFFFFFF == 0
        dx = ((x + 0.5) / W * 2 - 1) / cam.proj[0][0]
        dy = ((y + 0.5) / H * 2 - 1) / cam.proj[1][1]
        d = np.array([dx, dy, -1.0])
        c = (view @ np.array([0, 0, 0, 1.0]))[:3]
        b, a, cc = d @ c, d @ d, c @ c - 1.0
        t = (b - math.sqrt(b * b - a * cc)) / a
        p = d * t
        N = (p - c) / np.linalg.norm(p - c)
        V = -p / np.linalg.norm(p)
        sunL = -(view[:3, :3] @ sun_dir)
        sunL /= np.linalg.norm(sunL)
        color = np.array([0.03, 0.03, 0.04]) * base + _brdf64(N, V, sunL, base, metallic, max(roughness, 0.04), [1.0, 0.95, 0.9])
        color = color + np.array([0.01, 0.0, 0.02])
        lv = (view @ np.array([2.0, 2.0, 3.0, 1.0]))[:3]
        toL = lv - p
        dist = np.linalg.norm(toL)
        w = min(max(1.0 - (dist / 8.0) ** 4, 0.0), 1.0)
        rad = np.array([1.0, 0.5, 0.25]) * 3.0 * (w * w / (dist * dist + 1.0))
        color = color + _brdf64(N, V, toL / dist, base, metallic, max(roughness, 0.04), rad)
        if reinhard:
            color = color / (color + 1.0)
        assert np.allclose(hdr[y, x, :3], color, rtol=2e-3, atol=2e-4), (x, y, hdr[y, x, :3], color)
        assert hdr[y, x, 3] == 1.0
        checked += 1
    assert checked == 4
    bg = hdr[0, 0, :3]
    assert np.allclose(bg, [0.02, 0.02, 0.03], rtol=1e-3)                     # background: the clear colour, unshaded


def test_tile_lights_flag_and_instance_colour():
    W, H = 64, 48
    cam, objs, _, fp = _scene(W, H, grid=(4, 3, 3), seed=5)
    lights = synth.make_lights(16, extent=6.0, seed=2, center=(0.0, 0.0, 4.0), n_directional=1)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    lit = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fp).astype(np.float32)
    fp0 = fp.copy()
    fp0["flags"] = 0
    unlit = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fp0).astype(np.float32)
    hit = (keys & np.uint64(0xFFFFFFFF)) != np.uint64(NO_OBJ)
    assert (lit[hit][:, :3] >= unlit[hit][:, :3]).all() and (lit[hit][:, :3] > unlit[hit][:, :3]).any()
    assert np.array_equal(lit[~hit], unlit[~hit])
    inst = np.zeros(len(objs), oracle.INSTANCE_DTYPE)
    inst["color"] = [0.0, 1.0, 0.0, 1.0]                                       # green only
    fpc = fp0.copy()
    fpc["flags"] = T.FWD_INSTANCE_COLOR
    fpc["ambient_color"][0] = [1, 1, 1, 1]
    fpc["light_color"][0] = [0, 0, 0, 1]
    g = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fpc, instances=inst).astype(np.float32)
    assert (g[hit][:, 0] == 0).all() and (g[hit][:, 2] == 0).all() and (g[hit][:, 1] > 0.5).all()
