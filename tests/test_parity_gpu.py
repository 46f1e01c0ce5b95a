"""Parity of the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars: Hi-Z mips, tile min/max, visibility masks, visible counts, ordered draw
commands and per-tile light lists BIT-EXACT; RGBA16F images bit-exact (the kernels execute the
oracle's IEEE operation sequence) with a fall-back tolerance of 1 fp16 ulp stated where used;
8-bit output within 1 LSB (powf differs by ulps between glibc and CUDA).
"""
import os

import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_small.npz")


def _ctx(backend, W, H, **cfg):
    cfg.setdefault("max_objects", 131072)
    cfg.setdefault("max_lights", 8192)
    return backend.Context(W, H, **cfg)


def _taa_pc(W, H, blend=0.9, clamp=1.0, vel=1.0, jitter=(0.25, -0.16666667), prev=(0.0, 0.0)):
    pc = np.zeros(1, T.TAA_PC)
    w, h = np.float32(W), np.float32(H)
    pc["screen_params"][0] = [w, h, np.float32(1) / w, np.float32(1) / h]
    pc["params"][0] = [blend, clamp, vel, 0.0]
    pc["jitter"][0], pc["prev_jitter"][0] = jitter, prev
    return pc


def _post_pc(exposure=1.0, gamma=2.2, bloom_intensity=0.5):
    pc = np.zeros(1, T.POST_PC)
    pc["exposure"], pc["gamma"], pc["bloom_intensity"] = exposure, gamma, bloom_intensity
    return pc


def _bloom_params(threshold=0.8, mips=5, intensity=1.0):
    bp = np.zeros(1, T.BLOOM_PARAMS)
    bp["threshold"], bp["upsample_intensity"], bp["mip_count"] = threshold, intensity, mips
    return bp


def _assert_half_images_equal(got, ref, what):
    g, r = np.ascontiguousarray(got).view(np.uint16), np.ascontiguousarray(ref).view(np.uint16)
    if np.array_equal(g, r):
        return
    # report in fp16 ulps (sign-magnitude -> monotone integer)
    def mono(u):
        u = u.astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u)
    d = np.abs(mono(g) - mono(r))
    raise AssertionError(f"{what}: {int((d > 0).sum())} of {d.size} halfs differ, max {int(d.max())} fp16 ulp")


# ------------------------------------------------------------------ Hi-Z + tile min/max
@pytest.mark.parametrize("W,H", [(64, 64), (67, 45), (1, 1), (2, 1), (1, 7), (130, 70), (255, 257),
                                 (1920, 1080), (3840, 2160), (2048, 1024), (1000, 3)])
def test_hiz_pyramid_bit_exact(backend, oracle, W, H):
    depth = synth.random_depth(W, H, seed=W + H)
    mips, _ = oracle.hiz_build(depth)
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.hiz_build()
        assert c.hiz_level_count() == len(mips) + 1
        assert np.array_equal(c.hiz_mip(0), depth)
        for l, m in enumerate(mips, start=1):
            got = c.hiz_mip(l)
            assert got.shape == m.shape
            assert np.array_equal(got.view(np.uint32), m.view(np.uint32)), f"mip {l}"
        # tile min/max == light_culling.comp:130-143 on float bits
        lights = synth.make_lights(1, seed=1)
        cam = synth.Camera((0, 0, 5), (0, 0, 0), W, H)
        _, _, mm = oracle.light_cull(lights, cam.light_pc(1), cam.camera_data(), depth)
        got = c.readback(T.BUF_TILE_MINMAX, np.uint32).reshape(-1, 2)
        assert np.array_equal(got, mm)


def test_hiz_device_pointer_input_and_rebuild(backend, oracle):
    import torch
    W, H = 512, 256
    d0, d1 = synth.random_depth(W, H, seed=1), synth.random_depth(W, H, seed=2)
    t0, t1 = torch.from_numpy(d0).cuda(), torch.from_numpy(d1).cuda()
    torch.cuda.synchronize()
    with _ctx(backend, W, H, enable_post=0) as c:
        for t, d in ((t0, d0), (t1, d1), (t0, d0)):
            c.set_depth(t)                       # borrowed, zero-copy
            c.hiz_build()
            mips, _ = oracle.hiz_build(d)
            for l, m in enumerate(mips, start=1):
                assert np.array_equal(c.hiz_mip(l), m)


# ------------------------------------------------------------------ occlusion cull
def _cull_scene(W, H, grid, seed, eye=(0.0, 2.0, 22.0), disabled=0.05, spacing=2.0):
    cam = synth.Camera(eye, (0, 0, 0), W, H)
    objs, draws = synth.make_objects(grid, spacing=spacing, seed=seed, jitter=0.3, disabled_frac=disabled)
    depth = synth.splat_depth(objs, cam)
    return cam, objs, draws, depth


@pytest.mark.parametrize("W,H,grid,levels", [
    (320, 180, (6, 6, 6), 8), (200, 120, (7, 5, 3), 3), (1920, 1080, (10, 10, 10), 10),
    (3840, 2160, (25, 20, 20), 12), (3840, 2160, (25, 20, 20), 10), (1024, 1024, (40, 50, 50), 11),
    (67, 45, (3, 3, 3), 1)])
def test_occlusion_cull_bit_exact(backend, oracle, W, H, grid, levels):
    cam, objs, draws, depth = _cull_scene(W, H, grid, seed=W + grid[0])
    pc = cam.cull_pc(len(objs), levels)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    with _ctx(backend, W, H, enable_post=0, hiz_levels=levels) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.hiz_build()
        c.occlusion_cull(pc)
        got = c.visibility()
        assert np.array_equal(got, vis)
        assert c.visible_count() == cnt
        assert np.array_equal(c.draws_out(cnt), dout)                       # PIN-7 ascending order
        bits = c.readback(T.BUF_VISIBILITY_BITS, np.uint32)
        assert np.array_equal(np.unpackbits(bits.view(np.uint8), bitorder="little")[:len(vis)], vis.astype(np.uint8))
        st = c.stats()
        assert st["total_objects"] == len(objs) and st["visible_objects"] == cnt
    assert 0 < cnt < len(objs)


def test_occlusion_cull_edge_cases(backend, oracle):
    W, H = 128, 96
    cam = synth.Camera((0, 0, 10), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((1, 1, 1), seed=1)
    objs = np.repeat(objs, 70)                      # 70: not a multiple of the warp / CTA size
    objs["model"][1][3][:3] = [0, 0, 50]
    objs["model"][2][3][:3] = [0, 0, 9.9]           # clip.w <= 0 -> visible
    objs["flags"][3] = 0
    objs["flags"][4] = 2
    objs["extents"][5] = 0
    objs["model"][6][0][:3] = [0.6, 0.8, 0.0]       # rotated model: extents use column lengths only
    objs["model"][6][1][:3] = [-0.8, 0.6, 0.0]
    objs["extents"][7][:3] = [1e30, 1e30, 1e30]     # overflow to inf/NaN inside the projection
    objs["center"][8][:3] = [np.nan, 0, 0]
    for depth_val in (1.0, 0.0, 0.97):
        depth = np.full((H, W), depth_val, np.float32)
        pc = cam.cull_pc(len(objs), 8)
        vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
        with _ctx(backend, W, H, enable_post=0) as c:
            c.set_depth(depth)
            c.set_objects(objs, draws)
            c.occlusion_cull(pc)                    # builds the pyramid on demand
            assert np.array_equal(c.visibility(), vis)
            assert c.visible_count() == cnt
            assert np.array_equal(c.draws_out(cnt), dout)
            # objectCount smaller than the uploaded array; then zero objects
            pc2 = cam.cull_pc(33, 8)
            vis2, dout2, cnt2 = oracle.occlusion_cull(objs, draws, pc2, depth)
            c.occlusion_cull(pc2)
            assert c.visible_count() == cnt2 and np.array_equal(c.visibility()[:33], vis2)
            c.occlusion_cull(cam.cull_pc(0, 8))
            assert c.visible_count() == 0


# ------------------------------------------------------------------ row f-2: LOD selection + instance batching
def _instancing_scene(W, H, grid, seed, n_batches):
    cam, objs, draws, depth = _cull_scene(W, H, grid, seed=seed)
    meshes = synth.make_lod_meshes(6, seed=seed + 1)
    inst, batch, mesh = synth.make_instances(objs, n_batches, 6, seed=seed + 2)
    return cam, objs, draws, depth, meshes, inst, batch, mesh


def _compare_instancing(c, ref, n, n_batches):
    sel, out, offsets, draws, stats = ref
    got_sel = c.lod_selection()[:n]
    assert np.array_equal(got_sel.view(np.uint32), sel.view(np.uint32)), "LodSelection (level, blend bits, culled, coverage bits)"
    nb = n_batches * 8
    got_off = c.bucket_offsets()
    assert np.array_equal(got_off[:nb + 1], offsets)
    assert np.array_equal(c.instance_draws()[:nb], draws)
    got_out = c.instances_out(int(offsets[-1]))
    assert got_out.tobytes() == out.tobytes(), "instance streams (bucket order, ascending object index inside)"
    gs = c.instancing_stats()
    for k in ("objects_per_lod", "objects_culled", "total_instances", "batch_count", "triangles_before",
              "triangles_after", "instances_dropped"):
        assert np.array_equal(gs[k], stats[k]), k
    assert np.float32(gs["avg_instances_per_batch"]).view(np.uint32) == np.float32(stats["avg_instances_per_batch"]).view(np.uint32)


@pytest.mark.parametrize("mode,use_vis,grid", [
    ("ScreenSize", True, (12, 10, 9)), ("ScreenSize", False, (40, 50, 50)), ("Distance", True, (25, 20, 20)),
    (("Force", 1), True, (12, 10, 9)), (("Force", 7), False, (7, 5, 3)), ("ScreenSize", True, (1, 1, 1))])
def test_instancing_bit_exact(backend, oracle, mode, use_vis, grid):
    from ash_renderer_b200 import host
    W, H, n_batches = 1920, 1080, 9
    cam, objs, draws, depth, meshes, inst, batch, mesh = _instancing_scene(W, H, grid, seed=sum(grid), n_batches=n_batches)
    n = len(objs)
    cfg = host.LodConfig(mode=mode, global_bias=0.3 if mode == "Distance" else 0.0).record(use_visibility=use_vis)
    pc = cam.cull_pc(n, 10)
    vis, _, _ = oracle.occlusion_cull(objs, draws, pc, depth)
    ip = synth.instancing_params(cam, n)
    ref = oracle.instancing(inst, batch, mesh, meshes, ip, cfg, n_batches, vis if use_vis else None)
    with _ctx(backend, W, H, enable_post=0, max_batches=n_batches) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.set_instances(inst, batch, mesh)
        c.set_lod_meshes(meshes)
        c.set_lod_config(cfg)
        c.hiz_build()
        c.occlusion_cull(pc)
        c.instancing_finalize(ip)
        _compare_instancing(c, ref, n, n_batches)
        c.instancing_finalize(ip)                        # idempotent: stats and streams are rebuilt, not accumulated
        _compare_instancing(c, ref, n, n_batches)
    if n > 100:
        assert ref[4]["batch_count"] > 1 and 0 < ref[4]["total_instances"] <= n


def test_instancing_batch_cap_and_errors(backend, oracle):
    from ash_renderer_b200 import host
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), 640, 360)
    n = 70000
    inst = np.zeros(n, T.INSTANCE_DATA)
    inst["model_row3"][:, 3] = 1.0
    inst["custom"][:, 0] = np.arange(n)
    zero = np.zeros(n, np.uint32)
    meshes = synth.make_lod_meshes(1, seed=1)
    cfg = host.LodConfig(cull_threshold=0.0).record(use_visibility=False)
    ip = synth.instancing_params(cam, n)
    ref = oracle.instancing(inst, zero, zero, meshes, ip, cfg, 2)
    assert ref[4]["total_instances"] == 65536 and ref[4]["instances_dropped"] == n - 65536      # instancing.rs:219-223
    with _ctx(backend, 640, 360, enable_post=0, max_batches=2) as c:
        with pytest.raises(backend.AshError) as e:
            c.instancing_finalize(ip)
        assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED            # FeatureNotInitialized: nothing set yet
        c.set_instances(inst, zero, zero)
        with pytest.raises(backend.AshError):
            c.instancing_finalize(ip)                                     # no LOD meshes
        c.set_lod_meshes(meshes)
        c.set_lod_config(cfg)
        c.instancing_finalize(ip)
        _compare_instancing(c, ref, n, 2)
        with pytest.raises(backend.AshError):
            c.set_instances(inst, zero + 2, zero)                         # batch id >= max_batches


def test_view_run_with_instancing_stage(backend, oracle):
    from ash_renderer_b200 import host
    W, H, n_batches = 1280, 720, 4
    cam, objs, draws, depth, meshes, inst, batch, mesh = _instancing_scene(W, H, (14, 12, 10), seed=21, n_batches=n_batches)
    n = len(objs)
    pc = cam.cull_pc(n, 10)
    vis, _, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    cfg = host.LodConfig().record(use_visibility=True)
    ref = oracle.instancing(inst, batch, mesh, meshes, synth.instancing_params(cam, n), cfg, n_batches, vis)
    vp = np.zeros(1, T.VIEW_PARAMS)
    vp["cull"], vp["camera"], vp["light"] = pc, cam.camera_data(), cam.light_pc(0)
    vp["stages"] = T.STAGE_HIZ | T.STAGE_CULL | T.STAGE_INSTANCING
    with _ctx(backend, W, H, enable_post=0, max_batches=n_batches) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.set_instances(inst, batch, mesh)
        c.set_lod_meshes(meshes)
        c.set_lod_config(cfg)
        c.view_run(vp)
        assert c.visible_count() == cnt
        _compare_instancing(c, ref, n, n_batches)
        assert c.instancing_stats()["objects_culled"] + int(c.instancing_stats()["objects_per_lod"].sum()) == cnt


# ------------------------------------------------------------------ light cull
def _compare_tile_lists(got, ref):
    assert got.shape == ref.shape
    assert np.array_equal(got[:, 0], ref[:, 0]), "per-tile counts"
    cnt = ref[:, 0].astype(np.int64)
    mask = np.arange(256)[None, :] < cnt[:, None]
    assert np.array_equal(got[:, 1:][mask], ref[:, 1:][mask]), "per-tile index lists"   # PIN-8: valid prefix only


@pytest.mark.parametrize("W,H,n_lights,extent,eye", [
    (200, 120, 300, 10.0, (0.0, 2.0, 22.0)), (1920, 1080, 256, 10.0, (0.0, 2.0, 22.0)),
    (3840, 2160, 4096, 30.0, (0.0, 2.0, 40.0)), (3840, 2160, 8192, 30.0, (5.0, 3.0, 12.0)),
    (67, 45, 33, 5.0, (0.0, 0.0, 6.0)), (1000, 3, 100, 5.0, (0.0, 0.0, 6.0))])
def test_light_cull_bit_exact(backend, oracle, W, H, n_lights, extent, eye):
    cam = synth.Camera(eye, (0, 0, 0), W, H)
    objs, _ = synth.make_objects((6, 6, 6), seed=3)
    depth = synth.splat_depth(objs, cam)
    # centre half the lights on the camera so that the (raw-depth) near/far planes keep many
    lights = synth.make_lights(n_lights, extent=extent, seed=W + n_lights, center=(0, 0, 0))
    half = n_lights // 2
    lights["position"][:half, :3] = (lights["position"][:half, :3] * np.float32(0.2) + np.asarray(eye, np.float32))
    pc, cd = cam.light_pc(n_lights), cam.camera_data()
    ref, true_counts, _ = oracle.light_cull(lights, pc, cd, depth)
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.upload_lights(lights)
        c.upload_camera(cd)
        c.hiz_build()
        c.light_cull(pc)
        _compare_tile_lists(c.tile_data(), ref)
        st = c.stats()
        assert st["tiles"] == len(ref)
        assert st["sum_tile_lights"] == int(ref[:, 0].sum())
        assert st["max_tile_lights"] == int(ref[:, 0].max())
        assert st["overflow_tiles"] == int((true_counts > 256).sum())
    assert ref[:, 0].max() > 0


@pytest.mark.parametrize("case", ["near_tiny", "ortho", "random_matrix", "offcentre", "weird_lights", "flat_depth"])
def test_light_cull_prefilter_is_conservative(backend, oracle, case):
    """The per-cell prefilter (light_cull.cu k_light_cells) must never drop a light the shader keeps:
    cameras whose tile planes degenerate (shader zeroes the normal, :92-95), non-perspective and
    arbitrary inverse projections, non-finite / negative-radius lights, constant depth."""
    W, H = 1000, 600
    rng = np.random.default_rng(11)
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H, near=0.002 if case == "near_tiny" else 0.5)
    objs, _ = synth.make_objects((6, 6, 6), seed=3)
    depth = synth.splat_depth(objs, cam)
    n_lights = 1500
    lights = synth.make_lights(n_lights, extent=12.0, seed=5, center=(0, 0, 0))
    cd = cam.camera_data()
    # raw depth in [0,1] bounds the z slab (SURVEY PIN-2), so lights that the SIDE planes can separate
    # sit within ~1 unit of the eye in view z and are spread laterally
    n = 1200
    vp = np.stack([rng.uniform(-1.2, 1.2, n), rng.uniform(-0.8, 0.8, n), rng.uniform(-1.6, 0.3, n), np.ones(n)], 1)
    wp = (np.linalg.inv(cd["view"][0].astype(np.float64).T) @ vp.T).T
    lights["position"][:n, :3] = wp[:, :3].astype(np.float32)
    lights["position"][:n, 3] = rng.uniform(0.01, 0.25, n).astype(np.float32)
    if case == "ortho":
        ip = np.diag([8.0, 4.5, -50.0, 1.0]).astype(np.float32)
        ip[3, 2] = -0.5                                    # column-major rows: translation in z
        cd["inv_projection"][0] = ip
    elif case == "random_matrix":
        cd["inv_projection"][0] = rng.normal(size=(4, 4)).astype(np.float32)
    elif case == "offcentre":
        ip = cd["inv_projection"][0].copy()
        ip[3, 0] += np.float32(0.37); ip[3, 1] -= np.float32(0.21)     # sheared frustum
        cd["inv_projection"][0] = ip
    elif case == "weird_lights":
        p = lights["position"]
        p[0, 0] = np.nan; p[1, 3] = np.nan; p[2, 1] = np.inf; p[3, 3] = np.inf; p[4, 3] = -3.0
        p[5, :3] = 1e30; p[6, 3] = 0.0; p[7, 3] = -np.inf; p[8, 2] = -np.inf
    elif case == "flat_depth":
        depth = np.full((H, W), 0.9995, np.float32)
    pc = cam.light_pc(n_lights)
    ref, true_counts, _ = oracle.light_cull(lights, pc, cd, depth)
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.upload_lights(lights)
        c.upload_camera(cd)
        c.hiz_build()
        c.light_cull(pc)
        _compare_tile_lists(c.tile_data(), ref)
        assert c.stats()["overflow_tiles"] == int((true_counts > 256).sum())
    if case in ("offcentre", "weird_lights", "flat_depth"):
        assert ref[:, 0].min() < ref[:, 0].max()        # the side planes do separate tiles here


def test_light_cull_dense_scene_at_4k_with_overflow_tiles(backend, oracle):
    """The per-tile machinery at scale (the bench scene keeps ~5 lights per tile: most of its 8192 lights fall outside the
    raw-depth near / far planes, light_culling.comp:96-97): 8000 of 8192 lights inside the z slab of a 4K view, ~228
    per tile, about a thousand tiles past MAX_LIGHTS_PER_TILE = 256 (light_culling.comp:175-178) — lists, order and
    the overflow count bit-exact."""
    W, H = 3840, 2160
    cam, lights, cd, depth = synth.dense_light_scene(W, H)
    pc = cam.light_pc(len(lights))
    ref, true_counts, _ = oracle.light_cull(lights, pc, cd, depth)
    assert ref[:, 0].mean() > 150 and (true_counts > 256).sum() > 100 and ref[:, 0].max() == 256
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.upload_lights(lights)
        c.upload_camera(cd)
        c.hiz_build()
        for _ in range(2):                                # the second call runs with the tile planes cached
            c.light_cull(pc)
            _compare_tile_lists(c.tile_data(), ref)
            st = c.stats()
            assert st["sum_tile_lights"] == int(ref[:, 0].sum())
            assert st["overflow_tiles"] == int((true_counts > 256).sum())


@pytest.mark.parametrize("n_lights", [1, 1023, 1025, 3072, 20000, 65537])
def test_light_prepass_orders_candidates_across_ctas(backend, oracle, n_lights):
    """k_light_prepass is one CTA per 1024 lights with a look-back over the published counts of the CTAs before it: the
    candidates must come out in ascending light index whatever the CTA count (1 .. 65) and however full the last CTA is —
    the tile lists (PIN-7 order), the candidate count and lights_culled say so.  Three calls in a row: the tickets reset."""
    W, H = 320, 180
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, _ = synth.make_objects((6, 6, 6), seed=3)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(n_lights, extent=12.0, seed=n_lights, center=(0, 0, 0), n_directional=min(4, n_lights))
    half = n_lights // 2                                   # half the lights near the eye: inside the raw-depth slab
    lights["position"][:half, :3] = (lights["position"][:half, :3] * np.float32(0.05) + np.asarray((0.0, 2.0, 22.0), np.float32))
    lights["position"][:half, 3] *= np.float32(0.1)
    pc, cd = cam.light_pc(n_lights), cam.camera_data()
    ref, true_counts, _ = oracle.light_cull(lights, pc, cd, depth)
    with _ctx(backend, W, H, enable_post=0, max_lights=max(n_lights, 16)) as c:
        c.set_depth(depth)
        c.upload_lights(lights)
        c.upload_camera(cd)
        c.hiz_build()
        for _ in range(3):
            c.light_cull(pc)
            _compare_tile_lists(c.tile_data(), ref)
            st = c.stats()
            assert st["sum_tile_lights"] == int(ref[:, 0].sum())
            assert st["overflow_tiles"] == int((true_counts > 256).sum())
            assert 0 < int(st["candidate_lights"]) <= n_lights


def test_light_cull_overflow_zero_lights_and_disabled(backend, oracle):
    W, H = 64, 48
    cam = synth.Camera((0, 0, 5), (0, 0, 0), W, H)
    depth = synth.random_depth(W, H, seed=4)
    lights = synth.make_lights(700, extent=0.2, rmin=50, rmax=60, seed=9, n_directional=3, enabled_p=0.9,
                               center=(0, 0, 5))
    cd = cam.camera_data()
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.upload_camera(cd)
        c.upload_lights(lights)
        for n in (700, 256, 257, 1, 0):
            pc = cam.light_pc(n)
            ref, true_counts, _ = oracle.light_cull(lights[:n], pc, cd, depth)
            c.light_cull(pc)
            _compare_tile_lists(c.tile_data(), ref)
            assert c.stats()["overflow_tiles"] == int((true_counts > 256).sum())
        allowed = lights.copy()
        allowed["params"][:, 3] = 0.0                # all disabled
        c.upload_lights(allowed)
        c.light_cull(cam.light_pc(700))
        assert np.all(c.tile_data()[:, 0] == 0)


# ------------------------------------------------------------------ TAA
@pytest.mark.parametrize("W,H,with_vel", [(200, 120, True), (200, 120, False), (67, 45, True),
                                          (1920, 1080, False), (33, 9, False), (1920, 1080, True), (33, 9, True)])
def test_taa_bit_exact(backend, oracle, W, H, with_vel):
    cur, hist = synth.random_hdr(W, H, seed=1), synth.random_hdr(W, H, seed=2)
    vel = synth.random_velocity(W, H, seed=3) if with_vel else None
    with _ctx(backend, W, H) as c:
        for kw in (dict(), dict(clamp=0.0), dict(vel=0.0, jitter=(-0.4, 0.3), prev=(0.45, -0.45)), dict(blend=0.0)):
            pc = _taa_pc(W, H, **kw)
            # frame 0: history = current colour (PIN-12); frame 1: history = previous resolve
            c.taa_reset_history()
            c.set_color(hist, None)
            c.taa_resolve(_taa_pc(W, H, blend=1.0, clamp=0.0, vel=0.0, jitter=(0, 0)))   # out = hist exactly
            assert np.array_equal(c.taa_out()[..., :3], hist[..., :3])
            ref_hist = c.taa_out()
            c.set_color(cur, vel)
            c.taa_resolve(pc)
            ref = oracle.taa_resolve(cur, ref_hist, vel, pc)
            _assert_half_images_equal(c.taa_out(), ref, f"taa {kw}")


# BASELINE config 4 at a size the oracle follows in seconds: a 64-frame jittered sequence (four periods of the 16-entry
# Halton table of temporal_aa.rs:42-58), every resolve compared, the history recurrence fed by the oracle's own output.
@pytest.mark.parametrize("W,H,frames", [(160, 96, 18), (416, 240, 64)])
def test_taa_sequence_halton_16_frames(backend, oracle, W, H, frames):
    from ash_renderer_b200.host import TemporalAA
    seq = synth.HdrSequence(W, H, seed=4, n_discs=20)
    taa = TemporalAA()
    hist = None
    with _ctx(backend, W, H) as c:
        for t in range(frames):
            taa.begin_frame()
            j = taa.current_jitter()
            frame = seq.frame(float(j[0]), float(j[1]))
            pc = taa.push_constants(W, H)
            c.set_color(frame, None)
            c.taa_resolve(pc)
            ref = oracle.taa_resolve(frame, frame if hist is None else hist, None, pc)
            _assert_half_images_equal(c.taa_out(), ref, f"frame {t}")
            hist = ref


# ------------------------------------------------------------------ bloom + tonemap
# (1984, 1096): level 1 = 992 x 548 — box downsample with its right lane past the edge (992 % 64 == 32) and a partial last
# row of tiles (548 % 8 == 4); pair upsample whose second lane runs out of rows in the last row of tiles (548 % 16 == 4)
@pytest.mark.parametrize("W,H,mips", [(200, 120, 5), (67, 45, 5), (1920, 1080, 5), (64, 64, 3), (33, 9, 2), (256, 256, 8), (1984, 1096, 3)])
def test_bloom_chain_bit_exact(backend, oracle, W, H, mips):
    hdr = synth.random_hdr(W, H, seed=W)
    U, D = oracle.bloom_chain(hdr, 0.8, mips)
    with _ctx(backend, W, H, bloom_mips=mips) as c:
        c.set_color(hdr, None)
        c.bloom(_bloom_params(0.8, mips))
        for k in range(mips + 1):
            _assert_half_images_equal(c.bloom_mip(k), U[k], f"U_{k}")


@pytest.mark.parametrize("W,H", [(200, 120), (1920, 1080), (67, 45)])
def test_tonemap_within_one_lsb(backend, oracle, W, H):
    hdr = synth.random_hdr(W, H, seed=5)
    pc = _post_pc()
    with _ctx(backend, W, H) as c:
        c.set_color(hdr, None)
        c.tonemap(pc)                                   # no bloom computed: bloom term absent
        ref = oracle.tonemap(hdr, None, pc)
        d = np.abs(c.ldr8().astype(np.int16) - ref.astype(np.int16))
        assert d.max() <= 1 and (d > 0).mean() < 0.01
        c.bloom(_bloom_params())
        c.tonemap(pc)
        U, _ = oracle.bloom_chain(hdr, 0.8, 5)
        ref = oracle.tonemap(hdr, U[0], pc)
        d = np.abs(c.ldr8().astype(np.int16) - ref.astype(np.int16))
        assert d.max() <= 1 and (d > 0).mean() < 0.01
        assert np.all(c.ldr8()[..., 3] == 255)


@pytest.mark.parametrize("op", [T.TONEMAP_ACES, T.TONEMAP_REINHARD, T.TONEMAP_UNCHARTED2, T.TONEMAP_NONE])
def test_tonemap_operators_within_one_lsb(backend, oracle, op):
    """TonemappingConfig.operator (features/tonemapping.rs:9-20, row f-4): staged call and whole view, <= 1 LSB."""
    W, H = 320, 180
    hdr = synth.random_hdr(W, H, seed=15)
    pc = _post_pc(exposure=1.3)
    U, _ = oracle.bloom_chain(hdr, 0.8, 5)
    with _ctx(backend, W, H) as c:
        c.set_tonemap_operator(op)
        c.set_color(hdr, None)
        c.tonemap(pc)
        d = np.abs(c.ldr8().astype(np.int16) - oracle.tonemap(hdr, None, pc, op).astype(np.int16))
        assert d.max() <= 1 and (d > 0).mean() < 0.01
        cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
        vp = _view_params(cam, 0, 0, 1, W, H, _taa_pc(W, H), stages=T.STAGE_BLOOM | T.STAGE_TONEMAP)
        vp["post"] = pc
        for _ in range(2):                                # second call replays the captured graph
            c.view_run(vp)
            d = np.abs(c.ldr8().astype(np.int16) - oracle.tonemap(hdr, U[0], pc, op).astype(np.int16))
            assert d.max() <= 1 and (d > 0).mean() < 0.01
        with pytest.raises(backend.AshError):
            c.set_tonemap_operator(7)


# ------------------------------------------------------------------ whole view
def _view_params(cam, n_objs, n_lights, hiz_levels, W, H, taa_pc, stages=T.STAGE_ALL, view_id=7):
    vp = np.zeros(1, T.VIEW_PARAMS)
    vp["cull"] = cam.cull_pc(n_objs, hiz_levels)
    vp["light"] = cam.light_pc(n_lights)
    vp["camera"] = cam.camera_data()
    vp["taa"] = taa_pc
    vp["bloom"] = _bloom_params()
    vp["post"] = _post_pc()
    vp["view_id"], vp["stages"] = view_id, stages
    return vp


def _oracle_view(oracle, cam, objs, draws, lights, depth, frames, taa_pcs, hiz_levels):
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), hiz_levels), depth)
    td, true_counts, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    hist = None
    for f, pc in zip(frames, taa_pcs):
        hist = oracle.taa_resolve(f, f if hist is None else hist, None, pc)
    U, _ = oracle.bloom_chain(hist, 0.8, 5)
    ldr = oracle.tonemap(hist, U[0], _post_pc())
    return vis, dout, cnt, td, true_counts, hist, U[0], ldr


@pytest.mark.parametrize("W,H,keep", [(320, 180, True), (1920, 1080, True), (1920, 1080, False), (200, 120, False),
                                      (67, 45, True)])
def test_view_run_matches_oracle(backend, oracle, W, H, keep):
    from ash_renderer_b200.host import TemporalAA
    cam, objs, draws, depth = _cull_scene(W, H, (8, 8, 8), seed=5)
    lights = synth.make_lights(256, extent=10.0, seed=5, center=(0.0, 2.0, 22.0))
    seq = synth.HdrSequence(W, H, seed=6, n_discs=30)
    taa = TemporalAA()
    frames, pcs = [], []
    for _ in range(3):
        taa.begin_frame()
        j = taa.current_jitter()
        frames.append(seq.frame(float(j[0]), float(j[1])))
        pcs.append(taa.push_constants(W, H))
    levels = min(10, len(oracle.mip_extents(W, H)))
    vis, dout, cnt, td, true_counts, taa_ref, bloom_ref, ldr_ref = _oracle_view(
        oracle, cam, objs, draws, lights, depth, frames, pcs, levels)
    with _ctx(backend, W, H) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_depth(depth)
        for f, pc in zip(frames, pcs):
            c.set_color(f, None)
            c.view_run(_view_params(cam, len(objs), len(lights), levels, W, H, pc,
                                    stages=T.STAGE_ALL | (T.STAGE_KEEP_BLOOM if keep else 0)))
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt
        assert np.array_equal(c.draws_out(cnt), dout)
        _compare_tile_lists(c.tile_data(), td)
        _assert_half_images_equal(c.taa_out(), taa_ref, "taa")
        if keep:
            _assert_half_images_equal(c.bloom_mip(0), bloom_ref, "bloom U_0")
        d = np.abs(c.ldr8().astype(np.int16) - ldr_ref.astype(np.int16))
        assert d.max() <= 1
        st = c.stats()
        assert st["view_id"] == 7 and st["visible_objects"] == cnt and st["sum_tile_lights"] == int(td[:, 0].sum())
        assert c.kernel_launches() > 0


def test_view_run_post_chain_with_velocity_buffer(backend, oracle):
    """Fused TAA + threshold with a per-pixel velocity buffer (fast kernel, history gathered per pixel)."""
    from ash_renderer_b200.host import TemporalAA
    W, H = 640, 360
    seq = synth.HdrSequence(W, H, seed=9, n_discs=30)
    vel = synth.random_velocity(W, H, seed=10, still_frac=0.3, max_px=6.0)
    taa = TemporalAA()
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    hist, ref_bloom = None, None
    with _ctx(backend, W, H) as c:
        for t in range(3):
            taa.begin_frame()
            j = taa.current_jitter()
            f = seq.frame(float(j[0]), float(j[1]))
            pc = taa.push_constants(W, H)
            ref = oracle.taa_resolve(f, f if hist is None else hist, vel, pc)
            U, _ = oracle.bloom_chain(ref, 0.8, 5)
            c.set_color(f, vel)
            c.view_run(_view_params(cam, 0, 0, 1, W, H, pc, stages=T.STAGE_TAA | T.STAGE_BLOOM | T.STAGE_TONEMAP | T.STAGE_KEEP_BLOOM))
            _assert_half_images_equal(c.taa_out(), ref, f"taa+velocity frame {t}")
            _assert_half_images_equal(c.bloom_mip(0), U[0], f"bloom frame {t}")
            hist = ref


# ------------------------------------------------------------------ row f-3: two-phase culling + CullingStats
@pytest.mark.parametrize("graph", [0, 1])
def test_two_phase_culling_matches_oracle_procedure(backend, oracle, graph):
    """Phase 1 against the previous view's pyramid, draw, Hi-Z, phase 2 for the rest, draw the late objects:
    visibility, ordered draws, keys, final depth and CullingStats against the same procedure built from the
    single-phase oracle pieces, over a camera path (first view: no previous pyramid)."""
    W, H, levels = 512, 256, 8
    objs, draws = synth.make_objects((12, 8, 10), spacing=2.0, seed=61, jitter=0.3, disabled_frac=0.05)
    objs["model"][:, 0, 0] *= 1.5; objs["model"][:, 1, 1] *= 1.5; objs["model"][:, 2, 2] *= 1.5      # bigger spheres: real occlusion
    lights = synth.make_lights(64, extent=10.0, seed=62, center=(0.0, 2.0, 4.0))
    fp = synth.forward_params(near=0.5, flags=0)
    stages = T.STAGE_HIZ | T.STAGE_CULL | T.STAGE_LIGHTS | T.STAGE_FORWARD_DEPTH | T.STAGE_TWO_PHASE
    prev_depth = None
    late_total = 0
    with _ctx(backend, W, H, enable_post=0, graph_replay=graph) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_forward_params(fp)
        for t in range(6):
            cam = synth.Camera((2.5 * t - 6.0, 1.0 + 0.5 * t, 26.0 - 1.5 * t), (0.5 * t, 0, 0), W, H, near=0.5)
            pc = cam.cull_pc(len(objs), levels)
            ref = oracle.two_phase_cull(objs, draws, pc, cam.camera_data(), fp, W, H, prev_depth)
            vp = _view_params(cam, len(objs), len(lights), levels, W, H, _taa_pc(W, H), stages=stages, view_id=t)
            c.view_run(vp)
            assert np.array_equal(c.visibility(), ref["visibility"]), f"view {t}: visibility"
            assert c.visible_count() == ref["count"]
            assert np.array_equal(c.draws_out(ref["count"]), ref["draws_out"]), f"view {t}: draws"
            assert np.array_equal(c.forward_keys(), ref["keys"]), f"view {t}: keys"
            assert np.array_equal(c.depth().view(np.uint32), ref["depth"].view(np.uint32)), f"view {t}: depth"
            cs = c.culling_stats()
            _, _, after_frustum = oracle.occlusion_cull(objs, draws, pc, np.ones((H, W), np.float32))
            assert int(cs["objects_submitted"]) == len(objs) and int(cs["after_frustum"]) == after_frustum
            assert int(cs["early_visible"]) == ref["early_count"] and int(cs["late_visible"]) == ref["late_count"]
            assert int(cs["after_occlusion"]) == ref["count"] == ref["early_count"] + ref["late_count"]
            assert int(cs["lights_culled"]) == len(lights) - int(c.stats()["candidate_lights"])
            if t == 0:
                assert ref["early_count"] == after_frustum and ref["late_count"] == 0      # first view: frustum only
            late_total += ref["late_count"]
            # the light cull of the same view saw the final depth
            td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), ref["depth"])
            _compare_tile_lists(c.tile_data(), td)
            prev_depth = ref["depth"]
        assert late_total > 0, "the camera path never produced a late object: the test scene is too easy"
        cap, rep = c.graph_stats()
        assert ((cap == 2 and rep == 4) if graph else (cap == 0 and rep == 0)), (cap, rep)


def test_culling_stats_single_phase_with_instancing(backend, oracle):
    from ash_renderer_b200 import host
    W, H, n_batches = 640, 360, 4
    cam, objs, draws, depth, meshes, inst, batch, mesh = _instancing_scene(W, H, (12, 10, 8), seed=71, n_batches=n_batches)
    lights = synth.make_lights(200, extent=10.0, seed=72, center=(0.0, 2.0, 22.0))
    n = len(objs)
    pc = cam.cull_pc(n, 9)
    _, _, after_frustum = oracle.occlusion_cull(objs, draws, pc, np.ones((H, W), np.float32))
    _, _, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    vp = _view_params(cam, n, len(lights), 9, W, H, _taa_pc(W, H), stages=T.STAGE_HIZ | T.STAGE_CULL | T.STAGE_LIGHTS | T.STAGE_INSTANCING)
    with _ctx(backend, W, H, enable_post=0, max_batches=n_batches) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_instances(inst, batch, mesh)
        c.set_lod_meshes(meshes)
        c.set_lod_config(host.LodConfig().record(use_visibility=True))
        c.view_run(vp)
        cs, ist = c.culling_stats(), c.instancing_stats()
        assert (int(cs["objects_submitted"]), int(cs["after_frustum"]), int(cs["after_occlusion"])) == (n, after_frustum, cnt)
        assert int(cs["lights_culled"]) == len(lights) - int(c.stats()["candidate_lights"])
        assert int(cs["lod_triangles_saved"]) == int(ist["triangles_before"]) - int(ist["triangles_after"]) > 0
        assert int(cs["early_visible"]) == 0 and int(cs["late_visible"]) == 0
        rec = host.StatsCollector().record_device(cs)            # StatsCollector::record_culling / record_lod_savings
        assert (rec.objects_submitted, rec.after_frustum, rec.after_occlusion) == (n, after_frustum, cnt) and rec.lod_triangles_saved > 0
        # the staged call fills the cull's own fields too
        c.occlusion_cull(pc)
        cs = c.culling_stats()
        assert (int(cs["objects_submitted"]), int(cs["after_frustum"]), int(cs["after_occlusion"])) == (n, after_frustum, cnt)


# ------------------------------------------------------------------ row f-1: depth / HDR producers
def _forward_scene(W, H, eye, target, grid, seed, n_lights=200, **fpkw):
    cam = synth.Camera(eye, target, W, H, near=0.5)
    objs, draws = synth.make_objects(grid, spacing=2.0, seed=seed, jitter=0.3, disabled_frac=0.1)
    lights = synth.make_lights(n_lights, extent=10.0, seed=seed + 1, center=(0.0, 2.0, 4.0))
    return cam, objs, draws, lights, synth.forward_params(near=0.5, **fpkw)


@pytest.mark.parametrize("W,H,eye,target,grid", [
    (640, 360, (0.0, 2.0, 22.0), (0, 0, 0), (10, 8, 8)),
    (333, 187, (3.0, 1.0, 9.0), (0, 0, 0), (8, 6, 6)),            # odd extent, close: big rectangles
    (320, 180, (0.3, 0.2, 0.1), (5, 1, 4), (8, 6, 6)),            # eye inside the grid: near-plane crossings, full-screen rectangles
    (1920, 1080, (-14.0, 9.0, 30.0), (0, 0, 0), (24, 16, 16)),
])
def test_forward_depth_bit_exact(backend, oracle, W, H, eye, target, grid):
    cam, objs, draws, lights, fp = _forward_scene(W, H, eye, target, grid, seed=41)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    with _ctx(backend, W, H) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_forward_params(fp)
        c.forward_depth(cam.camera_data(), len(objs))
        assert np.array_equal(c.forward_keys(), keys), "(depth bits, object index) keys"
        assert np.array_equal(c.depth().view(np.uint32), depth.view(np.uint32)), "depth buffer"
        # and the depth buffer is live for the rest of the path
        levels = min(9, len(oracle.mip_extents(W, H)))
        vis, dout, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), levels), depth)
        c.hiz_build()
        c.occlusion_cull(cam.cull_pc(len(objs), levels))
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt


@pytest.mark.parametrize("flags", [T.FWD_TILE_LIGHTS, T.FWD_TILE_LIGHTS | T.FWD_REINHARD | T.FWD_INSTANCE_COLOR, 0])
def test_forward_shade_bit_exact(backend, oracle, flags):
    W, H = 640, 360
    cam, objs, draws, lights, fp = _forward_scene(W, H, (0.0, 2.0, 22.0), (0, 0, 0), (10, 8, 8), seed=43, n_lights=400,
                                                  metallic=0.3, roughness=0.35, flags=flags)
    meshes = synth.make_lod_meshes(6, seed=3)
    inst, batch, mesh = synth.make_instances(objs, 4, 6, seed=4)
    rng = np.random.default_rng(5)
    inst["color"][:, :3] = rng.uniform(0.1, 1.0, (len(inst), 3)).astype(np.float32)
    keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
    td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    ref = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fp, instances=inst)
    with _ctx(backend, W, H, max_batches=4) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_instances(inst, batch, mesh)
        c.set_forward_params(fp)
        with pytest.raises(backend.AshError):
            c.forward_shade()                                      # no forward_depth yet
        c.forward_depth(cam.camera_data(), len(objs))
        c.upload_camera(cam.camera_data())
        c.light_cull(cam.light_pc(len(lights)))
        _compare_tile_lists(c.tile_data(), td)
        c.forward_shade()
        _assert_half_images_equal(c.color(), ref, "forward HDR colour")
        lit = c.color()[..., :3].astype(np.float32)
        assert np.isfinite(lit).all() and lit.max() > 0.2


@pytest.mark.parametrize("graph", [0, 1])
def test_view_run_from_scene_to_ldr_with_forward_stages(backend, oracle, graph):
    """The whole widened view: objects + lights + camera in, depth and HDR colour made on the device, then Hi-Z,
    cull, light cull, TAA, bloom, tonemap — nothing but the scene crosses the host boundary; eager and graph replay."""
    from ash_renderer_b200.host import TemporalAA
    W, H = 640, 360
    objs, draws = synth.make_objects((10, 8, 8), spacing=2.0, seed=51, jitter=0.3)
    lights = synth.make_lights(300, extent=10.0, seed=52, center=(0.0, 2.0, 4.0))
    fp = synth.forward_params(near=0.5, flags=T.FWD_TILE_LIGHTS)
    taa = TemporalAA()
    hist = None
    stages = T.STAGE_ALL | T.STAGE_KEEP_BLOOM | T.STAGE_FORWARD_DEPTH | T.STAGE_FORWARD_SHADE
    with _ctx(backend, W, H, graph_replay=graph) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_forward_params(fp)
        for t in range(5):
            taa.begin_frame()
            pc = taa.push_constants(W, H)
            cam = synth.Camera((0.4 * t - 1.0, 2.0, 22.0 - 0.7 * t), (0, 0, 0), W, H, near=0.5)
            c.view_run(_view_params(cam, len(objs), len(lights), 9, W, H, pc, stages=stages, view_id=t))
            if t in (0, 4):                                        # first view (captured) and a replayed one
                keys, depth = oracle.forward_depth(objs, cam.camera_data(), fp, W, H)
                vis, dout, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), 9), depth)
                td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
                hdr = oracle.forward_shade(keys, objs, lights, td, cam.camera_data(), fp)
                assert np.array_equal(c.forward_keys(), keys)
                assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt
                _compare_tile_lists(c.tile_data(), td)
                _assert_half_images_equal(c.color(), hdr, f"HDR view {t}")
                assert int(c.stats()["view_id"]) == t
            else:
                hdr = c.color()
            ref = oracle.taa_resolve(hdr, hdr if hist is None else hist, None, pc)
            _assert_half_images_equal(c.taa_out(), ref, f"taa view {t}")
            hist = ref
        U, _ = oracle.bloom_chain(hist, 0.8, 5)
        _assert_half_images_equal(c.bloom_mip(0), U[0], "bloom")
        d = np.abs(c.ldr8().astype(np.int16) - oracle.tonemap(hist, U[0], _post_pc()).astype(np.int16))
        assert d.max() <= 1
        cap, rep = c.graph_stats()
        assert (cap > 0 and rep > 0) if graph else (cap == 0 and rep == 0)


@pytest.mark.parametrize("graph", [0, 1])
def test_forward_pass_velocity_feeds_the_taa(backend, oracle, graph):
    """Velocity output of the forward pass (RG16F, from previous-frame transforms) consumed by the TAA's velocity branch
    (taa_resolve.comp:94-102), end to end through view_run with a moving camera: velocity bit-exact vs the oracle, the
    resolve bit-exact vs the oracle's TAA fed with that velocity; the static-camera view yields velocity 0 on every hit."""
    from ash_renderer_b200.host import TemporalAA
    W, H = 640, 360
    objs, draws = synth.make_objects((10, 8, 8), spacing=2.0, seed=61, jitter=0.3)
    lights = synth.make_lights(200, extent=10.0, seed=62, center=(0.0, 2.0, 4.0))
    fp = synth.forward_params(near=0.5, flags=T.FWD_TILE_LIGHTS)
    taa = TemporalAA()
    stages = T.STAGE_ALL | T.STAGE_FORWARD_DEPTH | T.STAGE_FORWARD_SHADE
    eye = lambda t: (0.05 * t - 1.0, 2.0 + 0.02 * t, 22.0 - 0.08 * t) if t != 3 else eye(2)     # view 3 repeats view 2's camera
    hist, prev_cam = None, None
    with _ctx(backend, W, H, graph_replay=graph) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_forward_params(fp)
        for t in range(5):
            taa.begin_frame()
            pc = taa.push_constants(W, H)
            cam = synth.Camera(eye(t), (0, 0, 0), W, H, near=0.5)
            reproj = synth.reprojection(cam, prev_cam if prev_cam is not None else cam)
            c.set_forward_reprojection(reproj)
            c.view_run(_view_params(cam, len(objs), len(lights), 9, W, H, pc, stages=stages, view_id=t))
            keys = c.forward_keys()
            vel = oracle.forward_velocity(keys, objs, cam.camera_data(), fp, reproj)
            got = c.velocity()
            assert np.array_equal(got.view(np.uint16), vel.view(np.uint16)), f"velocity of view {t}"
            hit = (keys & np.uint64(0xFFFFFFFF)) != np.uint64(T.FWD_NO_OBJECT)
            assert hit.mean() > 0.2
            mag = np.abs(vel.astype(np.float32)).max(axis=2)
            if t in (0, 3):                                        # same camera as the previous view: rounding noise only
                assert (mag[hit] * W).max() < 0.02
            else:                                                  # the camera moved: sub-pixel to few-pixel motion on the hits
                assert (mag[hit] * W).max() > 0.5 and np.all(mag[~hit] == 0)
            hdr = c.color()
            ref = oracle.taa_resolve(hdr, hdr if hist is None else hist, vel, pc)
            _assert_half_images_equal(c.taa_out(), ref, f"taa view {t}")
            hist, prev_cam = ref, cam
        cap, rep = c.graph_stats()
        assert (cap > 0 and rep > 0) if graph else (cap == 0 and rep == 0)
        c.set_forward_reprojection(None)                           # off again: the jitter-delta path, no velocity buffer
        taa.begin_frame()
        pc = taa.push_constants(W, H)
        c.view_run(_view_params(cam, len(objs), len(lights), 9, W, H, pc, stages=stages, view_id=9))
        _assert_half_images_equal(c.taa_out(), oracle.taa_resolve(c.color(), hist, None, pc), "taa without velocity")
        with pytest.raises(backend.AshError):
            c.velocity()


@pytest.mark.parametrize("velocity", [False, True])
def test_view_run_graph_replay_equals_eager(backend, velocity):
    """view_run through CUDA-graph replay (per-view kernel arguments patched) against the eager launches:
    a moving camera, TAA jitter sequence, instancing stage; every output of every view bit-identical."""
    from ash_renderer_b200 import host
    from ash_renderer_b200.host import TemporalAA
    W, H, n_batches, n_views = 640, 360, 4, 8
    cam0, objs, draws, depth, meshes, inst, batch, mesh = _instancing_scene(W, H, (10, 8, 8), seed=31, n_batches=n_batches)
    lights = synth.make_lights(300, extent=10.0, seed=31, center=(0.0, 2.0, 22.0))
    seq = synth.HdrSequence(W, H, seed=32, n_discs=20)
    vel = synth.random_velocity(W, H, seed=33, still_frac=0.3, max_px=4.0) if velocity else None
    taa = TemporalAA()
    views = []
    for t in range(n_views):
        taa.begin_frame()
        j = taa.current_jitter()
        cam = synth.Camera((0.6 * t - 2.0, 2.0 + 0.1 * t, 22.0 - 0.5 * t), (0.1 * t, 0, 0), W, H)
        vp = _view_params(cam, len(objs), len(lights), 9, W, H, taa.push_constants(W, H),
                          stages=T.STAGE_ALL | T.STAGE_KEEP_BLOOM | T.STAGE_INSTANCING, view_id=100 + t)
        views.append((seq.frame(float(j[0]), float(j[1])), vp))
    outs = {}
    for mode in (0, 1):
        with _ctx(backend, W, H, max_batches=n_batches, graph_replay=mode) as c:
            c.set_objects(objs, draws)
            c.upload_lights(lights)
            c.set_depth(depth)
            c.set_instances(inst, batch, mesh)
            c.set_lod_meshes(meshes)
            c.set_lod_config(host.LodConfig().record(use_visibility=True))
            per_view = []
            for f, vp in views:
                c.set_color(f, vel)
                c.view_run(vp)
                cnt = c.visible_count()
                st = c.instancing_stats()
                per_view.append(dict(vis=c.visibility(), cnt=cnt, draws=c.draws_out(cnt), tiles=c.tile_data(), stats=c.stats().tobytes(),
                                     taa=c.taa_out(), bloom=c.bloom_mip(0), ldr=c.ldr8(), sel=c.lod_selection().tobytes(),
                                     off=c.bucket_offsets(), idraws=c.instance_draws(), istats=st.tobytes(),
                                     iout=c.instances_out(int(st["total_instances"])).tobytes()))
            outs[mode] = per_view
            cap, rep = c.graph_stats()
            if mode == 0:
                assert (cap, rep) == (0, 0)
            else:      # parity 0 first frame (no history), parity 1, parity 0 with history, then replays
                assert cap == 3 and rep == n_views - 3, (cap, rep)
            launches = c.kernel_launches()
        outs[("launches", mode)] = launches
    assert outs[("launches", 0)] == outs[("launches", 1)]
    for t, (a, b) in enumerate(zip(outs[0], outs[1])):
        assert a["cnt"] == b["cnt"] and a["stats"] == b["stats"], f"view {t}: stats"
        for k in ("vis", "draws", "tiles", "off", "idraws"):
            assert np.array_equal(a[k], b[k]), f"view {t}: {k}"
        for k in ("sel", "istats", "iout"):
            assert a[k] == b[k], f"view {t}: {k}"
        for k in ("taa", "bloom", "ldr"):
            assert np.array_equal(np.asarray(a[k]).view(np.uint8), np.asarray(b[k]).view(np.uint8)), f"view {t}: {k}"
    assert len({v["stats"] for v in outs[1]}) == n_views      # the views really differ (view id, counts)


def test_view_io_host_path(backend, oracle):
    """upload_view / download_view (the e2e path): host buffers in, host buffers out."""
    W, H = 320, 180
    cam, objs, draws, depth = _cull_scene(W, H, (6, 6, 6), seed=8)
    lights = synth.make_lights(64, extent=10.0, seed=8, center=(0.0, 2.0, 22.0))
    hdr = synth.random_hdr(W, H, seed=8)
    pc = _taa_pc(W, H)
    vis, dout, cnt, td, _, taa_ref, _, ldr_ref = _oracle_view(oracle, cam, objs, draws, lights, depth, [hdr], [pc], 9)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    o_vis, o_td = np.zeros(len(objs), np.uint32), np.zeros((tiles, 257), np.uint32)
    o_ldr, o_st, o_cnt = np.zeros((H, W, 4), np.uint8), np.zeros(1, T.VIEW_STATS), np.zeros(1, np.uint32)
    with _ctx(backend, W, H) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.upload_view(depth, hdr)
        c.view_run(_view_params(cam, len(objs), len(lights), 9, W, H, pc, view_id=3))
        c.download_view(o_vis, o_td, o_ldr, o_st, o_cnt)
        c.sync()
        # the packed (CSR) form of the tile lists; then a destination that is too small, then a malformed call
        p_vis, p_off, p_lights = np.zeros(len(objs), np.uint32), np.zeros(tiles + 1, np.uint32), np.full(tiles * 8, 0xFFFFFFFF, np.uint32)
        p_ldr, p_cnt = np.zeros((H, W, 4), np.uint8), np.zeros(1, np.uint32)
        c.download_view_packed(p_vis, p_off, p_lights, p_ldr, None, p_cnt)
        c.sync()
        s_off, small = np.zeros(tiles + 1, np.uint32), np.zeros(4, np.uint32)
        c.download_view_packed(None, s_off, small)
        c.sync()
        with pytest.raises(backend.AshError):
            c.download_view_packed(None, p_off, None)
    assert np.array_equal(o_vis, vis) and o_cnt[0] == cnt and o_st["view_id"][0] == 3
    counts = td[:, 0].astype(np.int64)
    expect = np.concatenate([td[t, 1:1 + counts[t]] for t in range(tiles)]).astype(np.uint32)
    assert np.array_equal(p_off, np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)) and int(p_off[-1]) <= len(p_lights)
    assert np.array_equal(p_lights[:int(p_off[-1])], expect)        # same lists as the 257-stride table, back to back
    assert np.array_equal(p_vis, vis) and p_cnt[0] == cnt and np.array_equal(p_ldr, o_ldr)
    assert int(s_off[-1]) == counts.sum() > 4 and np.array_equal(small, expect[:4])   # truncated, total still reported
    _compare_tile_lists(o_td, td)
    assert np.abs(o_ldr.astype(np.int16) - ldr_ref.astype(np.int16)).max() <= 1


def test_golden_vectors_through_cuda(backend):
    from tests.golden.make_golden import inputs, W, H
    g = np.load(GOLDEN)
    i = inputs()
    cam = i["cam"]
    with _ctx(backend, W, H) as c:
        c.set_depth(i["depth"])
        c.set_objects(i["objs"], i["draws"])
        c.upload_lights(i["lights"])
        c.upload_camera(cam.camera_data())
        c.hiz_build()
        flat = np.concatenate([c.hiz_mip(l).ravel() for l in range(1, c.hiz_level_count())])
        assert np.array_equal(flat, g["hiz_flat"])
        c.occlusion_cull(cam.cull_pc(len(i["objs"]), min(10, c.hiz_level_count())))
        assert np.array_equal(c.visibility(), g["visibility"]) and c.visible_count() == g["visible_count"][0]
        assert np.array_equal(c.draws_out(c.visible_count()).view(np.uint32), g["draws_out"])
        c.light_cull(cam.light_pc(len(i["lights"])))
        td = c.tile_data()
        assert np.array_equal(td[:, 0], g["tile_counts"])
        assert np.array_equal(np.concatenate([td[t, 1:1 + td[t, 0]] for t in range(len(td))]), g["tile_lists"])
        assert np.array_equal(c.readback(T.BUF_TILE_MINMAX, np.uint32).reshape(-1, 2), g["tile_minmax"])
        # TAA with an explicit history: seed the history buffer by resolving `hist` with blend 1
        c.set_color(i["hist"], None)
        c.taa_resolve(_taa_pc(W, H, blend=1.0, clamp=0.0, vel=0.0, jitter=(0, 0)))
        c.set_color(i["hdr"], i["vel"])
        c.taa_resolve(i["taa_pc"])
        # the seeded history has alpha 1 and identical rgb, so the golden (history = hist) applies
        assert np.array_equal(c.taa_out().view(np.uint16), g["taa"])
        c.bloom(_bloom_params())
        assert np.array_equal(c.bloom_mip(0).view(np.uint16), g["bloom_u0"])
        c.tonemap(i["post_pc"])
        assert np.abs(c.ldr8().astype(np.int16) - g["ldr8"].astype(np.int16)).max() <= 1


# ------------------------------------------------------------------ size-independent properties at full size
def test_full_size_properties_4k(backend):
    import torch
    W, H = 3840, 2160
    g = torch.Generator(device="cuda").manual_seed(1)
    depth = torch.rand((H, W), generator=g, device="cuda", dtype=torch.float32)
    hdr = (torch.rand((H, W, 4), generator=g, device="cuda") * 4.0).to(torch.float16)
    hdr[..., 3] = 1.0
    torch.cuda.synchronize()
    with _ctx(backend, W, H) as c:
        c.set_depth(depth)
        c.hiz_build()
        # every mip texel equals the max over its footprint (levels with clean 2^l footprints)
        for l in (1, 3, 4, 7):
            w, h = W >> l, H >> l
            ref = depth[:h << l, :w << l].reshape(h, 1 << l, w, 1 << l).amax(dim=(1, 3)).cpu().numpy()
            assert np.array_equal(c.hiz_mip(l), ref)
        assert c.hiz_mip(11)[0, 0] == c.hiz_mip(10)[:2, :2].max()     # floor sizing drops the odd third column (3x2 -> 1x1)
        mm = c.readback(T.BUF_TILE_MINMAX, np.uint32).reshape(135, 240, 2).view(np.float32)
        tiles = depth.reshape(135, 16, 240, 16)
        assert np.array_equal(mm[..., 0], tiles.amin(dim=(1, 3)).cpu().numpy())
        assert np.array_equal(mm[..., 1], tiles.amax(dim=(1, 3)).cpu().numpy())
        # TAA with blend 0 is the identity on rgb; tonemap is monotone in exposure
        c.set_color(hdr, None)
        c.taa_resolve(_taa_pc(W, H, blend=0.0))
        out = c.taa_out()
        assert np.array_equal(out[..., :3], hdr[..., :3].cpu().numpy())
        c.tonemap(_post_pc(exposure=0.5))
        lo = c.ldr8().copy()
        c.tonemap(_post_pc(exposure=1.0))
        assert np.all(c.ldr8()[..., :3] >= lo[..., :3])
        # bloom is linear in neither input nor threshold, but a frame below threshold-knee blooms to 0
        dark = (hdr * 0.07).contiguous()
        c.set_color(dark, None)
        c.bloom(_bloom_params())
        assert not c.bloom_mip(0)[..., :3].any()


def test_post_chain_edge_sizes_bit_exact():
    """tools/sweep_post_sizes.py: bloom chain and TAA at eleven extents chosen for the edges of the tiled kernels (partial
    tiles, second packed lane out of range, halo clamps, 4K itself), every image bit-exact against the oracle."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sweep_post_sizes", os.path.join(os.path.dirname(__file__), "..", "tools", "sweep_post_sizes.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.run(verbose=False) == 0


# ------------------------------------------------------------------ BASELINE configs at their full sizes
def test_config3_view_run_at_4k_matches_oracle(backend, oracle):
    """BASELINE config 3 as written: 3840x2160, 10 000 instances (25x20x20 grid), 4096 lights, full 12-level pyramid,
    occlusion cull + light cull, plus the post chain of the same view (two frames).  Everything against the oracle."""
    from ash_renderer_b200.host import TemporalAA
    W, H = 3840, 2160
    cam = synth.Camera((0.0, 6.0, 70.0), (0, 0, 0), W, H, far=300.0)
    objs, draws = synth.make_objects((25, 20, 20), seed=3, jitter=0.0, disabled_frac=0.02)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(4096, extent=30.0, rmin=1.0, rmax=6.0, seed=3)
    seq = synth.HdrSequence(W, H, seed=4, n_discs=200)
    taa = TemporalAA()
    frames, pcs = [], []
    for _ in range(2):
        taa.begin_frame()
        j = taa.current_jitter()
        frames.append(seq.frame(float(j[0]), float(j[1])))
        pcs.append(taa.push_constants(W, H))
    vis, dout, cnt, td, true_counts, taa_ref, bloom_ref, ldr_ref = _oracle_view(oracle, cam, objs, draws, lights, depth, frames, pcs, 12)
    assert 0 < cnt < len(objs) and int(td[:, 0].sum()) > 10000
    with _ctx(backend, W, H, hiz_levels=12) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_depth(depth)
        for f, pc in zip(frames, pcs):
            c.set_color(f, None)
            c.view_run(_view_params(cam, len(objs), len(lights), 12, W, H, pc, stages=T.STAGE_ALL | T.STAGE_KEEP_BLOOM))
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt and np.array_equal(c.draws_out(cnt), dout)
        _compare_tile_lists(c.tile_data(), td)
        mips, _ = oracle.hiz_build(depth)
        for l in (1, 5, 11):
            assert np.array_equal(c.hiz_mip(l), mips[l - 1])
        _assert_half_images_equal(c.taa_out(), taa_ref, "taa")
        _assert_half_images_equal(c.bloom_mip(0), bloom_ref, "bloom U_0")
        assert np.abs(c.ldr8().astype(np.int16) - ldr_ref.astype(np.int16)).max() <= 1


def test_config5_cull_100k_objects_at_4k(backend, oracle):
    """BASELINE config 5's cull: 100 000 instances against the full pyramid of a 3840x2160 depth buffer."""
    W, H = 3840, 2160
    cam = synth.Camera((20.0, 25.0, 48.0), (2.0, -1.0, 3.0), W, H, near=0.5, far=200.0)
    objs, draws = synth.make_objects((50, 50, 40), spacing=2.0, seed=5, jitter=0.4)
    depth = synth.splat_depth(objs[::8], cam, shrink=0.7)
    pc = cam.cull_pc(len(objs), 12)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    assert 1000 < cnt < len(objs) - 1000
    with _ctx(backend, W, H, hiz_levels=12, enable_post=0) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.occlusion_cull(pc)
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt and np.array_equal(c.draws_out(cnt), dout)


@pytest.mark.parametrize("precision", [T.PRECISION_EXACT, T.PRECISION_FAST])
def test_config4_64_frame_sequence_at_4k(backend, oracle, precision):
    """BASELINE config 4 as written: the 4K post chain over a 64-frame jittered sequence.  The frame
    generator applies the Halton sub-pixel shift analytically; the library carries the TAA recurrence through all 64
    frames.  Frames 0, 1, 15, 16 and 63 are checked against the oracle evaluated on the library's own history of that
    frame (so the oracle runs 5 frames, not 64), then bloom U_0 and the 8-bit image of frame 63.
    exact mode: bit-exact images, LDR8 <= 1 LSB; fast mode: every texel within 1 fp16 ulp, LDR8 <= 1 LSB."""
    from ash_renderer_b200.host import TemporalAA
    from tests.test_precision_fast_gpu import assert_within_ulps
    W, H = 3840, 2160
    seq = synth.HdrSequence(W, H, seed=4, n_discs=200)
    taa = TemporalAA()
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    check = {0, 1, 15, 16, 63}
    base = seq.frame(0.0, 0.0)
    with _ctx(backend, W, H, precision=precision) as c:
        prev = None
        for t in range(64):
            taa.begin_frame()
            j = taa.current_jitter()
            # full re-render only for the checked frames and their predecessors' inputs; the others reuse a cheap
            # integer-shifted copy of the base frame (their resolves are inputs of the recurrence, not checked themselves)
            f = seq.frame(float(j[0]), float(j[1])) if (t in check or t + 1 in check) else np.roll(base, t % 3 - 1, axis=1)
            pc = taa.push_constants(W, H)
            c.set_color(f, None)
            vp = _view_params(cam, 0, 0, 1, W, H, pc, stages=T.STAGE_TAA | T.STAGE_BLOOM | T.STAGE_TONEMAP | T.STAGE_KEEP_BLOOM)
            c.view_run(vp)
            if t in check:
                out = c.taa_out()
                ref = oracle.taa_resolve(f, f if prev is None else prev, None, pc)
                if precision == T.PRECISION_EXACT:
                    _assert_half_images_equal(out, ref, f"frame {t}")
                else:
                    assert_within_ulps(out, ref, 1, f"frame {t}")
                if t == 63:
                    U, _ = oracle.bloom_chain(out, 0.8, 5)
                    if precision == T.PRECISION_EXACT:
                        _assert_half_images_equal(c.bloom_mip(0), U[0], "bloom U_0 of frame 63")
                    else:
                        # U_0 chains the fused threshold (<= 1 ulp of threshold(resolve)) through the bloom levels:
                        # <= 2 ulp, beyond 1 ulp on fewer than one half in a million (measured: 1 of 33 M)
                        assert_within_ulps(c.bloom_mip(0), U[0], 2, "bloom U_0 of frame 63", frac_beyond_one=1e-6)
                    ldr = oracle.tonemap(out, U[0], _post_pc())
                    assert np.abs(c.ldr8().astype(np.int16) - ldr.astype(np.int16)).max() <= 1
            prev = c.taa_out() if (t + 1 in check) else None


# ------------------------------------------------------------------ round-1 review findings
def test_reference_default_hiz_levels_on_a_small_extent(backend, oracle):
    """OcclusionCulling::push_constants always sends HIZ_LEVELS = 10 (occlusion_culling.rs:21,262); at 320x180 the
    pyramid has 9 levels.  textureLod clamps to the last mip: accepted, and equal to the oracle."""
    W, H = 320, 180
    cam, objs, draws, depth = _cull_scene(W, H, (8, 8, 8), seed=3)
    assert len(oracle.mip_extents(W, H)) == 9
    pc = cam.cull_pc(len(objs), 10)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    with _ctx(backend, W, H) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.occlusion_cull(pc)
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt and np.array_equal(c.draws_out(cnt), dout)


def test_draw_index_past_the_template_buffer_reads_zeros(backend, oracle):
    """drawCommandsIn[obj.drawIndex] with drawIndex >= the buffer's size: zeros (robustBufferAccess), not a fault."""
    W, H = 320, 180
    cam, objs, draws, depth = _cull_scene(W, H, (6, 6, 6), seed=4, disabled=0.0)
    objs["draw_index"][::7] = len(draws) + 5
    objs["draw_index"][3] = 0xFFFFFFFF
    pc = cam.cull_pc(len(objs), 9)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    assert cnt > 20 and (dout["index_count"] == 0).any() and (dout["index_count"] != 0).any()
    with _ctx(backend, W, H) as c:
        c.set_depth(depth)
        c.set_objects(objs, draws)
        c.occlusion_cull(pc)
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt and np.array_equal(c.draws_out(cnt), dout)
        c.sync()                                           # no sticky error: the context still works
        c.occlusion_cull(pc)
        assert c.visible_count() == cnt


@pytest.mark.parametrize("W,H", [(4096, 1), (8192, 2), (1, 4096), (3000, 1)])
def test_hiz_degenerate_aspect_ratios(backend, oracle, W, H):
    """Levels whose height is clamped to 1 only halve: the shared-memory tail of k_pyramid_tail must hold W / 2 texels."""
    depth = synth.random_depth(W, H, seed=W + H)
    mips, _ = oracle.hiz_build(depth)
    with _ctx(backend, W, H, enable_post=0) as c:
        c.set_depth(depth)
        c.hiz_build()
        assert c.hiz_level_count() == len(mips) + 1
        for l, m in enumerate(mips, start=1):
            assert np.array_equal(c.hiz_mip(l), m), f"mip {l}"


# ------------------------------------------------------------------ error behaviour (AshError mirror)
def test_error_codes(backend):
    W, H = 64, 64
    with _ctx(backend, W, H, max_objects=16, max_lights=4) as c:
        cam = synth.Camera((0, 0, 5), (0, 0, 0), W, H)
        with pytest.raises(backend.AshError) as e:
            c.hiz_build()
        assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED
        with pytest.raises(backend.AshError) as e:
            c.taa_resolve(_taa_pc(W, H))
        assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED
        objs, draws = synth.make_objects((3, 3, 3), seed=1)
        with pytest.raises(backend.AshError) as e:
            c.set_objects(objs, draws)                    # 27 > max_objects
        assert e.value.code == backend.ASH_ERR_INVALID
        c.set_depth(synth.random_depth(W, H))
        with pytest.raises(backend.AshError) as e:
            c.light_cull(cam.light_pc(1))                 # no camera uploaded
        assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED
        with pytest.raises(backend.AshError) as e:
            c.get(12345)
        assert e.value.code == backend.ASH_ERR_NOT_FOUND
        with pytest.raises(backend.AshError) as e:
            c.occlusion_cull(cam.cull_pc(0, 0))           # hizLevels == 0
        assert e.value.code == backend.ASH_ERR_INVALID
        c.occlusion_cull(cam.cull_pc(0, 99))              # hizLevels > levels_built is legal: the sampler clamps
    with _ctx(backend, W, H, enable_post=0) as c:
        with pytest.raises(backend.AshError) as e:
            c.set_color(synth.random_hdr(W, H), None)
        assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED


# ------------------------------------------------------------------ fp32_intermediates storage (SURVEY PIN-6, second mode)
@pytest.mark.parametrize("W,H,frames", [(416, 240, 8), (135, 77, 3), (1920, 1080, 2)])
def test_fp32_intermediates_mode_within_1e5(backend, oracle, W, H, frames):
    """BASELINE.json north_star: "post-processed images within max-abs 1e-5 in fp32".  With AshCudaConfig.fp32_intermediates
    the TAA output / history and the bloom targets are RGBA32F; the config-4 recurrence (jittered frames, staged calls and
    view_run) is compared with the oracle running the same storage mode: max-abs <= 1e-5 on every image of every frame
    (measured: 0 — the kernels execute the oracle's IEEE sequence), LDR8 <= 1 LSB."""
    from ash_renderer_b200.host import TemporalAA
    seq = synth.HdrSequence(W, H, seed=4, n_discs=30)
    taa = TemporalAA()
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    hist = None
    worst = 0.0
    with _ctx(backend, W, H, fp32_intermediates=1) as c:
        for t in range(frames):
            taa.begin_frame()
            j = taa.current_jitter()
            f = seq.frame(float(j[0]), float(j[1]))
            pc = taa.push_constants(W, H)
            c.set_color(f, None)
            if t % 2 == 0:                                   # staged calls and the fused entry point alternate
                c.taa_resolve(pc)
                c.bloom(_bloom_params())
                c.tonemap(_post_pc())
            else:
                c.view_run(_view_params(cam, 0, 0, 1, W, H, pc, stages=T.STAGE_TAA | T.STAGE_BLOOM | T.STAGE_TONEMAP | T.STAGE_KEEP_BLOOM))
            ref = oracle.taa_resolve(f, f if hist is None else hist, None, pc, out32=True)
            got = c.taa_out()
            assert got.dtype == np.float32 and ref.dtype == np.float32
            worst = max(worst, float(np.abs(got - ref).max()))
            U, D = oracle.bloom_chain(ref, 0.8, 5, out32=True)
            for k in range(6):
                worst = max(worst, float(np.abs(c.bloom_mip(k) - U[k]).max()))
            assert np.abs(c.ldr8().astype(np.int16) - oracle.tonemap(ref, U[0], _post_pc()).astype(np.int16)).max() <= 1
            hist = ref
        assert worst <= 1e-5, f"max-abs {worst:.3e} in fp32"
        assert c.graph_stats() == (0, 0)
