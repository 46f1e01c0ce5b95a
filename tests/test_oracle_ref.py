"""Pins the CPU oracle (oracle/ash_oracle.cpp) to the REFERENCE'S OWN SHADER TEXT.

oracle/_ref/libashref.so is /root/reference/shaders/{hiz_generate,occlusion_cull,light_culling,taa_resolve}.comp and
{bloom_threshold,bloom_downsample,bloom_upsample,tonemapping}.frag compiled as C++ by g++ behind
oracle/ref/glsl_compat.hpp (recipe: oracle/ref/Makefile + strip_glsl.py; no graphics stack involved).  Every test here
demands the restatement to equal that library BIT FOR BIT — Hi-Z mips, visibility, ordered draw commands, per-tile
light lists, every RGBA16F image, the 8-bit output — on the golden inputs, BASELINE config 2 at reduced extent, ragged
extents and the edge cases the domain has.  What the shaders leave to the implementation (sampler filtering, fp
evaluation order, image stores) is pinned by SURVEY.md §8.0 and implemented once in glsl_compat.hpp.
CPU only (not gpu-marked); skipped when neither /root/reference nor a prebuilt oracle/_ref is present."""
import hashlib
import os

import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T

ref = pytest.importorskip("oracle.ref")
pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built and /root/reference absent")

HERE = os.path.dirname(os.path.abspath(__file__))
SHADERS = ["hiz_generate.comp", "occlusion_cull.comp", "light_culling.comp", "taa_resolve.comp",
           "bloom_threshold.frag", "bloom_downsample.frag", "bloom_upsample.frag", "tonemapping.frag"]


def bits(a):
    a = np.ascontiguousarray(a)
    if a.dtype.fields is not None:                     # structured records (draw commands): raw words
        return a.view(np.uint32)
    return a.view({2: np.uint16, 4: np.uint32, 1: np.uint8, 8: np.uint64}[a.dtype.itemsize])


def same(a, b):
    return np.array_equal(bits(a), bits(b))


def taa_pc(W, H, blend=0.9, clamp=1.0, vel=1.0, jitter=(0.25, -0.16666667), prev=(0.0, 0.0)):
    pc = np.zeros(1, T.TAA_PC)
    w, h = np.float32(W), np.float32(H)
    pc["screen_params"][0] = [w, h, np.float32(1) / w, np.float32(1) / h]
    pc["params"][0] = [blend, clamp, vel, 0.0]
    pc["jitter"][0], pc["prev_jitter"][0] = jitter, prev
    return pc


def post_pc(exposure=1.0, gamma=2.2, bloom=0.5):
    pc = np.zeros(1, T.POST_PC)
    pc["exposure"], pc["gamma"], pc["bloom_intensity"] = exposure, gamma, bloom
    return pc


def compare_tiles(o_td, o_true, r_td, r_true):
    assert np.array_equal(o_true, r_true)
    assert np.array_equal(o_td[:, 0], r_td[:, 0])
    valid = np.arange(T.TILE_STRIDE)[None, :] < (1 + o_td[:, :1])
    # fibres run the 256 invocations in ascending gl_LocalInvocationIndex, so the shader's own atomics yield
    # ascending light order: the lists are compared as sequences, not as sets
    assert np.array_equal(o_td[valid], r_td[valid])
    assert np.all(r_td[~valid] == 0xFFFFFFFF)            # PIN-8: the shader writes nothing beyond 1 + count


# ------------------------------------------------------------------ provenance
def test_reference_shader_hashes_recorded():
    """The golden file names the exact shader text it was generated from."""
    g = np.load(os.path.join(HERE, "golden", "ref_small.npz"))
    recorded = dict(zip(g["shader_names"].tolist(), g["shader_sha256"].tolist()))
    assert sorted(recorded) == sorted(SHADERS)
    if os.path.isdir(ref.REFERENCE_SHADERS):
        for s in SHADERS:
            with open(os.path.join(ref.REFERENCE_SHADERS, s), "rb") as f:
                assert hashlib.sha256(f.read()).hexdigest() == recorded[s], s


def test_fp16_conversion_of_the_compat_layer_equals_numpy_and_the_oracle(oracle):
    """imageStore's rgba16f rounding (gcc _Float16) == numpy.float16 == the oracle's bit-twiddling f2h."""
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-70000, 70000, 4000), rng.uniform(-1e-4, 1e-4, 4000), rng.standard_normal(4000),
                        [0.0, -0.0, 65504.0, 65519.9, 65520.0, 5.96e-8, 2.98e-8, 2.9802325e-8, 6.1e-5]]).astype(np.float32)
    lib = ref.lib()
    got = np.array([lib.ash_ref_f2h(float(v)) for v in x], np.uint16)
    with np.errstate(over="ignore"):
        assert np.array_equal(got, x.astype(np.float16).view(np.uint16))
    assert np.array_equal(got, np.array([oracle.lib().ash_oracle_f2h(float(v)) for v in x], np.uint16))


# ------------------------------------------------------------------ the golden inputs: all eight shaders
def test_golden_inputs_all_shaders(oracle):
    from tests.golden.make_golden import inputs
    i = inputs()
    cam = i["cam"]
    mips, flat = oracle.hiz_build(i["depth"])
    _, rflat = ref.hiz_build(i["depth"])
    assert same(flat, rflat)
    pc = cam.cull_pc(len(i["objs"]), min(10, len(mips) + 1))
    vis, dout, cnt = oracle.occlusion_cull(i["objs"], i["draws"], pc, i["depth"], flat)
    rvis, rdout, rcnt = ref.occlusion_cull(i["objs"], i["draws"], pc, i["depth"], rflat)
    assert cnt == rcnt > 0 and np.array_equal(vis, rvis) and same(dout, rdout)
    lpc, cd = cam.light_pc(len(i["lights"])), cam.camera_data()
    td, true_counts, _ = oracle.light_cull(i["lights"], lpc, cd, i["depth"])
    rtd, rtrue = ref.light_cull(i["lights"], lpc, cd, i["depth"])
    compare_tiles(td, true_counts, rtd, rtrue)
    assert int(true_counts.sum()) > 1000
    for vel in (i["vel"], None):
        taa = oracle.taa_resolve(i["hdr"], i["hist"], vel, i["taa_pc"])
        assert same(taa, ref.taa_resolve(i["hdr"], i["hist"], vel, i["taa_pc"]))
    U, D = oracle.bloom_chain(taa, 0.8, 5)
    RU, RD = ref.bloom_chain(taa, 0.8, 5)
    for k in range(6):
        assert same(D[k], RD[k]) and same(U[k], RU[k]), k
    assert np.array_equal(oracle.tonemap(taa, U[0], i["post_pc"]), ref.tonemap(taa, RU[0], i["post_pc"]))


# ------------------------------------------------------------------ BASELINE config 2 at reduced extent
def test_config2_reduced_extent(oracle):
    """1000 instances on a 10^3 grid, 256 lights, 16x16 tiles, Hi-Z + cull + light cull at
    960x540; then the post chain on a cfg-4 frame of the same extent."""
    W, H = 960, 540
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((10, 10, 10), seed=2)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(256, extent=10.0, seed=2)
    mips, flat = oracle.hiz_build(depth)
    _, rflat = ref.hiz_build(depth)
    assert same(flat, rflat)
    pc = cam.cull_pc(len(objs), 10)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth, flat)
    rvis, rdout, rcnt = ref.occlusion_cull(objs, draws, pc, depth, rflat)
    assert cnt == rcnt and 0 < cnt < len(objs) and np.array_equal(vis, rvis) and same(dout, rdout)
    td, tc, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    rtd, rtc = ref.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    compare_tiles(td, tc, rtd, rtc)
    seq = synth.HdrSequence(W, H, seed=4, n_discs=60)
    f0, f1 = seq.frame(0.0, -0.16666667), seq.frame(-0.25, 0.16666667)
    pc1 = taa_pc(W, H, jitter=(-0.25, 0.16666667), prev=(0.0, -0.16666667))
    taa = oracle.taa_resolve(f1, f0, None, pc1)
    assert same(taa, ref.taa_resolve(f1, f0, None, pc1))
    U, D = oracle.bloom_chain(taa, 0.8, 5)
    RU, RD = ref.bloom_chain(taa, 0.8, 5)
    for k in range(6):
        assert same(D[k], RD[k]) and same(U[k], RU[k]), k
    assert np.array_equal(oracle.tonemap(taa, U[0], post_pc()), ref.tonemap(taa, RU[0], post_pc()))


# ------------------------------------------------------------------ per shader: ragged extents and edge cases
@pytest.mark.parametrize("W,H", [(67, 45), (1, 1), (2, 1), (1, 7), (1000, 3), (255, 257), (130, 70)])
def test_hiz_ragged_extents(oracle, W, H):
    depth = synth.random_depth(W, H, seed=W * 31 + H)
    a, b = oracle.hiz_build(depth)[0], ref.hiz_build(depth)[0]
    assert len(a) == len(b) == len(oracle.mip_extents(W, H)) - 1
    assert all(same(x, y) for x, y in zip(a, b))


@pytest.mark.parametrize("hiz_levels", [1, 3, 8, 10])
def test_cull_levels_and_camera_inside_the_grid(oracle, hiz_levels):
    """Eye inside the grid: corners with clip.w <= 0 (occlusion_cull.comp:145-147), objects straddling planes."""
    W, H = 512, 288
    cam = synth.Camera((0.3, 0.2, 1.1), (4.0, 1.0, -6.0), W, H)
    objs, draws = synth.make_objects((12, 12, 12), seed=5, jitter=0.4, disabled_frac=0.1)
    depth = synth.splat_depth(objs, cam)
    pc = cam.cull_pc(len(objs), hiz_levels)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    rvis, rdout, rcnt = ref.occlusion_cull(objs, draws, pc, depth)
    assert cnt == rcnt and np.array_equal(vis, rvis) and same(dout, rdout)


def test_cull_hiz_levels_above_the_pyramid_clamps_like_the_sampler(oracle):
    """The reference pushes hizLevels = 10 whatever the extent (occlusion_culling.rs:21,262); textureLod clamps."""
    W, H = 200, 120                                  # 8 levels
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((7, 7, 7), seed=6, jitter=0.3)
    depth = synth.splat_depth(objs, cam)
    for levels in (10, 12, 99):
        pc = cam.cull_pc(len(objs), levels)
        vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
        rvis, rdout, rcnt = ref.occlusion_cull(objs, draws, pc, depth)
        assert cnt == rcnt and np.array_equal(vis, rvis) and same(dout, rdout)


def test_cull_non_finite_and_degenerate_objects(oracle):
    W, H = 320, 180
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((5, 5, 5), seed=9, jitter=0.2)
    objs["center"][3, 0] = np.nan
    objs["extents"][7, :3] = np.inf
    objs["model"][11, 3, :3] = [np.inf, 0, 0]
    objs["extents"][13, :3] = 0.0
    objs["model"][17] = 0.0
    objs["extents"][19, :3] = -0.5
    objs["flags"][23] = 2            # bit 0 clear: disabled
    depth = synth.splat_depth(objs[24:], cam)
    pc = cam.cull_pc(len(objs), 9)
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    rvis, rdout, rcnt = ref.occlusion_cull(objs, draws, pc, depth)
    assert cnt == rcnt and np.array_equal(vis, rvis) and same(dout, rdout)


def test_cull_100k_objects_and_the_log2_caveat(oracle):
    """100 k objects (BASELINE config 5's count) at 1024^2.  The oracle pins ceil(log2 s) to the exact integer form
    (SURVEY §8.1 a-2); the compat layer evaluates the shader's `ceil(log2(x))` with glibc log2f.  They may differ
    only where log2f rounds 2^k(1+eps) to exactly k; such objects are reported, and none may differ otherwise."""
    W = H = 1024
    cam = synth.Camera((0.0, 30.0, 90.0), (0, 0, 0), W, H, far=400.0)
    objs, draws = synth.make_objects((50, 50, 40), seed=5, jitter=0.5)
    depth = synth.splat_depth(objs, cam)
    pc = cam.cull_pc(len(objs), 11)
    vis, _, cnt = oracle.occlusion_cull(objs, draws, pc, depth)
    rvis, _, rcnt = ref.occlusion_cull(objs, draws, pc, depth)
    diff = np.nonzero(vis != rvis)[0]
    print(f"100k objects: {cnt} visible (oracle) / {rcnt} (reference text), {len(diff)} differ (log2 knife edges)")
    assert len(diff) <= 2 and 0 < cnt < len(objs)


def test_light_overflow_disabled_directional_and_zero_lights(oracle):
    W, H = 64, 48
    cam = synth.Camera((0.0, 0.0, 8.0), (0, 0, 0), W, H)
    depth = synth.random_depth(W, H, seed=3)
    lights = synth.make_lights(600, extent=2.0, rmin=50.0, rmax=80.0, seed=3, n_directional=8, enabled_p=0.8)
    td, tc, _ = oracle.light_cull(lights, cam.light_pc(600), cam.camera_data(), depth)
    rtd, rtc = ref.light_cull(lights, cam.light_pc(600), cam.camera_data(), depth)
    assert tc.max() > 256                                   # sharedLightCount overflows MAX_LIGHTS_PER_TILE
    compare_tiles(td, tc, rtd, rtc)
    td0, tc0, _ = oracle.light_cull(lights, cam.light_pc(0), cam.camera_data(), depth)
    rtd0, rtc0 = ref.light_cull(lights, cam.light_pc(0), cam.camera_data(), depth)   # early-out path :160-165
    compare_tiles(td0, tc0, rtd0, rtc0)
    assert np.all(rtd0[:, 0] == 0)


@pytest.mark.parametrize("W,H,seed", [(200, 120, 1), (67, 45, 2), (33, 9, 3)])
def test_light_ragged_extents_and_odd_cameras(oracle, W, H, seed):
    cam = synth.Camera((1.0, 3.0, 12.0), (0.5, 0.0, -1.0), W, H, fov_deg=70.0, near=0.1, far=50.0)
    objs, _ = synth.make_objects((5, 5, 5), seed=seed, jitter=0.3)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(300, extent=8.0, seed=seed, center=(1.0, 3.0, 12.0))
    lights["position"][5] = [np.nan, 0, 0, 1.0]
    lights["position"][6, 3] = -1.0
    td, tc, _ = oracle.light_cull(lights, cam.light_pc(300), cam.camera_data(), depth)
    rtd, rtc = ref.light_cull(lights, cam.light_pc(300), cam.camera_data(), depth)
    compare_tiles(td, tc, rtd, rtc)


@pytest.mark.parametrize("blend,clamp,vel,jitter", [(0.9, 1.0, 1.0, (0.25, -0.16666667)), (0.9, 0.0, 1.0, (0.0, 0.0)),
                                                      (0.5, 1.0, 0.0, (-0.375, 0.38888889)), (1.0, 1.0, 1.0, (3.5, -2.25))])
@pytest.mark.parametrize("with_velocity", [False, True])
def test_taa_parameter_space(oracle, blend, clamp, vel, jitter, with_velocity):
    W, H = 97, 61
    hdr, hist = synth.random_hdr(W, H, seed=11), synth.random_hdr(W, H, seed=12)
    v = synth.random_velocity(W, H, seed=13) if with_velocity else None
    pc = taa_pc(W, H, blend, clamp, vel, jitter)
    assert same(oracle.taa_resolve(hdr, hist, v, pc), ref.taa_resolve(hdr, hist, v, pc))


@pytest.mark.parametrize("W,H", [(135, 77), (64, 64), (33, 9), (5, 3), (1, 1), (270, 135)])
def test_bloom_chain_and_tonemap_odd_extents(oracle, W, H):
    """135 -> 67: the odd level whose last row / column the floor sizing drops (bilinear taps off texel centres)."""
    hdr = synth.random_hdr(W, H, seed=W + H, hot_frac=0.05)
    U, D = oracle.bloom_chain(hdr, 0.8, 5)
    RU, RD = ref.bloom_chain(hdr, 0.8, 5)
    for k in range(6):
        assert same(D[k], RD[k]) and same(U[k], RU[k]), k
    for pc in (post_pc(), post_pc(2.0, 2.4, 1.0), post_pc(0.25, 1.0, 0.0)):
        assert np.array_equal(oracle.tonemap(hdr, U[0], pc), ref.tonemap(hdr, RU[0], pc))
    assert np.array_equal(oracle.tonemap(hdr, None, post_pc()), ref.tonemap(hdr, None, post_pc()))


# ------------------------------------------------------------------ fp32_intermediates storage (PIN-6, second mode)
@pytest.mark.parametrize("W,H", [(200, 120), (135, 77)])
def test_fp32_intermediates_mode(oracle, W, H):
    """TAA output / history and the bloom targets stored as RGBA32F instead of RGBA16F (SURVEY PIN-6's fp32_intermediates):
    the oracle equals the reference's shader text bit for bit in this mode too, over a 3-frame recurrence + bloom + tonemap;
    rounding the fp32 first-frame resolve to fp16 gives the RGBA16F mode's resolve exactly."""
    seq = synth.HdrSequence(W, H, seed=4, n_discs=25)
    jit = [(0.0, -0.16666667), (-0.25, 0.16666667), (0.25, -0.38888889)]
    hist = rhist = None
    prev = (0.0, 0.0)
    for t, j in enumerate(jit):
        f = seq.frame(*j)
        pc = taa_pc(W, H, jitter=j, prev=prev)
        out = oracle.taa_resolve(f, f if hist is None else hist, None, pc, out32=True)
        rout = ref.taa_resolve(f, f if rhist is None else rhist, None, pc, out32=True)
        assert out.dtype == np.float32 and same(out, rout), t
        if t == 0:
            assert same(out.astype(np.float16), oracle.taa_resolve(f, f, None, pc))
        hist, rhist, prev = out, rout, j
    U, D = oracle.bloom_chain(hist, 0.8, 5, out32=True)
    RU, RD = ref.bloom_chain(rhist, 0.8, 5, out32=True)
    for k in range(6):
        assert D[k].dtype == np.float32 and same(D[k], RD[k]) and same(U[k], RU[k]), k
    assert np.array_equal(oracle.tonemap(hist, U[0], post_pc()), ref.tonemap(rhist, RU[0], post_pc()))


# ------------------------------------------------------------------ the PIN-6 weight model
def test_exact_weight_mode_also_matches_and_bounds_the_pin6_change(oracle):
    """SURVEY PIN-6 first pinned exact fp32 bilinear weights; the library pins 8-bit sub-texel weights.
    Both oracles implement both; they agree with each other in either mode, and the two MODELS are compared on the
    config-4 frame generator so the size of the change is a number, not an argument."""
    W, H = 480, 270
    seq = synth.HdrSequence(W, H, seed=4, n_discs=40)
    f0, f1 = seq.frame(0.0, -0.16666667), seq.frame(-0.25, 0.16666667)
    pc1 = taa_pc(W, H, jitter=(-0.25, 0.16666667), prev=(0.0, -0.16666667))
    out = {}
    try:
        for mode in (oracle.WEIGHTS_Q8, oracle.WEIGHTS_EXACT):
            oracle.set_weight_mode(mode)
            ref.set_weight_mode(mode)
            taa = oracle.taa_resolve(f1, f0, None, pc1)
            assert same(taa, ref.taa_resolve(f1, f0, None, pc1))
            U, D = oracle.bloom_chain(taa, 0.8, 5)
            RU, RD = ref.bloom_chain(taa, 0.8, 5)
            for k in range(6):
                assert same(D[k], RD[k]) and same(U[k], RU[k]), (mode, k)
            ldr = oracle.tonemap(taa, U[0], post_pc())
            assert np.array_equal(ldr, ref.tonemap(taa, RU[0], post_pc()))
            out[mode] = (taa, U[0], ldr)
    finally:
        oracle.set_weight_mode(oracle.WEIGHTS_Q8)
        ref.set_weight_mode(ref.WEIGHTS_Q8)
    (t_q, u_q, l_q), (t_e, u_e, l_e) = out[oracle.WEIGHTS_Q8], out[oracle.WEIGHTS_EXACT]

    def ulps(a, b):
        a, b = a.view(np.int16).astype(np.int32), b.view(np.int16).astype(np.int32)
        return int(np.abs(a - b).max())
    d_taa = float(np.abs(t_q[..., :3].astype(np.float32) - t_e[..., :3].astype(np.float32)).max())
    d_u0 = float(np.abs(u_q[..., :3].astype(np.float32) - u_e[..., :3].astype(np.float32)).max())
    print(f"PIN-6 exact vs 8-bit weights on cfg-4 frames: TAA max-abs {d_taa:.3e} ({ulps(t_q, t_e)} fp16 ulp), "
          f"bloom U_0 max-abs {d_u0:.3e} ({ulps(u_q, u_e)} fp16 ulp), LDR8 max {int(np.abs(l_q.astype(int) - l_e.astype(int)).max())} LSB")
    # a weight error of at most 2^-9 per axis on texels of radiance <= 60 (disc + base): the models differ by a few
    # fp16 ulps at disc edges (measured: TAA 6.6e-2 = 17 ulp at radiance ~30, U_0 7.8e-2) and by <= 1 LSB after the
    # tonemap; DESIGN.md §2 quotes these numbers
    assert d_taa <= 60.0 * 2.0 / 512.0 and d_u0 <= 60.0 * 2.0 / 512.0
    assert int(np.abs(l_q.astype(int) - l_e.astype(int)).max()) <= 1
