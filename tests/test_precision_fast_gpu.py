"""The TOLERANCE precision mode (AshCudaConfig.precision = ASH_PRECISION_FAST) of the post chain against the CPU oracle.

BASELINE.json's north_star grants the post-processed images "max-abs 1e-5 in fp32 / 1 LSB on 8-bit output"; with
RGBA16F storage (PIN-6) the fp32 criterion reads "within 1 fp16 ulp" (SURVEY.md §7).  The fast mode spends exactly
that: approximate reciprocals and FMA contraction in the colour arithmetic.  Sampling positions and weights,
visibility and light lists stay exact.  Bars stated in each test: every RGBA16F texel within 1 fp16 ulp of the oracle
evaluated on the SAME inputs, 8-bit output within 1 LSB.  The exact mode (the library default) is covered by
tests/test_parity_gpu.py and stays bit-exact."""
import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T
from tests.test_parity_gpu import _ctx, _taa_pc, _post_pc, _bloom_params, _view_params, _cull_scene, _compare_tile_lists

pytestmark = pytest.mark.gpu
FAST = dict(precision=T.PRECISION_FAST)


def half_ulp_distance(got, ref):
    g = np.ascontiguousarray(got).view(np.uint16).astype(np.int32)
    r = np.ascontiguousarray(ref).view(np.uint16).astype(np.int32)
    mono = lambda u: np.where(u & 0x8000, -(u & 0x7FFF), u)     # sign-magnitude -> monotone integer
    return np.abs(mono(g) - mono(r))


def assert_within_ulps(got, ref, ulps, what, max_frac=None, frac_beyond_one=None):
    d = half_ulp_distance(got, ref)
    assert d.max() <= ulps, f"{what}: max {int(d.max())} fp16 ulp ({int((d > ulps).sum())} of {d.size} halfs beyond {ulps})"
    if frac_beyond_one is not None:
        assert (d > 1).mean() <= frac_beyond_one, f"{what}: {int((d > 1).sum())} of {d.size} halfs beyond 1 ulp"
    if max_frac is not None:
        assert (d > 0).mean() <= max_frac, f"{what}: {(d > 0).mean():.4f} of the halfs differ"
    return float((d > 0).mean())


def test_history_regularity_check_separates_the_two_variants(backend):
    """Host-side plan (exact_taps.cu): Halton deltas are regular, a delta on an 8-bit weight boundary is not."""
    lib = backend.load_library()
    reg = lambda n, d: lib.ashcuda_debug_exact_taps(4, 0, n, np.float32(n), np.float32(d))
    for n in (200, 120, 1920, 1080, 3840, 2160):
        assert reg(n, 0.25) and reg(n, -0.16666667) and reg(n, 0.0) and reg(n, -0.85)
        assert not reg(n, 0.001953125) and not reg(n, 0.998046875)


@pytest.mark.parametrize("W,H", [(200, 120), (67, 45), (33, 9), (1920, 1080), (1000, 3)])
def test_taa_fast_within_one_ulp_white_noise(backend, oracle, W, H):
    """White-noise HDR with hot texels (the worst case for the neighbourhood clip), four parameter sets."""
    cur, hist = synth.random_hdr(W, H, seed=1), synth.random_hdr(W, H, seed=2)
    with _ctx(backend, W, H, **FAST) as c:
        for kw in (dict(), dict(clamp=0.0), dict(vel=0.0, jitter=(-0.4, 0.3), prev=(0.45, -0.45)), dict(blend=0.0),
                   dict(jitter=(0.0, 0.0), prev=(0.0, 0.0)),
                   # a delta of 1/512 px sits on a rounding boundary of the 8-bit weight: single columns flip to the
                   # neighbouring weight, the frame is NOT regular and the per-tile table variant runs
                   dict(jitter=(0.001953125, 0.3), prev=(0.0, 0.0)), dict(jitter=(0.1, 0.998046875), prev=(0.0, 0.0))):
            pc = _taa_pc(W, H, **kw)
            c.taa_reset_history()
            c.set_color(hist, None)
            c.taa_resolve(_taa_pc(W, H, blend=1.0, clamp=0.0, vel=0.0, jitter=(0, 0)))   # out = hist (history = colour, blend 1)
            ref_hist = c.taa_out()
            assert_within_ulps(ref_hist[..., :3], hist[..., :3], 1, "history seed")
            c.set_color(cur, None)
            c.taa_resolve(pc)
            ref = oracle.taa_resolve(cur, ref_hist, None, pc)
            assert_within_ulps(c.taa_out(), ref, 1, f"fast taa {kw}")


def test_taa_fast_64_frame_sequence(backend, oracle):
    """BASELINE config 4's recurrence at a size the oracle follows in seconds: 64 jittered frames, every resolve within
    1 fp16 ulp of the oracle evaluated on the SAME history (the library's previous output)."""
    from ash_renderer_b200.host import TemporalAA
    W, H = 416, 240
    seq = synth.HdrSequence(W, H, seed=4, n_discs=20)
    taa = TemporalAA()
    hist = None
    worst = 0.0
    with _ctx(backend, W, H, **FAST) as c:
        for t in range(64):
            taa.begin_frame()
            j = taa.current_jitter()
            frame = seq.frame(float(j[0]), float(j[1]))
            pc = taa.push_constants(W, H)
            c.set_color(frame, None)
            c.taa_resolve(pc)
            ref = oracle.taa_resolve(frame, frame if hist is None else hist, None, pc)
            worst = max(worst, assert_within_ulps(c.taa_out(), ref, 1, f"frame {t}"))
            hist = c.taa_out()
    assert worst < 0.05      # a few percent of the halfs differ (by one ulp) at most


@pytest.mark.parametrize("W,H", [(320, 180), (1920, 1080), (67, 45)])
def test_view_run_fast_mode(backend, oracle, W, H):
    """Whole view in fast mode: integer outputs bit-exact; TAA and the fused threshold within 1 ulp of the oracle on
    the same inputs; the 8-bit image within 1 LSB of the oracle's own end-to-end result."""
    from ash_renderer_b200.host import TemporalAA
    cam, objs, draws, depth = _cull_scene(W, H, (8, 8, 8), seed=5)
    lights = synth.make_lights(256, extent=10.0, seed=5, center=(0.0, 2.0, 22.0))
    seq = synth.HdrSequence(W, H, seed=6, n_discs=30)
    taa = TemporalAA()
    levels = min(10, len(oracle.mip_extents(W, H)))
    vis, dout, cnt = oracle.occlusion_cull(objs, draws, cam.cull_pc(len(objs), levels), depth)
    td, _, _ = oracle.light_cull(lights, cam.light_pc(len(lights)), cam.camera_data(), depth)
    hist_o = hist_c = None
    with _ctx(backend, W, H, **FAST) as c:
        c.set_objects(objs, draws)
        c.upload_lights(lights)
        c.set_depth(depth)
        for t in range(3):
            taa.begin_frame()
            j = taa.current_jitter()
            f = seq.frame(float(j[0]), float(j[1]))
            pc = taa.push_constants(W, H)
            c.set_color(f, None)
            c.view_run(_view_params(cam, len(objs), len(lights), levels, W, H, pc, stages=T.STAGE_ALL | T.STAGE_KEEP_BLOOM))
            ref = oracle.taa_resolve(f, f if hist_c is None else hist_c, None, pc)      # same history as the library
            assert_within_ulps(c.taa_out(), ref, 1, f"taa frame {t}")
            hist_o = oracle.taa_resolve(f, f if hist_o is None else hist_o, None, pc)   # the oracle's own recurrence
            hist_c = c.taa_out()
        assert np.array_equal(c.visibility(), vis) and c.visible_count() == cnt and np.array_equal(c.draws_out(cnt), dout)
        _compare_tile_lists(c.tile_data(), td)
        U, _ = oracle.bloom_chain(hist_o, 0.8, 5)
        ldr_ref = oracle.tonemap(hist_o, U[0], _post_pc())
        d = np.abs(c.ldr8().astype(np.int16) - ldr_ref.astype(np.int16))
        assert d.max() <= 1, f"LDR8 differs by {int(d.max())} LSB"
        # bloom of the library's own resolve, stage by stage: U_0 within 1 ulp of the oracle chain on that resolve
        Uc, _ = oracle.bloom_chain(hist_c, 0.8, 5)
        # two chained stages (the fused threshold, within 1 ulp of threshold(resolve), feeds the exact bloom chain):
        # 2 ulp at most, and beyond 1 ulp on fewer than one texel in a million
        assert_within_ulps(c.bloom_mip(0), Uc[0], 2, "bloom U_0", frac_beyond_one=1e-6)


@pytest.mark.parametrize("W,H", [(1920, 1080), (3840, 2160), (1984, 1096), (200, 120)])
def test_bloom_chain_fast_mode(backend, oracle, W, H):
    """Staged bloom in fast mode: the large upsample levels run the 4x4-convolution form of the tent (post_up_pair.cu).
    Every level within 2 fp16 ulp of the oracle's chain, beyond 1 ulp on fewer than 1e-5 of the halfs (the levels chain)."""
    hdr = synth.random_hdr(W, H, seed=W + 1)
    U, D = oracle.bloom_chain(hdr, 0.8, 5)
    with _ctx(backend, W, H, **FAST) as c:
        c.set_color(hdr, None)
        c.bloom(_bloom_params(0.8, 5))
        for k in range(5, -1, -1):
            assert_within_ulps(c.bloom_mip(k), U[k], 2, f"U_{k}", frac_beyond_one=1e-5)
        c.tonemap(_post_pc())
        d = np.abs(c.ldr8().astype(np.int16) - oracle.tonemap(hdr, U[0], _post_pc()).astype(np.int16))
        assert d.max() <= 1


def test_fast_threshold_is_of_the_stored_texel(backend, oracle):
    """The fused bloom threshold reads the resolve as stored (fp16): D_0 within 1 ulp of threshold(resolve)."""
    W, H = 640, 360
    hdr, hist = synth.random_hdr(W, H, seed=7, hot_frac=0.1), synth.random_hdr(W, H, seed=8, hot_frac=0.1)
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    with _ctx(backend, W, H, enable_post=1, **FAST) as c:
        c.set_color(hist, None)
        c.taa_resolve(_taa_pc(W, H, blend=1.0, clamp=0.0, vel=0.0, jitter=(0, 0)))
        c.set_color(hdr, None)
        vp = _view_params(cam, 0, 0, 1, W, H, _taa_pc(W, H), stages=T.STAGE_TAA | T.STAGE_BLOOM)
        vp["bloom"]["mip_count"] = 0
        c.view_run(vp)
        d0 = c.readback(T.BUF_BLOOM_THRESHOLD, np.float16).reshape(H, W, 4)
        assert_within_ulps(d0, oracle.bloom_threshold(c.taa_out(), 0.8), 1, "D_0")
