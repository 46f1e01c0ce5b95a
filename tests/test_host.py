"""Host-side mirror of the reference's CPU state (ash_renderer_b200/host.py) and view sharding.
The cases follow the reference's own in-file unit tests for this boundary."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from ash_renderer_b200 import host, types as T, views

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- occlusion_culling.rs:290-314
def test_bounding_box_from_min_max():
    c, e = host.cull_bounding_box_from_min_max((-1.0, -1.0, -1.0), (1.0, 1.0, 1.0))
    assert np.array_equal(c, [0, 0, 0, 0]) and np.array_equal(e, [1, 1, 1, 0])
    c, e = host.cull_bounding_box_from_min_max((0.0, 2.0, -4.0), (2.0, 3.0, 0.0))
    assert np.array_equal(c[:3], [1.0, 2.5, -2.0]) and np.array_equal(e[:3], [1.0, 0.5, 2.0])


def test_cull_rate():
    s = host.OcclusionStats()
    assert s.cull_rate() == 0.0
    s.total_objects, s.after_occlusion = 100, 30
    assert abs(s.cull_rate() - 0.7) < 1e-9
    assert "30/100" in s.format()


def test_occlusion_culling_state():
    oc = host.OcclusionCulling(max_objects=3)
    ident = np.eye(4, dtype=np.float32)
    for i in range(5):                      # the 4th and 5th are dropped with a warning (add_object :233-238)
        oc.add_object(host.cull_bounding_box_from_center_extents((i, 0, 0), (0.5, 0.5, 0.5)), ident, i)
    assert oc.object_count() == 3
    d = oc.object_data()
    assert d.dtype.itemsize == 112 and np.all(d["flags"] == 1) and list(d["draw_index"]) == [0, 1, 2]
    assert np.array_equal(d["model"][1], ident)                   # columns, as to_cols_array_2d()
    pc = oc.push_constants(ident, 1920, 1080)
    assert pc["object_count"][0] == 3 and pc["hiz_levels"][0] == 10
    assert np.allclose(pc["screen_params"][0], [1920, 1080, 1 / 1920, 1 / 1080])
    oc.update_stats(1)
    assert oc.stats().draws_culled == 2 and oc.stats().after_occlusion == 1
    oc.begin_frame()
    assert oc.object_count() == 0 and oc.stats().total_objects == 0


# ---- features/light_culling.rs:257-276
def test_gpu_light_packing_and_tiles():
    g = host.gpu_light_from_point_light(host.PointLight((1.0, 2.0, 3.0), (1.0, 0.5, 0.25), 2.0, 10.0))
    assert g["position"][3] == 10.0 and g["color"][3] == 2.0 and g["params"][3] == 1.0
    d = host.gpu_light_from_directional_light(host.DirectionalLight())
    assert d["position"][3] == np.finfo(np.float32).max and d["direction"][3] == 2.0
    p = host.LightCullingPass()
    assert not p.is_enabled()
    p.calculate_tiles(1920, 1080)
    assert (p.tiles_x, p.tiles_y) == (120, 68) and p.get_dispatch_dimensions() == (120, 68, 1)
    p.update_lights([host.PointLight()] * 2000, [host.DirectionalLight()])
    assert p.light_count() == 1024 and p.is_enabled()             # MAX_LIGHTS cap
    assert p.get_push_constants(1920, 1080)["light_count"][0] == 1024
    big = host.LightCullingPass(max_lights=8192)
    big.update_lights([host.PointLight()] * 2000, [host.DirectionalLight()])
    assert big.light_count() == 2001 and big.get_light_buffer_data()["direction"][-1][3] == 2.0


def test_create_camera_data():
    view = np.eye(4, dtype=np.float32)
    from ash_renderer_b200 import synth
    proj = synth.perspective_rh(np.radians(45.0), 16 / 9, 0.5, 100.0).astype(np.float32)
    cd = host.create_camera_data(view, proj, (1.0, 2.0, 3.0))
    assert np.array_equal(cd["camera_pos"][0], [1, 2, 3, 1])       # light_culling_integration.rs:224-230
    prod = cd["inv_projection"][0].astype(np.float64).T @ proj.astype(np.float64).T
    assert np.allclose(prod, np.eye(4), atol=1e-5)


# ---- temporal_aa.rs:220-238
def test_temporal_aa():
    taa = host.TemporalAA()
    proj = np.eye(4, dtype=np.float32)
    assert np.array_equal(taa.jitter_projection(proj, 1920, 1080), proj)     # zero initial jitter
    taa.begin_frame()
    j1 = taa.current_jitter()
    taa.begin_frame()
    j2 = taa.current_jitter()
    assert j1 != j2 and all(-0.5 <= v <= 0.5 for v in (*j1, *j2))
    assert float(j1[0]) == 0.0 and abs(float(j1[1]) - (1 / 3 - 0.5)) < 1e-7  # halton(1,2)-.5, halton(1,3)-.5
    pc = taa.push_constants(1920, 1080)
    assert np.array_equal(pc["params"][0], np.array([0.9, 1, 1, 0], np.float32))
    assert tuple(pc["jitter"][0]) == j2 and tuple(pc["prev_jitter"][0]) == j1
    jp = taa.jitter_projection(proj, 1920, 1080)
    assert jp[3][0] == np.float32(j2[0]) * np.float32(2) / np.float32(1920)
    taa.reset_history()
    assert taa.current_jitter() == (0.0, 0.0) and taa.halton.index == 0
    off = host.TemporalAA(host.TaaConfig(enabled=False))
    off.begin_frame()
    assert np.array_equal(off.jitter_projection(proj, 64, 64), proj)


def test_post_configs_mirror_reference_defaults():
    b, t = host.BloomConfig(), host.TonemappingConfig()
    assert (b.threshold, b.intensity, b.mip_count, b.soft_knee) == (0.8, 0.5, 5, 0.5)     # bloom.rs:23-33
    assert (t.exposure, t.gamma, t.operator) == (1.0, 2.2, "Aces")                         # tonemapping.rs:35-44
    pc = t.push_constants(b.intensity)
    assert pc.dtype.itemsize == 16 and pc["bloom_intensity"][0] == 0.5


# ---- features/feature_trait.rs:36-98
def test_feature_manager_order_and_uniqueness():
    log = []

    class A(host.RenderFeature):
        def name(self): return "A"
        def on_added(self, d): log.append(("added", "A", d))
        def before_frame(self, c): log.append("A.before")
        def render(self, c): log.append("A.render")
        def on_removed(self, d): log.append("A.removed")

    class B(A):
        def name(self): return "B"
        def on_added(self, d): log.append(("added", "B", d))
        def before_frame(self, c): log.append("B.before")
        def render(self, c): log.append("B.render")
        def on_removed(self, d): log.append("B.removed")

    fm = host.FeatureManager()
    fm.add_feature(A())                      # no device yet: on_added not called
    fm.set_device("dev")
    fm.add_feature(B())
    fm.add_feature(A())                      # one instance per type
    fm.before_frame(None)
    fm.render(None)
    fm.cleanup()
    assert log == [("added", "B", "dev"), "A.before", "B.before", "A.render", "B.render", "A.removed", "B.removed"]


def test_cuda_feature_fails_loudly_before_on_added():
    from ash_renderer_b200 import backend
    f = host.CudaComputeFeature(64, 64)
    with pytest.raises(backend.AshError) as e:
        f.before_frame()
    assert e.value.code == backend.ASH_ERR_NOT_INITIALIZED


# ---- view sharding
def test_shard_views_partitions_exactly():
    for n, w in ((1024, 8), (1024, 3), (5, 8), (0, 2), (17, 4)):
        for mode in ("block", "cyclic"):
            owned = [views.shard_views(n, w, r, mode) for r in range(w)]
            flat = sorted(v for o in owned for v in o)
            assert flat == list(range(n))
            assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
            for r, o in enumerate(owned):
                assert all(views.owner_of(v, n, w, mode) == r for v in o)
    assert views.shard_views(1024, 8, 3) == list(range(384, 512))


def test_stats_wire_format_roundtrip():
    st = np.zeros(3, T.VIEW_STATS)
    st["view_id"] = [7, 0, 4000000000]
    st["visible_objects"] = [1, 2, 3]
    w = views.stats_to_words(st)
    assert w.shape == (3, 8) and w.dtype == np.int32
    assert np.array_equal(views.words_to_stats(w), st)


def test_all_gather_stats_gloo_world_size_2(tmp_path):
    """The N>1 host path on CPU: 2 ranks, gloo, uneven shards (5 views), ordered result on every rank."""
    script = tmp_path / "worker.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {ROOT!r})
        import numpy as np, torch.distributed as dist
        from ash_renderer_b200 import views, types as T
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        dist.init_process_group("gloo", rank=rank, world_size=world)
        mine = views.shard_views(5, world, rank)
        st = np.zeros(len(mine), T.VIEW_STATS)
        st["view_id"] = mine
        st["visible_objects"] = [100 + v for v in mine]
        st["total_objects"] = 1000
        allst = views.all_gather_stats(st, 5, world, rank)
        assert list(allst["view_id"]) == [0, 1, 2, 3, 4], allst
        assert list(allst["visible_objects"]) == [100, 101, 102, 103, 104]
        dist.barrier()
        dist.destroy_process_group()
        print("rank", rank, "ok", mine)
    """))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "[0, 1, 2]" in outs[0] and "[3, 4]" in outs[1]
