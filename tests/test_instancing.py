"""Row f-2 on the CPU: the instancing / LOD oracle (oracle/ash_oracle_instancing.cpp) and the host mirror
(ash_renderer_b200/host.py) against the reference's own unit tests (instancing.rs:279-334,
lod_system.rs:396-428), and against each other on random scenes."""
import numpy as np
import pytest

from ash_renderer_b200 import host, synth, types as T


# ---- instancing.rs:283-289 test_instance_data
def test_instance_data_position():
    model = np.eye(4, dtype=np.float32)
    model[3][:3] = [1.0, 2.0, 3.0]                       # Mat4::from_translation, columns
    inst = host.instance_data_from_matrix(model)
    assert inst.dtype.itemsize == 96
    assert host.instance_position(inst) == (1.0, 2.0, 3.0)
    assert tuple(inst["color"]) == (1.0, 1.0, 1.0, 1.0) and not inst["custom"].any()


# ---- instancing.rs:291-305 test_batching, :307-320 test_multiple_batches
def test_batching_and_multiple_batches():
    m = host.InstancingManager()
    m.begin_frame()
    key = host.BatchKey(1, 1)
    for i in range(100):
        model = np.eye(4, dtype=np.float32)
        model[3][0] = i
        m.add_instance(key, host.instance_data_from_matrix(model))
    m.finalize()
    assert m.stats().total_instances == 100 and m.stats().batch_count == 1
    assert m.stats().avg_instances_per_batch == 100.0 and m.batch_ids().tolist() == [0] * 100
    m = host.InstancingManager()
    m.begin_frame()
    for mesh_id in range(5):
        m.add_instance(host.BatchKey(mesh_id, 0), np.zeros((), T.INSTANCE_DATA))
    m.finalize()
    assert m.stats().batch_count == 5 and m.batch_ids().tolist() == [0, 1, 2, 3, 4]
    m.begin_frame()                                       # batches are cleared, keys keep their ids
    m.add_instance(host.BatchKey(3, 0), np.zeros((), T.INSTANCE_DATA))
    m.finalize()
    assert m.stats().batch_count == 1 and m.batch_ids().tolist() == [3]


def test_batch_cap_is_max_instances_per_batch():
    m = host.InstancingManager()
    m.begin_frame()
    z = np.zeros((), T.INSTANCE_DATA)
    m._counts, m._keys = [host.MAX_INSTANCES_PER_BATCH - 1], {host.BatchKey(0, 0): 0}
    m.add_instance(host.BatchKey(0, 0), z)
    m.add_instance(host.BatchKey(0, 0), z)                # dropped, instancing.rs:219-223
    assert m.stats().total_instances == 1


# ---- lod_system.rs:400-404 test_single_lod, :406-415 test_screen_coverage, :417-425 test_reduction_rate
def test_reference_lod_unit_tests(oracle):
    mesh = host.LodMesh.single(1000, 1.0)
    assert mesh.level_count() == 1 and mesh.levels[0].index_count == 3000
    cam = synth.Camera((0.0, 0.0, 10.0), (0, 0, 0), 1920, 1080, fov_deg=45.0, near=0.1, far=100.0, flip_y=False)
    vp = cam.view_proj.astype(np.float32)
    cov_host = host.LodManager.calculate_screen_coverage((0.0, 0.0, 0.0), 1.0, vp, 1920.0, 1080.0)
    cov_oracle = oracle.screen_coverage(vp, (0.0, 0.0, 0.0), 1.0, 1920.0, 1080.0)
    assert 0.0 < cov_oracle < 1.0
    assert np.float32(cov_host) == np.float32(cov_oracle)
    # closed form: clip.w = 10, pixel radius = 0.1 * 1920 / 2 = 96, area = pi * 96^2 / (1920*1080)
    assert abs(cov_oracle - np.pi * 96.0 ** 2 / (1920 * 1080)) < 1e-6
    assert oracle.screen_coverage(vp, (0.0, 0.0, 20.0), 1.0, 1920.0, 1080.0) == 0.0        # behind the camera
    assert oracle.screen_coverage(vp, (0.0, 0.0, 9.9), 5.0, 1920.0, 1080.0) == 1.0         # clamped
    st = host.LodStats()
    st.triangles_before, st.triangles_after = 1000, 250
    assert abs(st.reduction_rate() - 0.75) < 0.001


def _scene(seed, n_batches=5, grid=(9, 8, 7)):
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), 1920, 1080)
    objs, _ = synth.make_objects(grid, seed=seed, jitter=0.3)
    meshes = synth.make_lod_meshes(6, seed=seed + 1)
    inst, batch, mesh = synth.make_instances(objs, n_batches, 6, seed=seed + 2)
    return cam, inst, batch, mesh, meshes


@pytest.mark.parametrize("mode", ["ScreenSize", "Distance", ("Force", 2), ("Force", 99)])
def test_oracle_agrees_with_the_scalar_host_manager(oracle, mode):
    cam, inst, batch, mesh, meshes = _scene(3)
    n = len(inst)
    cfg = host.LodConfig(mode=mode, global_bias=0.25)
    ip = synth.instancing_params(cam, n)
    sel, out, offsets, draws, stats = oracle.instancing(inst, batch, mesh, meshes, ip, cfg.record(use_visibility=False), 5)
    mgr = host.LodManager(cfg)
    mgr.begin_frame()
    vp = cam.view_proj.astype(np.float32)
    eye = cam.eye.astype(np.float32)
    kept = {}
    for i in range(n):
        rec = meshes[mesh[i]]
        lm = host.LodMesh([host.LodLevel(*[x.item() for x in rec["levels"][l]]) for l in range(rec["level_count"])], float(rec["bounding_radius"]))
        lm.lod_bias = float(rec["lod_bias"])
        s = mgr.select_lod(lm, host.instance_position(inst[i]), vp, eye, 1920.0, 1080.0)
        assert (s.level, bool(s.culled)) == (int(sel[i]["level"]), bool(sel[i]["culled"])), i
        assert np.float32(s.blend) == sel[i]["blend"] and np.float32(s.screen_coverage) == sel[i]["screen_coverage"], i
        if not s.culled:
            kept.setdefault(int(batch[i]) * 8 + s.level, []).append(i)
    hs = mgr.stats()
    assert hs.objects_per_lod == stats["objects_per_lod"].tolist() and hs.objects_culled == stats["objects_culled"]
    assert hs.triangles_before == stats["triangles_before"] and hs.triangles_after == stats["triangles_after"]
    # streams: ascending object index inside each bucket, buckets back to back
    for b in range(40):
        ids = out[offsets[b]:offsets[b + 1]]["custom"][:, 0].astype(int).tolist()
        assert ids == kept.get(b, []), b
        assert draws[b]["instance_count"] == len(ids) and draws[b]["first_instance"] == offsets[b]
    assert stats["total_instances"] == offsets[-1] == sum(len(v) for v in kept.values())
    assert stats["batch_count"] == len(kept)
    assert stats["avg_instances_per_batch"] == np.float32(offsets[-1]) / np.float32(len(kept))


def test_oracle_draw_ranges_visibility_and_cap(oracle):
    cam, inst, batch, mesh, meshes = _scene(5)
    n = len(inst)
    vis = (np.arange(n) % 3 != 0).astype(np.uint32)
    cfg = host.LodConfig().record(use_visibility=True)
    sel, out, offsets, draws, stats = oracle.instancing(inst, batch, mesh, meshes, synth.instancing_params(cam, n), cfg, 5, vis)
    assert np.all(sel["level"][vis == 0] == T.LOD_NOT_VISIBLE) and np.all(sel["level"][vis == 1] != T.LOD_NOT_VISIBLE)
    assert not np.isin(out["custom"][:, 0].astype(int), np.nonzero(vis == 0)[0]).any()
    for b in np.nonzero(draws["instance_count"])[0]:
        m, l = meshes[mesh[np.nonzero(batch == b // 8)[0][0]]], b % 8
        assert draws[b]["index_count"] == m["levels"][l]["index_count"] and draws[b]["first_index"] == m["levels"][l]["index_offset"]
    assert stats["objects_culled"] + stats["objects_per_lod"].sum() == vis.sum()
    # MAX_INSTANCES_PER_BATCH: 70 000 identical instances in one bucket keep the first 65 536
    big = np.repeat(inst[:1], 70000)
    big["custom"][:, 0] = np.arange(70000)
    one = np.zeros(70000, np.uint32)
    single = meshes[:1].copy()
    single[0]["level_count"], single[0]["levels"][0]["screen_threshold"] = 1, 0.0
    k = host.LodConfig(cull_threshold=0.0).record(use_visibility=False)
    sel, out, offsets, draws, stats = oracle.instancing(big, one, one, single, synth.instancing_params(cam, 70000), k, 1)
    assert stats["total_instances"] == 65536 and stats["instances_dropped"] == 70000 - 65536 and stats["batch_count"] == 1
    assert np.array_equal(out["custom"][:, 0], np.arange(65536, dtype=np.float32))
