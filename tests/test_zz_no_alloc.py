"""ashcuda_cull_compact_static: the allocation-free variant of the ordered draw-command compaction (no counterpart in
the reference — include/ashcuda.h, csrc/cull_compact_static.cu).

Called through the C ABI on caller-owned device buffers and compared BIT-EXACT (integer outputs: the bar of every cull
test in this repository) with what the existing entry point, ashcuda_occlusion_cull, leaves in ASH_BUF_DRAWS_OUT /
ASH_BUF_VISIBLE_COUNT for the same visibility — on two calls in a row: the second, smaller call finds the first call's
contents in the static scratch array and must not be affected by them.  Collected last (file name), self-contained:
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from ash_renderer_b200 import synth, types as T

pytestmark = pytest.mark.gpu


def _scene(W, H, grid, seed, eye):
    cam = synth.Camera(eye, (0, 0, 0), W, H)
    objs, draws = synth.make_objects(grid, seed=seed, jitter=0.3, disabled_frac=0.05)
    objs["draw_index"][::7] = len(draws) + 3           # past the template buffer: zeros, as the existing kernel writes
    depth = synth.splat_depth(objs, cam)
    return cam, objs, draws, depth


def test_static_compaction_equals_occlusion_cull_on_two_calls_in_a_row(backend):
    import torch
    W, H = 640, 360
    # call 1: 27 000 objects (106 CTAs of scratch); call 2: 1 331 objects (6 CTAs) — stale entries 6.. stay in the array
    cases = [_scene(W, H, (30, 30, 30), 31, (0.0, 6.0, 60.0)), _scene(W, H, (11, 11, 11), 32, (3.0, 2.0, 24.0))]
    stream = torch.cuda.Stream()
    for call, (cam, objs, draws, depth) in enumerate(cases):
        n = len(objs)
        with backend.Context(W, H, max_objects=n, max_lights=16, enable_post=0) as c:
            c.set_depth(depth)
            c.set_objects(objs, draws)
            c.occlusion_cull(cam.cull_pc(n, 9))                      # the existing entry point
            vis, cnt = c.visibility(), c.visible_count()
            want = c.draws_out(cnt)
        assert 0 < cnt < n and (want["index_count"] == 0).any()
        # caller-owned device buffers, made BEFORE the call; sentinel-filled so that unwritten slots show
        d_objs = torch.from_numpy(objs.view(np.uint8)).cuda()
        d_draws = torch.from_numpy(draws.view(np.uint8)).cuda()
        d_vis = torch.from_numpy(vis.view(np.int32)).cuda()
        d_out = torch.full((n * 5,), -1, dtype=torch.int32, device="cuda")
        d_cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        backend.cull_compact_static(d_objs, d_draws, d_vis, n, d_out, d_cnt, device=0, stream=stream.cuda_stream)
        stream.synchronize()
        if call == 1:                                                # (the first call may load the module lazily)
            assert torch.cuda.mem_get_info()[0] >= free0, "the entry point allocated device memory"
        got_cnt = int(d_cnt.cpu().numpy().view(np.uint32)[0])
        got = d_out.cpu().numpy().view(np.uint32).reshape(n, 5)
        assert got_cnt == cnt, f"call {call}: visible_count {got_cnt} != {cnt}"
        assert np.array_equal(got[:cnt], want.view(np.uint32).reshape(cnt, 5)), f"call {call}: draw commands differ"
        assert np.all(got[cnt:] == 0xFFFFFFFF), f"call {call}: slots beyond visible_count were written"


def test_static_compaction_limits_and_empty_input(backend):
    import torch
    lib = backend.load_library()
    err = C.create_string_buffer(256)
    d_cnt = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    # beyond the static scratch: rejected with a clear message, nothing launched
    rc = lib.ashcuda_cull_compact_static(0, None, None, None, 0, None, backend.STATIC_COMPACT_MAX_OBJECTS + 1, None,
                                         C.c_void_p(d_cnt.data_ptr()), err, len(err))
    assert rc == backend.ASH_ERR_INVALID and b"ASH_STATIC_COMPACT_MAX_OBJECTS" in err.value
    torch.cuda.synchronize()
    assert int(d_cnt.item()) == 7
    with pytest.raises(backend.AshError) as e:
        backend.cull_compact_static(d_cnt, d_cnt, d_cnt, backend.STATIC_COMPACT_MAX_OBJECTS + 256, d_cnt, d_cnt)
    assert e.value.code == backend.ASH_ERR_INVALID
    # no objects: the count is stored as zero (PIN-8), on the default stream
    rc = lib.ashcuda_cull_compact_static(0, None, None, None, 0, None, 0, None, C.c_void_p(d_cnt.data_ptr()), err, len(err))
    torch.cuda.synchronize()
    assert rc == backend.ASH_OK and int(d_cnt.item()) == 0
    # the largest size the array holds: every object visible, identity order
    n = backend.STATIC_COMPACT_MAX_OBJECTS
    objs = np.zeros(n, T.CULL_OBJECT)
    objs["draw_index"] = np.arange(n) % 5
    draws = np.zeros(5, T.DRAW_COMMAND)
    draws["index_count"] = 10 + np.arange(5)
    d_objs, d_draws = torch.from_numpy(objs.view(np.uint8)).cuda(), torch.from_numpy(draws.view(np.uint8)).cuda()
    d_vis = torch.ones(n, dtype=torch.int32, device="cuda")
    d_out = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
    backend.cull_compact_static(d_objs, d_draws, d_vis, n, d_out, d_cnt)
    torch.cuda.synchronize()
    assert int(d_cnt.item()) == n
    assert np.array_equal(d_out.cpu().numpy().reshape(n, 5)[:, 0], 10 + np.arange(n) % 5)
