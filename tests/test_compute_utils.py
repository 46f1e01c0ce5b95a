"""Row f-4: shaders/compute_utils.glsl — the CPU restatement's properties (CPU) and the CUDA header against it (GPU)."""
import numpy as np
import pytest

from ash_renderer_b200 import synth
from oracle import oracle as O


def _data(n, seed):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) * 10 ** rng.uniform(-3, 3, n)).astype(np.float32)
    x[rng.integers(0, n, 5)] = 0.0
    x[rng.integers(0, n, 3)] = -0.0
    return x


def _camera():
    cam = synth.Camera((3.0, 2.0, 12.0), (0, 0, 0), 640, 360, near=0.5, far=100.0)
    return cam, cam.view_proj.astype(np.float32).reshape(-1), synth.cols_inverse(cam.proj).astype(np.float32).reshape(-1)


def _shapes(seed, n=4096):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-25, 25, (n, 3)).astype(np.float32)
    r = rng.uniform(0.05, 4.0, (n, 1)).astype(np.float32)
    return np.concatenate([c, r], 1), np.concatenate([c - r, c + r * rng.uniform(0.2, 1.0, (n, 3)).astype(np.float32)], 1)


def test_oracle_reductions_and_scan():
    x = _data(256 * 37, 1)
    g = x.reshape(-1, 256)
    assert np.array_equal(O.compute_utils(O.CU_MIN, x), g.min(1)) and np.array_equal(O.compute_utils(O.CU_MAX, x), g.max(1))
    s = O.compute_utils(O.CU_SUM, x)
    exact = g.astype(np.float64).sum(1)
    assert np.all(np.abs(s - exact) <= 8 * np.finfo(np.float32).eps * np.abs(g).astype(np.float64).sum(1))   # tree: log2(256) roundings
    # the tree pairs element i with i + stride: reproduce it with numpy float32 adds
    t = g.copy()
    for stride in (128, 64, 32, 16, 8, 4, 2, 1):
        t[:, :stride] = t[:, :stride] + t[:, stride:2 * stride]
    assert np.array_equal(s.view(np.uint32), t[:, 0].view(np.uint32))
    u = np.random.default_rng(2).integers(0, 2 ** 24, 256 * 9).astype(np.uint32)
    u[5] = 0xFFFFFFF0                                                    # wraps like uint
    ref = (np.cumsum(u.reshape(-1, 256).astype(np.uint64), 1) - u.reshape(-1, 256)).astype(np.uint32)
    assert np.array_equal(O.compute_utils(O.CU_SCAN, u).reshape(-1, 256), ref)


def test_oracle_frustum_and_depth_helpers():
    cam, vp, inv_proj = _camera()
    planes = O.compute_utils(O.CU_PLANES, params=vp).astype(np.float64)
    assert np.allclose(np.linalg.norm(planes[:, :3], axis=1), 1.0, atol=1e-6)
    m = vp.reshape(4, 4).astype(np.float64)                                               # m[c] = column c; the GLSL indexes COLUMNS (mat[3] + mat[0]),
    raw = np.stack([m[3] + m[0], m[3] - m[0], m[3] + m[1], m[3] - m[1], m[3] + m[2], m[3] - m[2]])   # like occlusion_cull.comp:82-100
    assert np.allclose(planes, raw / np.linalg.norm(raw[:, :3], axis=1, keepdims=True), rtol=1e-5, atol=1e-6)
    spheres, boxes = _shapes(3)
    vis = O.compute_utils(O.CU_SPHERE_IN_FRUSTUM, spheres, vp)
    d = spheres[:, :3].astype(np.float64) @ planes[:, :3].T + planes[:, 3]
    expect = (d >= -spheres[:, 3:4].astype(np.float64)).all(1)
    margin = np.abs(d + spheres[:, 3:4]).min(1) > 1e-4
    assert np.array_equal(vis[margin].astype(bool), expect[margin]) and 0.02 < vis.mean() < 0.9
    bvis = O.compute_utils(O.CU_AABB_IN_FRUSTUM, boxes, vp)
    pv = np.where(planes[:, None, :3] >= 0, boxes[None, :, 3:], boxes[None, :, :3])        # 6, n, 3
    bd = (pv * planes[:, None, :3]).sum(2) + planes[:, 3:4]
    bmargin = np.abs(bd).min(0) > 1e-4
    assert np.array_equal(bvis[bmargin].astype(bool), (bd >= 0).all(0)[bmargin])
    # depth helpers: project view-space points, then invert
    rng = np.random.default_rng(4)
    pts = np.stack([rng.uniform(-3, 3, 500), rng.uniform(-2, 2, 500), -rng.uniform(0.6, 90, 500)], 1)
    clip = np.concatenate([pts, np.ones((500, 1))], 1) @ cam.proj                           # cam.proj[c] = column c
    ndc = clip[:, :3] / clip[:, 3:4]
    uvd = np.stack([ndc[:, 0] * 0.5 + 0.5, ndc[:, 1] * 0.5 + 0.5, ndc[:, 2]], 1).astype(np.float32)
    out = O.compute_utils(O.CU_DEPTH, uvd, np.concatenate([inv_proj, [0.5, 100.0]]))
    assert np.allclose(out[:, 0], -pts[:, 2], rtol=2e-3)                                    # linearizeDepth = view distance
    assert np.allclose(out[:, 1:], pts, rtol=2e-3, atol=2e-3)                               # uvDepthToView


@pytest.mark.gpu
def test_cuda_header_bit_exact_against_the_oracle():
    from ash_renderer_b200 import backend as B
    x = _data(256 * 301, 7)
    for op in (B.CU_MIN, B.CU_MAX, B.CU_SUM):
        assert np.array_equal(B.compute_utils(op, x).view(np.uint32), O.compute_utils(op, x).view(np.uint32)), op
    u = np.random.default_rng(8).integers(0, 2 ** 31, 256 * 64).astype(np.uint32)
    assert np.array_equal(B.compute_utils(B.CU_SCAN, u), O.compute_utils(O.CU_SCAN, u))
    cam, vp, inv_proj = _camera()
    assert np.array_equal(B.compute_utils(B.CU_PLANES, params=vp).view(np.uint32), O.compute_utils(O.CU_PLANES, params=vp).view(np.uint32))
    spheres, boxes = _shapes(9, n=50000)
    assert np.array_equal(B.compute_utils(B.CU_SPHERE_IN_FRUSTUM, spheres, vp), O.compute_utils(O.CU_SPHERE_IN_FRUSTUM, spheres, vp))
    assert np.array_equal(B.compute_utils(B.CU_AABB_IN_FRUSTUM, boxes, vp), O.compute_utils(O.CU_AABB_IN_FRUSTUM, boxes, vp))
    rng = np.random.default_rng(10)
    uvd = rng.uniform(0, 1, (10000, 3)).astype(np.float32)
    p = np.concatenate([inv_proj, [0.5, 100.0]]).astype(np.float32)
    assert np.array_equal(B.compute_utils(B.CU_DEPTH, uvd, p).view(np.uint32), O.compute_utils(O.CU_DEPTH, uvd, p).view(np.uint32))
    with pytest.raises(B.AshError):
        B.compute_utils(B.CU_SUM, x[:100])                                                  # not a multiple of the workgroup
