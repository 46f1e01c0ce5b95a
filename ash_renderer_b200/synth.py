"""Synthetic inputs for the hot path (numpy only; used by tests, smoke and bench).

Rasterisation is out of scope, so depth and HDR colour are generated here: matrices follow glam
(`perspective_rh` with [0,1] depth, `look_at_rh`, examples/02_cube.rs:87-91 incl. the Y flip),
built in float64 and rounded once to fp32.  Recipes follow SURVEY.md §8.4.
"""
import math

import numpy as np

from . import types as T


# ------------------------------------------------------------------ matrices (column-major)
def perspective_rh(fov_y, aspect, near, far, flip_y=True):
    """glam::Mat4::perspective_rh (depth 0..1). Returns M with M[c] = column c (float64)."""
    h = math.cos(0.5 * fov_y) / math.sin(0.5 * fov_y)
    w = h / aspect
    r = far / (near - far)
    m = np.zeros((4, 4))
    m[0] = [w, 0, 0, 0]
    m[1] = [0, -h if flip_y else h, 0, 0]      # examples/02_cube.rs:89-91 flips Y
    m[2] = [0, 0, r, -1]
    m[3] = [0, 0, r * near, 0]
    return m


def look_at_rh(eye, center, up=(0.0, 1.0, 0.0)):
    eye, center, up = (np.asarray(v, np.float64) for v in (eye, center, up))
    f = center - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, up)
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.zeros((4, 4))
    m[0] = [s[0], u[0], -f[0], 0]
    m[1] = [s[1], u[1], -f[1], 0]
    m[2] = [s[2], u[2], -f[2], 0]
    m[3] = [-s.dot(eye), -u.dot(eye), f.dot(eye), 1]
    return m


def cols_mul(a, b):
    """(a*b) for column-array matrices (X[c] = column c)."""
    return (a.T @ b.T).T


def cols_inverse(a):
    return np.linalg.inv(a.T).T


class Camera:
    def __init__(self, eye, target, W, H, fov_deg=45.0, near=0.5, far=100.0, flip_y=True):
        self.eye = np.asarray(eye, np.float64)
        self.W, self.H = W, H
        self.view = look_at_rh(eye, target)
        self.proj = perspective_rh(math.radians(fov_deg), W / H, near, far, flip_y)
        self.view_proj = cols_mul(self.proj, self.view)       # renderer.rs:1645 projection*view

    def cull_pc(self, n_objects, hiz_levels):
        pc = np.zeros(1, T.CULLING_PC)
        pc["view_proj"][0] = self.view_proj.astype(np.float32)
        w, h = np.float32(self.W), np.float32(self.H)
        pc["screen_params"][0] = [w, h, np.float32(1.0) / w, np.float32(1.0) / h]   # occlusion_culling.rs:255-260
        pc["object_count"] = n_objects
        pc["hiz_levels"] = hiz_levels
        return pc

    def light_pc(self, n_lights):
        pc = np.zeros(1, T.LIGHT_PC)
        pc["screen_size"][0] = [self.W, self.H]
        pc["light_count"] = n_lights
        return pc

    def camera_data(self):
        """LightCullingIntegration::create_camera_data, light_culling_integration.rs:112-123."""
        cd = np.zeros(1, T.CAMERA_DATA)
        cd["view"][0] = self.view.astype(np.float32)
        cd["projection"][0] = self.proj.astype(np.float32)
        cd["inv_projection"][0] = cols_inverse(self.proj).astype(np.float32)
        cd["camera_pos"][0] = [*self.eye.astype(np.float32), 1.0]
        return cd


# ------------------------------------------------------------------ scene
def make_objects(grid, spacing=2.0, seed=2, jitter=0.0, disabled_frac=0.0, n_draws=8):
    """Instances on a grid: local AABB centre 0 extents 0.5, model = T * uniform scale U(.5,1.5)."""
    rng = np.random.default_rng(seed)
    gx, gy, gz = grid
    n = gx * gy * gz
    idx = np.stack(np.meshgrid(np.arange(gx), np.arange(gy), np.arange(gz), indexing="ij"), -1)
    pos = (idx.reshape(-1, 3) - (np.array(grid) - 1) / 2.0) * spacing
    if jitter:
        pos = pos + rng.uniform(-jitter, jitter, pos.shape)
    scale = rng.uniform(0.5, 1.5, n)
    objs = np.zeros(n, T.CULL_OBJECT)
    objs["extents"][:, :3] = 0.5
    m = np.zeros((n, 4, 4))
    m[:, 0, 0] = m[:, 1, 1] = m[:, 2, 2] = scale
    m[:, 3, :3] = pos
    m[:, 3, 3] = 1.0
    objs["model"] = m.astype(np.float32)
    objs["draw_index"] = np.arange(n) % n_draws
    objs["flags"] = 1
    if disabled_frac:
        objs["flags"][rng.random(n) < disabled_frac] = 0
    draws = np.zeros(n_draws, T.DRAW_COMMAND)
    draws["index_count"] = 36 * (1 + np.arange(n_draws))
    draws["instance_count"] = 1
    draws["first_index"] = 1000 * np.arange(n_draws)
    draws["vertex_offset"] = -7 * np.arange(n_draws)
    draws["first_instance"] = np.arange(n_draws)
    return objs, draws


def make_lights(n, extent=10.0, rmin=1.0, rmax=5.0, seed=2, n_directional=4, enabled_p=0.95,
                center=(0.0, 0.0, 0.0)):
    rng = np.random.default_rng(seed + 1000)
    lights = np.zeros(n, T.GPU_LIGHT)
    lights["position"][:, :3] = rng.uniform(-extent, extent, (n, 3)) + np.asarray(center)
    lights["position"][:, 3] = rng.uniform(rmin, rmax, n)
    lights["color"] = np.concatenate([rng.uniform(0, 1, (n, 3)), rng.uniform(0.5, 4, (n, 1))], 1)
    lights["params"] = [0.0, 0.0, 1.0, 1.0]                    # features/light_culling.rs:51
    lights["params"][rng.random(n) >= enabled_p, 3] = 0.0
    for i in rng.choice(n, min(n_directional, n), replace=False):   # from_directional_light :56-63
        lights["position"][i] = [0, 0, 0, np.finfo(np.float32).max]
        lights["direction"][i] = [0.3, -1.0, 0.2, 2.0]
    return lights


def dense_light_scene(W, H, n_lights=8192, n_near=8000, rmax=0.52, seed=21):
    """A view whose light lists are long: raw depth in [0, 1] bounds the z slab of light_culling.comp:96-97 to about one
    unit around the eye, so `n_near` lights are spread laterally INSIDE that slab (view space x, y, z in [-1.6, 1.6] x
    [-0.9, 0.9] x [-1.6, 0.3], radius up to `rmax`): ~228 lights per 16x16 tile at 4K, about a thousand tiles beyond 256."""
    rng = np.random.default_rng(seed)
    cam = Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, _ = make_objects((6, 6, 6), seed=3)
    depth = splat_depth(objs, cam)
    lights = make_lights(n_lights, extent=12.0, seed=5, center=(0, 0, 0))
    cd = cam.camera_data()
    vp = np.stack([rng.uniform(-1.6, 1.6, n_near), rng.uniform(-0.9, 0.9, n_near), rng.uniform(-1.6, 0.3, n_near), np.ones(n_near)], 1)
    wp = (np.linalg.inv(cd["view"][0].astype(np.float64).T) @ vp.T).T
    lights["position"][:n_near, :3] = wp[:, :3].astype(np.float32)
    lights["position"][:n_near, 3] = rng.uniform(0.01, rmax, n_near).astype(np.float32)
    return cam, lights, cd, depth


def splat_depth(objs, cam, shrink=0.5, background=1.0):
    """Coherent occluder z-buffer: each instance's projected AABB rectangle, shrunk about its
    centre, filled with the FARTHEST corner's ndc z over a `background` clear."""
    W, H = cam.W, cam.H
    depth = np.full((H, W), background, np.float32)
    n = len(objs)
    if n == 0:
        return depth
    model = objs["model"].astype(np.float64)
    c = objs["center"][:, :3].astype(np.float64)
    e = objs["extents"][:, :3].astype(np.float64)
    wc = np.einsum("nck,nc->nk", model[:, :3, :3], c) + model[:, 3, :3]
    we = e * np.linalg.norm(model[:, :3, :3], axis=2)
    sg = np.array([[sx, sy, sz] for sz in (-1, 1) for sy in (-1, 1) for sx in (-1, 1)], np.float64)
    corners = wc[:, None, :] + we[:, None, :] * sg[None]                       # n,8,3
    vp = cam.view_proj
    clip = corners @ vp[:3, :] + vp[3]                                         # n,8,4
    ok = (clip[..., 3] > 1e-6).all(1)
    w = np.where(clip[..., 3:4] > 1e-6, clip[..., 3:4], 1.0)
    ndc = clip[..., :3] / w
    sx = (ndc[..., 0] * 0.5 + 0.5) * W
    sy = (ndc[..., 1] * 0.5 + 0.5) * H
    x0, x1, y0, y1 = sx.min(1), sx.max(1), sy.min(1), sy.max(1)
    cx, cy, hx, hy = (x0 + x1) / 2, (y0 + y1) / 2, (x1 - x0) / 2 * shrink, (y1 - y0) / 2 * shrink
    zf = np.clip(ndc[..., 2].max(1), 0.0, 1.0).astype(np.float32)
    ix0 = np.clip(np.floor(cx - hx), 0, W).astype(int)
    ix1 = np.clip(np.ceil(cx + hx), 0, W).astype(int)
    iy0 = np.clip(np.floor(cy - hy), 0, H).astype(int)
    iy1 = np.clip(np.ceil(cy + hy), 0, H).astype(int)
    for i in np.nonzero(ok & (ix1 > ix0) & (iy1 > iy0) & (zf < background))[0]:
        blk = depth[iy0[i]:iy1[i], ix0[i]:ix1[i]]
        np.minimum(blk, zf[i], out=blk)
    return depth


def random_depth(W, H, seed=0, blocks=32):
    """Piecewise-smooth random depth in [0,1] (blocky occluders + noise) for kernel-level tests."""
    rng = np.random.default_rng(seed)
    by, bx = (H + blocks - 1) // blocks, (W + blocks - 1) // blocks
    coarse = rng.uniform(0.2, 1.0, (by, bx)).astype(np.float32)
    d = np.kron(coarse, np.ones((blocks, blocks), np.float32))[:H, :W]
    d = d - rng.uniform(0, 0.05, (H, W)).astype(np.float32)
    d[rng.random((H, W)) < 0.01] = 1.0
    return np.ascontiguousarray(np.clip(d, 0.0, 1.0).astype(np.float32))


# ------------------------------------------------------------------ HDR colour frames
class HdrSequence:
    """Procedural HDR scene F(x,y) sampled at jittered pixel centres: a smooth
    base in [0,2] plus emissive discs with radiance U(2,30). frame(jx,jy) -> (H,W,4) float16."""

    def __init__(self, W, H, seed=4, n_discs=200):
        rng = np.random.default_rng(seed)
        self.W, self.H = W, H
        self.waves = rng.uniform(0.002, 0.05, (6, 2)) * rng.choice([-1, 1], (6, 2))
        self.phase = rng.uniform(0, 2 * np.pi, (6, 3))
        self.dc = rng.uniform(0, 1, (n_discs, 2)) * [W, H]
        self.dr = rng.uniform(2.0, max(3.0, 0.02 * H), n_discs)
        self.dl = rng.uniform(2.0, 30.0, (n_discs, 1)) * rng.uniform(0.3, 1.0, (n_discs, 3))

    def frame(self, jx=0.0, jy=0.0):
        W, H = self.W, self.H
        x = (np.arange(W, dtype=np.float32) + np.float32(0.5 + jx))[None, :]
        y = (np.arange(H, dtype=np.float32) + np.float32(0.5 + jy))[:, None]
        img = np.ones((H, W, 3), np.float32)
        for k in range(6):
            arg = self.waves[k, 0] * x + self.waves[k, 1] * y
            for c in range(3):
                img[..., c] += np.float32(1.0 / 6.0) * np.sin(arg + np.float32(self.phase[k, c]))
        for i in range(len(self.dr)):
            r = self.dr[i]
            x0, x1 = max(0, int(self.dc[i, 0] - r - 2)), min(W, int(self.dc[i, 0] + r + 3))
            y0, y1 = max(0, int(self.dc[i, 1] - r - 2)), min(H, int(self.dc[i, 1] + r + 3))
            if x1 <= x0 or y1 <= y0:
                continue
            d = np.sqrt((x[:, x0:x1] - self.dc[i, 0]) ** 2 + (y[y0:y1] - self.dc[i, 1]) ** 2)
            a = np.clip(r + 0.5 - d, 0.0, 1.0).astype(np.float32)[..., None]
            img[y0:y1, x0:x1] += a * self.dl[i].astype(np.float32)
        out = np.empty((H, W, 4), np.float16)
        out[..., :3] = np.clip(img, 0.0, 60000.0)
        out[..., 3] = 1.0
        return out


def random_hdr(W, H, seed=0, hot_frac=0.02):
    """White-noise HDR (worst case for neighbourhood clamping) with a few very bright texels."""
    rng = np.random.default_rng(seed)
    img = rng.uniform(0.0, 2.0, (H, W, 4)).astype(np.float32)
    hot = rng.random((H, W)) < hot_frac
    img[hot, :3] *= rng.uniform(2.0, 30.0, (int(hot.sum()), 1)).astype(np.float32)
    img[..., 3] = 1.0
    return img.astype(np.float16)


def random_velocity(W, H, seed=0, still_frac=0.5, max_px=3.0):
    """RG16F velocity in UV units; `still_frac` of texels are exactly zero (jitter branch)."""
    rng = np.random.default_rng(seed)
    v = rng.uniform(-max_px, max_px, (H, W, 2)) / np.array([W, H])
    v[rng.random((H, W)) < still_frac] = 0.0
    return v.astype(np.float16)


# ------------------------------------------------------------------ row f-2: instances + LOD meshes
def make_lod_meshes(n_meshes=6, seed=7):
    """LodMesh records with 1..8 levels, descending screen thresholds / ascending distance thresholds."""
    rng = np.random.default_rng(seed)
    meshes = np.zeros(n_meshes, T.LOD_MESH)
    for m in range(n_meshes):
        n = 1 + (m * 3) % T.MAX_LOD_LEVELS if m else 1
        th = np.sort(rng.uniform(0.0005, 0.2, n))[::-1].copy()
        th[-1] = 0.0 if m % 2 == 0 else th[-1]
        dist = np.sort(rng.uniform(5.0, 120.0, n))
        tri = (20000 // (2 ** np.arange(n))).astype(np.uint32)
        for l in range(n):
            meshes[m]["levels"][l] = (th[l], dist[l], tri[l], tri[l] // 2 + 3, 100000 * m + 3 * int(tri[:l].sum()), 3 * tri[l])
        meshes[m]["level_count"] = n
        meshes[m]["bounding_radius"] = rng.uniform(0.3, 2.5)
        meshes[m]["lod_bias"] = [0.0, 0.0, 1.0, -1.0, 0.4, 2.6][m % 6]
    return meshes


def make_instances(objs, n_batches=5, n_meshes=6, seed=8):
    """One InstanceData per cull object (same model columns); a batch id and a LodMesh id per object
    (every object of a batch draws the same mesh, as a BatchKey names one mesh)."""
    rng = np.random.default_rng(seed)
    n = len(objs)
    inst = np.zeros(n, T.INSTANCE_DATA)
    m = objs["model"]
    inst["model_row0"], inst["model_row1"], inst["model_row2"], inst["model_row3"] = m[:, 0], m[:, 1], m[:, 2], m[:, 3]
    inst["color"] = rng.uniform(0, 1, (n, 4)).astype(np.float32)
    inst["custom"][:, 0] = np.arange(n)
    batch = rng.integers(0, n_batches, n).astype(np.uint32)
    mesh_of_batch = (np.arange(n_batches) * 7 + 1) % n_meshes
    return inst, batch, mesh_of_batch[batch].astype(np.uint32)


def instancing_params(cam, n):
    ip = np.zeros(1, T.INSTANCING_PARAMS)
    ip["view_proj"][0] = cam.view_proj.astype(np.float32)
    ip["camera_pos"][0] = [*cam.eye.astype(np.float32), 1.0]
    ip["screen_size"][0] = [cam.W, cam.H]
    ip["object_count"] = n
    return ip


def reprojection(cam, prev_cam):
    """prevViewProj * inverse(view) as columns (float64 -> fp32 once): maps a view-space point of `cam` to the previous
    frame's clip space (ashcuda_set_forward_reprojection)."""
    return cols_mul(prev_cam.view_proj, cols_inverse(cam.view)).astype(np.float32)


def forward_params(light_direction=(-0.3, -1.0, -0.2), light_color=(1.0, 0.95, 0.9), ambient=(0.03, 0.03, 0.04),
                   base_color=(0.8, 0.8, 0.8), emissive=(0.0, 0.0, 0.0), metallic=0.1, roughness=0.5,
                   background=(0.02, 0.02, 0.03), near=0.5, radius_scale=1.0, flags=T.FWD_TILE_LIGHTS):
    """AshForwardParams: the frag.frag MVP / Material uniform fields the compute forward pass (row f-1) uses."""
    p = np.zeros(1, T.FORWARD_PARAMS)
    p["light_direction"][0, :3] = light_direction
    p["light_color"][0] = [*light_color, 1.0]
    p["ambient_color"][0] = [*ambient, 1.0]
    p["base_color_factor"][0] = [*base_color, 1.0]
    p["emissive_factor"][0] = [*emissive, 1.0]
    p["parameters"][0] = [metallic, roughness, 1.0, 1.0]
    p["background"][0] = [*background, 1.0]
    p["near_plane"], p["radius_scale"], p["flags"] = near, radius_scale, flags
    return p

