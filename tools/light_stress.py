"""tools/light_stress.py (GPU box): the light-cull kernels on the dense scene of synth.dense_light_scene (8192 lights, ~228 per
tile at 4K, overflow tiles) beside the bench scene's ~5 per tile: per-class kernel times from the library's own profiler."""
import json
import sys
import numpy as np
sys.path.insert(0, ".")
from ash_renderer_b200 import backend, synth, types as T

W, H = 3840, 2160
out = {}
for name in ("dense", "bench_like"):
    if name == "dense":
        cam, lights, cd, depth = synth.dense_light_scene(W, H)
    else:
        cam = synth.Camera((0.0, 6.0, 60.0), (0, 0, 0), W, H)
        objs, _ = synth.make_objects((12, 12, 12), seed=3)
        depth = synth.splat_depth(objs, cam)
        lights = synth.make_lights(8192, extent=40.0, seed=7)
        cd = cam.camera_data()
    pc = cam.light_pc(len(lights))
    with backend.Context(W, H, max_objects=1024, max_lights=8192, enable_post=0) as c:
        c.set_depth(depth); c.upload_lights(lights); c.upload_camera(cd); c.hiz_build()
        for _ in range(3):
            c.light_cull(pc)
        c.profile_kernels((1 << T.KERNEL_LIGHT_PREPASS) | (1 << T.KERNEL_LIGHT_TILES))
        c.kernel_time_ms(T.KERNEL_LIGHT_PREPASS); c.kernel_time_ms(T.KERNEL_LIGHT_TILES)
        n = 20
        for _ in range(n):
            c.light_cull(pc)
        pre, npre = c.kernel_time_ms(T.KERNEL_LIGHT_PREPASS)
        til, ntil = c.kernel_time_ms(T.KERNEL_LIGHT_TILES)
        st = c.stats()
        tiles = st["tiles"]
        out[name] = {"candidate_lights": int(st["candidate_lights"]) if "candidate_lights" in st.dtype.names else None, "lights_per_tile_mean": float(st["sum_tile_lights"]) / float(tiles),
                     "max_tile_lights": int(st["max_tile_lights"]), "overflow_tiles": int(st["overflow_tiles"]),
                     "prepass_and_cells_us": 1e3 * pre / max(npre, 1), "tiles_us": 1e3 * til / max(ntil, 1),
                     "launches": [int(npre), int(ntil)]}
print(json.dumps(out, indent=1))
