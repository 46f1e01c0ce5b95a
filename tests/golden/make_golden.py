"""Regenerates tests/golden/oracle_small.npz:  python -m tests.golden.make_golden

The reference ships no golden vectors for this path and cannot be executed here (GLSL under
Vulkan; no graphics stack, no cargo), so these are REGRESSION vectors of the CPU oracle on small
seeded inputs: they pin the oracle against drift, not against the reference ("parity unpinned").
The GPU parity tests also replay them through the CUDA path (tests/test_parity_gpu.py).
"""
import os

import numpy as np

from ash_renderer_b200 import synth, types as T

HERE = os.path.dirname(os.path.abspath(__file__))
W, H = 200, 120


def inputs():
    cam = synth.Camera((0.0, 2.0, 22.0), (0, 0, 0), W, H)
    objs, draws = synth.make_objects((6, 6, 6), seed=21, jitter=0.3, disabled_frac=0.1)
    depth = synth.splat_depth(objs, cam)
    lights = synth.make_lights(300, extent=10.0, seed=21, center=(0.0, 2.0, 22.0))
    hdr = synth.random_hdr(W, H, seed=21)
    hist = synth.random_hdr(W, H, seed=22)
    vel = synth.random_velocity(W, H, seed=23)
    taa_pc = np.zeros(1, T.TAA_PC)
    w, h = np.float32(W), np.float32(H)
    taa_pc["screen_params"][0] = [w, h, np.float32(1) / w, np.float32(1) / h]
    taa_pc["params"][0] = [0.9, 1.0, 1.0, 0.0]
    taa_pc["jitter"][0], taa_pc["prev_jitter"][0] = (0.25, -0.16666667), (0.0, 0.0)
    post_pc = np.zeros(1, T.POST_PC)
    post_pc["exposure"], post_pc["gamma"], post_pc["bloom_intensity"] = 1.0, 2.2, 0.5
    return dict(cam=cam, objs=objs, draws=draws, depth=depth, lights=lights, hdr=hdr, hist=hist, vel=vel,
                taa_pc=taa_pc, post_pc=post_pc)


def compute(oracle):
    i = inputs()
    cam = i["cam"]
    mips, flat = oracle.hiz_build(i["depth"])
    pc = cam.cull_pc(len(i["objs"]), min(10, len(mips) + 1))
    vis, dout, cnt = oracle.occlusion_cull(i["objs"], i["draws"], pc, i["depth"], flat)
    td, true_counts, minmax = oracle.light_cull(i["lights"], cam.light_pc(len(i["lights"])), cam.camera_data(), i["depth"])
    counts = td[:, 0].copy()
    lists = np.concatenate([td[t, 1:1 + counts[t]] for t in range(len(td))])
    taa = oracle.taa_resolve(i["hdr"], i["hist"], i["vel"], i["taa_pc"])
    U, D = oracle.bloom_chain(taa, 0.8, 5)
    ldr = oracle.tonemap(taa, U[0], i["post_pc"])
    out = {"hiz_flat": flat, "visibility": vis, "visible_count": np.array([cnt], np.uint32),
           "draws_out": dout.view(np.uint32), "tile_counts": counts, "tile_lists": lists,
           "tile_minmax": minmax, "taa": taa.view(np.uint16), "bloom_u0": U[0].view(np.uint16),
           "ldr8": ldr}
    for k in range(1, 6):
        out[f"bloom_d{k}"] = D[k].view(np.uint16)
    return out


if __name__ == "__main__":
    from oracle import oracle
    np.savez_compressed(os.path.join(HERE, "oracle_small.npz"), **compute(oracle))
    print("wrote", os.path.join(HERE, "oracle_small.npz"))
