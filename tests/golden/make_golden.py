"""Regenerates tests/golden/ref_small.npz:  python -m tests.golden.make_golden   (needs /root/reference)

The vectors are outputs of the REFERENCE'S OWN SHADER SOURCES on small seeded inputs: oracle/_ref/libashref.so is
/root/reference/shaders/*.comp|*.frag compiled as C++ behind oracle/ref/glsl_compat.hpp (oracle/ref/Makefile).  The
file records the sha256 of every shader it was generated from.  tests/test_oracle.py checks the CPU oracle against
them, tests/test_parity_gpu.py replays them through the CUDA path; both run where /root/reference is absent.
(`tile_minmax` is the shader's shared-memory min/max, not an output: it comes from the oracle.)
"""
import hashlib
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


def compute(oracle, minmax_from=None):
    """`oracle`: oracle.oracle or oracle.ref (same call signatures)."""
    i = inputs()
    cam = i["cam"]
    mips, flat = oracle.hiz_build(i["depth"])
    pc = cam.cull_pc(len(i["objs"]), min(10, len(mips) + 1))
    vis, dout, cnt = oracle.occlusion_cull(i["objs"], i["draws"], pc, i["depth"], flat)
    res = oracle.light_cull(i["lights"], cam.light_pc(len(i["lights"])), cam.camera_data(), i["depth"])
    td = res[0]
    minmax = res[2] if len(res) > 2 else minmax_from.light_cull(i["lights"], cam.light_pc(len(i["lights"])),
                                                                 cam.camera_data(), i["depth"])[2]
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


SHADERS = ["hiz_generate.comp", "occlusion_cull.comp", "light_culling.comp", "taa_resolve.comp",
           "bloom_threshold.frag", "bloom_downsample.frag", "bloom_upsample.frag", "tonemapping.frag"]
META = ("shader_names", "shader_sha256", "generator")

if __name__ == "__main__":
    from oracle import oracle, ref
    out = compute(ref, minmax_from=oracle)
    out["shader_names"] = np.array(SHADERS)
    out["shader_sha256"] = np.array([hashlib.sha256(open(os.path.join(ref.REFERENCE_SHADERS, s), "rb").read()).hexdigest()
                                     for s in SHADERS])
    out["generator"] = np.array("oracle/_ref/libashref.so (reference GLSL compiled as C++, oracle/ref/)")
    np.savez_compressed(os.path.join(HERE, "ref_small.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_small.npz"))
