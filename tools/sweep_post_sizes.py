"""One-off GPU sweep: bloom chain and TAA (+threshold) bit-exact against the oracle at sizes that exercise the
edges of the tiled kernels (partial tiles, second lane out of range, halo clamps).  python tools/sweep_post_sizes.py"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ash_renderer_b200 import backend, synth, types as T
from oracle import oracle

def taa_pc(W, H, jitter, prev):
    pc = np.zeros(1, T.TAA_PC); w, h = np.float32(W), np.float32(H)
    pc["screen_params"][0] = [w, h, np.float32(1) / w, np.float32(1) / h]
    pc["params"][0] = [0.9, 1.0, 1.0, 0.0]; pc["jitter"][0], pc["prev_jitter"][0] = jitter, prev
    return pc

SIZES = [(2048, 64, 2), (64, 2048, 2), (4096, 40, 2), (1280, 720, 4), (2560, 1440, 3), (3840, 2160, 5), (1056, 1000, 3),
         (2112, 520, 3), (1088, 1928, 3), (96, 80, 2), (3904, 72, 2)]


def run(sizes=SIZES, verbose=True):
    bad = 0
    for W, H, mips in sizes:
        hdr = synth.random_hdr(W, H, seed=W + H)
        U, D = oracle.bloom_chain(hdr, 0.8, mips)
        bp = np.zeros(1, T.BLOOM_PARAMS); bp["threshold"], bp["upsample_intensity"], bp["mip_count"] = 0.8, 1.0, mips
        with backend.Context(W, H, bloom_mips=mips, max_objects=16, max_lights=16) as c:
            c.set_color(hdr, None)
            c.bloom(bp)
            ok = all(np.array_equal(np.ascontiguousarray(c.bloom_mip(k)).view(np.uint16), np.ascontiguousarray(U[k]).view(np.uint16)) for k in range(mips + 1))
            hist = synth.random_hdr(W, H, seed=W + H + 1)
            ok_t = True
            for jit, prev in (((0.25, -0.1666), (0.0, 0.0)), ((-0.4, 0.3), (0.45, -0.35)), ((0.0, 0.0), (0.0, 0.0))):
                pc = taa_pc(W, H, jit, prev)
                c.set_color(hist, None); c.taa_resolve(taa_pc(W, H, (0, 0), (0, 0)))      # history := something resolved
                h_dev = np.ascontiguousarray(c.taa_out()).copy()
                c.set_color(hdr, None); c.taa_resolve(pc)
                ref = oracle.taa_resolve(hdr, h_dev, None, pc)
                ok_t &= np.array_equal(np.ascontiguousarray(c.taa_out()).view(np.uint16), np.ascontiguousarray(ref).view(np.uint16))
        if verbose: print(f"{W}x{H} mips={mips}: bloom {'ok' if ok else 'MISMATCH'}, taa {'ok' if ok_t else 'MISMATCH'}", flush=True)
        bad += (not ok) + (not ok_t)
    return bad


if __name__ == "__main__":
    n_bad = run()
    print("FAILED" if n_bad else "all sizes bit-exact")
    sys.exit(1 if n_bad else 0)
