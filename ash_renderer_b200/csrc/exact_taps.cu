// exact_taps.cu — host-side validation of the sampler shortcuts the fast post kernels rely on.
//
// The pinned LINEAR sampler (PIN-6: 8-bit sub-texel weights) lands EXACTLY on known texels with
// known weights whenever a tap's ideal coordinate is a multiple of 1/4 texel and the fp32
// evaluation error of the coordinate stays below 1/512 texel.  That holds for the extents that
// matter (<= 4096, even halving), but instead of trusting an error bound the library evaluates
// the very same expressions (ash_math.cuh, compiled for the host without FP contraction) for
// every row/column of an extent once, and only then lets a fast kernel replace the general one.
// O(W + H) per check; results are cached in the context.
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

static inline int clampi(int v, int n) { return v < 0 ? 0 : (v > n - 1 ? n - 1 : v); }

// a tap that must reproduce texel `e` bit-exactly (weights 0/1, or both texels identical)
static inline bool tap_is_texel(float u, int n, int e) {
    int i0, i1; float f;
    bilinear_setup(u, n, i0, i1, f);
    return (i0 == e && i1 == e) || (i0 == e && f == 0.0f) || (i1 == e && f == 1.0f);
}

// u = (x + 0.5) / size : taa_resolve.comp:92, postprocess.vert:12 -> texel x
bool axis_center_exact(int n, float size_f) {
    for (int x = 0; x < n; ++x)
        if (!tap_is_texel(fdiv(fadd((float)x, 0.5f), size_f), n, x)) return false;
    return true;
}

// taa_resolve.comp:118: uv + vec2(d) * screenParams.zw, d in {-1,0,1} -> texel clamp(x + d)
bool axis_taa_taps_exact(int n, float size_f, float inv) {
    for (int x = 0; x < n; ++x) {
        const float u = fdiv(fadd((float)x, 0.5f), size_f);
        for (int d = -1; d <= 1; ++d)
            if (!tap_is_texel(fadd(u, fmul((float)d, inv)), n, clampi(x + d, n))) return false;
    }
    return true;
}

// bloom_downsample.frag:22-34 with texelSize = 1/src: taps u, u -+ d  (dst n = src/2)
bool axis_down_taps_exact(int sn, int dn) {
    if (sn != 2 * dn) return false;
    const float d = fdiv(1.0f, (float)sn);
    for (int x = 0; x < dn; ++x) {
        const float u = fdiv(fadd((float)x, 0.5f), (float)dn);
        for (int k = -1; k <= 1; ++k) {
            const float uu = fadd(u, k < 0 ? -d : (k > 0 ? d : 0.0f));
            int i0, i1; float f;
            bilinear_setup(uu, sn, i0, i1, f);
            if (i0 != clampi(2 * x + k, sn) || i1 != clampi(2 * x + k + 1, sn) || f != 0.5f) return false;
        }
    }
    return true;
}

// bloom_upsample.frag:18-29 with d = texelSize*0.5: taps u + {-2d, -d, 0, d, 2d}  (dst n = 2*src)
bool axis_up_taps_exact(int sn, int dn) {
    if (dn != 2 * sn) return false;
    const float d = fmul(fdiv(1.0f, (float)sn), 0.5f);
    for (int x = 0; x < dn; ++x) {
        const float u = fdiv(fadd((float)x, 0.5f), (float)dn);
        const float off[5] = {fmul(-d, 2.0f), -d, 0.0f, d, fmul(d, 2.0f)};
        for (int m = -2; m <= 2; ++m) {
            const int q = 2 * x - 1 + 2 * m;                     // 4 * ideal source coordinate
            const int i = (q >= 0) ? q / 4 : -((-q + 3) / 4);    // floor(q / 4)
            const float g = (float)(q - 4 * i) * 0.25f;          // 0.25 or 0.75
            int i0, i1; float f;
            bilinear_setup(fadd(u, off[m + 2]), sn, i0, i1, f);
            if (i0 != clampi(i, sn) || i1 != clampi(i + 1, sn) || f != g) return false;
        }
    }
    return true;
}

}  // namespace ash
