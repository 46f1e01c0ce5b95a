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
#include <cstring>
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

// taa_resolve.comp:92-108 without a velocity buffer: historyUV = clamp(uv - jitterDelta / size, 0, 1) on one axis.
// "Regular": every p samples texels clamp(p + d), clamp(p + d + 1) with ONE weight f (where both taps are the same texel
// the weight does not matter).  True for almost every jitter delta; false when the delta sits so close to a multiple of
// 1/256 texel (or to 0) that the fp32 noise of (p + 0.5) / n * n moves single columns to the neighbouring weight / texel.
bool axis_taa_history_regular(int n, float size_f, float jitter_delta, int* d_out, float* f_out) {
    if (jitter_delta == 0.0f && axis_center_exact(n, size_f)) {     // historyUV == uv and that tap IS texel p: weight 0 on p + 1
        *d_out = 0; *f_out = 0.0f;
        return true;
    }
    const float vel = fdiv(jitter_delta, size_f);
    int d = 0; float f0 = 0.0f; bool have = false;
    for (int pass = 0; pass < 2; ++pass)
        for (int p = 0; p < n; ++p) {
            const float u = fdiv(fadd((float)p, 0.5f), size_f);
            const float hu = gclamp(fsub(u, vel), 0.0f, 1.0f);
            int i0, i1; float f;
            bilinear_setup(hu, n, i0, i1, f);
            if (pass == 0) { if (i0 != i1) { d = i0 - p; f0 = f; have = true; break; } }
            else if (i0 != clampi(p + d, n) || i1 != clampi(p + d + 1, n) || (i0 != i1 && f != f0)) return false;
        }
    if (!have && n > 1) return false;
    if (d < -1 || d > 0) return false;                     // taps must stay inside the 1-texel halo of a tile
    *d_out = d; *f_out = f0;
    return true;
}

// the blend factor of a frame without velocity buffer (:145-150): velocity = jitter delta / size, uniform
float taa_uniform_blend(const AshTaaPushConstants& pc) {
    float blend = pc.params[0];
    if (pc.params[2] > 0.5f) {
        const float sw = pc.screen_params[0], sh = pc.screen_params[1];
        const float vx = fdiv(fsub(pc.jitter[0], pc.prev_jitter[0]), sw);
        const float vy = fdiv(fsub(pc.jitter[1], pc.prev_jitter[1]), sh);
        const float vl = fsqrt(fadd(fmul(fmul(vx, sw), fmul(vx, sw)), fmul(fmul(vy, sh), fmul(vy, sh))));
        blend = fmul(blend, gclamp(fsub(1.0f, fmul(vl, 0.1f)), 0.0f, 1.0f));
    }
    return blend;
}

}  // namespace ash

using namespace ash;

// precision = FAST: the argument block of k_taa_tol for these push constants; the O(W + H) regularity check of the history
// taps runs once per distinct (screenParams, jitter, prevJitter) and is cached (a Halton period has 16 of them)
const TaaTolArgs& ash_taa_tol_plan(ashcuda_ctx* c, const AshTaaPushConstants& pc, float threshold, TaaTolArgs* scratch) {
    for (TaaTolArgs& p : c->tol_plans)
        if (memcmp(&p.pc, &pc, sizeof(pc)) == 0) { *scratch = p; scratch->threshold = threshold; return *scratch; }
    TaaTolArgs a{};
    a.pc = pc;
    a.blend = taa_uniform_blend(pc);
    const bool rx = axis_taa_history_regular((int)c->W, pc.screen_params[0], fsub(pc.jitter[0], pc.prev_jitter[0]), &a.dx, &a.fx);
    const bool ry = axis_taa_history_regular((int)c->H, pc.screen_params[1], fsub(pc.jitter[1], pc.prev_jitter[1]), &a.dy, &a.fy);
    a.regular = (rx && ry) ? 1u : 0u;
    if (c->tol_plans.size() >= 64) c->tol_plans.erase(c->tol_plans.begin());
    c->tol_plans.push_back(a);
    *scratch = a; scratch->threshold = threshold;
    return *scratch;
}

