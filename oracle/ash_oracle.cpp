// ash_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// A scalar C++ restatement of the reference's GLSL for the per-frame compute path, used only by
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the
// checker and the timed CPU baseline.  Nothing under ash_renderer_b200/ may call into this file.
//
// PARITY PINNED TO THE REFERENCE'S SHADER TEXT: the reference (saptak7777/Ash-Renderer 0.3.8) holds no golden
// vectors or CPU implementation for these shaders and cannot run under Vulkan here, but its GLSL sources compile
// as C++ behind oracle/ref/glsl_compat.hpp (oracle/_ref/libashref.so, built from /root/reference/shaders where
// they lie).  tests/test_oracle_ref.py requires this restatement to equal that library bit for bit on all eight
// shaders, and tests/golden/ref_small.npz holds its outputs.  What the GLSL leaves to the implementation
// (sampler filtering, evaluation order, image stores) is fixed by the PIN table of SURVEY.md §8.0 / DESIGN.md §2 and
// is the same in both; each function cites the shader lines it follows (paths relative to the reference repo).
//
// Build: g++ -O2 -std=c++17 -fopenmp -ffp-contract=off -fno-fast-math -shared -fPIC  (oracle/Makefile)
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include "../include/ashcuda.h"

// ------------------------------------------------------------------------------------------
// PIN-1: GLSL built-ins with a fixed evaluation order (left to right, no FMA contraction).
// ------------------------------------------------------------------------------------------
namespace {

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct mat4 { vec4 c[4]; };  // c[i] = column i  (occlusion_culling.rs:96-102 uploads columns)

inline float gmin(float a, float b) { return (b < a) ? b : a; }   // GLSL min(x,y) = y<x ? y : x
inline float gmax(float a, float b) { return (a < b) ? b : a; }   // GLSL max(x,y) = x<y ? y : x
inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
inline float gsign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator+(vec3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
inline vec3 vmin(vec3 a, vec3 b) { return {gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)}; }
inline vec3 vmax(vec3 a, vec3 b) { return {gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)}; }

inline float dot3(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length3(vec3 v) { return sqrtf(dot3(v, v)); }
inline float length2(float x, float y) { return sqrtf(x * x + y * y); }
inline vec3 normalize3(vec3 v) { float l = length3(v); return {v.x / l, v.y / l, v.z / l}; }
// GLSL spec: cross(x,y) = (x1*y2 - y1*x2, x2*y0 - y2*x0, x0*y1 - y0*x1)
inline vec3 cross3(vec3 a, vec3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline vec3 xyz(vec4 v) { return {v.x, v.y, v.z}; }

// m * v = m[0]*v.x + m[1]*v.y + m[2]*v.z + m[3]*v.w, summed left to right
inline vec4 mul(const mat4& m, vec4 v) {
    vec4 r;
    r.x = ((m.c[0].x * v.x + m.c[1].x * v.y) + m.c[2].x * v.z) + m.c[3].x * v.w;
    r.y = ((m.c[0].y * v.x + m.c[1].y * v.y) + m.c[2].y * v.z) + m.c[3].y * v.w;
    r.z = ((m.c[0].z * v.x + m.c[1].z * v.y) + m.c[2].z * v.z) + m.c[3].z * v.w;
    r.w = ((m.c[0].w * v.x + m.c[1].w * v.y) + m.c[2].w * v.z) + m.c[3].w * v.w;
    return r;
}
inline mat4 load_mat4(const float m[4][4]) {
    mat4 r;
    for (int c = 0; c < 4; ++c) r.c[c] = {m[c][0], m[c][1], m[c][2], m[c][3]};
    return r;
}

// ------------------------------------------------------------------------------------------
// fp16 storage (PIN-6: RGBA16F targets, round-to-nearest-even on store)
// ------------------------------------------------------------------------------------------
inline float h2f(uint16_t h) {
    uint32_t s = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, bits;
    if (e == 0) {
        if (m == 0) bits = s;
        else { int sh = 0; while (!(m & 0x400u)) { m <<= 1; ++sh; }
               bits = s | ((uint32_t)(113 - sh) << 23) | ((m & 0x3FFu) << 13); }
    } else if (e == 31) bits = s | 0x7F800000u | (m << 13);
    else bits = s | ((e + 112u) << 23) | (m << 13);
    float f; memcpy(&f, &bits, 4); return f;
}
inline uint16_t f2h(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t s = (x >> 16) & 0x8000u, a = x & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? 0x200u : 0u));
    if (a >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);            // rounds to inf
    if (a < 0x33000001u) return (uint16_t)s;                          // rounds to zero
    int e = (int)(a >> 23) - 127; uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;
    uint32_t hm = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (hm & 1u))) ++hm;
    uint32_t he = (e < -14) ? 0u : (uint32_t)(e + 15);
    // hm carries its implicit bit for normals (bit 10): (he-1)<<10 + hm handles mantissa overflow
    uint32_t out = (e < -14) ? hm : (((he - 1u) << 10) + hm);
    return (uint16_t)(s | out);
}

// ------------------------------------------------------------------------------------------
// Samplers.
// PIN-5: depth / Hi-Z: NEAREST, CLAMP_TO_EDGE: texel = clamp(int(floor(u*W)), 0, W-1).
// PIN-6: colour: LINEAR, CLAMP_TO_EDGE, unnormalised x = u*W - 0.5, i0 = floor(x), weight
//        f = x - i0 quantised to 8 fractional bits (round to nearest) as texture units do
//        (Vulkan subTexelPrecisionBits = 8); lerp horizontally then vertically, a*(1-f)+b*f.
// ------------------------------------------------------------------------------------------
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

inline int nearest_texel(float u, int W) {
    float t = floorf(u * (float)W);
    if (!(t > -1.0f)) return 0;               // also catches NaN
    if (t >= (float)W) return W - 1;
    return (int)t;
}

// RGBA16F (4 halfs per texel) or, in the fp32_intermediates mode of PIN-6, RGBA32F (4 floats per texel, f32 = 1)
struct Img16 { const uint16_t* p; int W, H; int f32 = 0; };

inline vec4 fetch16(const Img16& im, int x, int y) {
    if (im.f32) {
        const float* t = reinterpret_cast<const float*>(im.p) + ((size_t)y * im.W + x) * 4;
        return {t[0], t[1], t[2], t[3]};
    }
    const uint16_t* t = im.p + ((size_t)y * im.W + x) * 4;
    return {h2f(t[0]), h2f(t[1]), h2f(t[2]), h2f(t[3])};
}
int g_exact_weights = 0;   // ash_oracle_set_weight_mode: 1 = SURVEY PIN-6 exact fp32 weights, 0 = 8-bit
inline void bilinear_setup(float u, int W, int& i0, int& i1, float& f) {
    float x = u * (float)W - 0.5f;
    float fl = floorf(x);
    float fr = x - fl;
    float fq = floorf(fr * 256.0f + 0.5f) / 256.0f;   // 8-bit sub-texel weight
    int i = (int)fl;
    i0 = clampi(i, 0, W - 1);
    i1 = clampi(i + 1, 0, W - 1);
    f = g_exact_weights ? fr : fq;
}
inline float lerp1(float a, float b, float f) { return a * (1.0f - f) + b * f; }
inline vec4 texture16(const Img16& im, float u, float v) {
    int x0, x1, y0, y1; float fx, fy;
    bilinear_setup(u, im.W, x0, x1, fx);
    bilinear_setup(v, im.H, y0, y1, fy);
    vec4 a = fetch16(im, x0, y0), b = fetch16(im, x1, y0);
    vec4 c = fetch16(im, x0, y1), d = fetch16(im, x1, y1);
    vec4 top = {lerp1(a.x, b.x, fx), lerp1(a.y, b.y, fx), lerp1(a.z, b.z, fx), lerp1(a.w, b.w, fx)};
    vec4 bot = {lerp1(c.x, d.x, fx), lerp1(c.y, d.y, fx), lerp1(c.z, d.z, fx), lerp1(c.w, d.w, fx)};
    return {lerp1(top.x, bot.x, fy), lerp1(top.y, bot.y, fy), lerp1(top.z, bot.z, fy),
            lerp1(top.w, bot.w, fy)};
}
inline void store16(uint16_t* img, int W, int x, int y, vec3 rgb, float a, int f32 = 0) {
    if (f32) {
        float* t = reinterpret_cast<float*>(img) + ((size_t)y * W + x) * 4;
        t[0] = rgb.x; t[1] = rgb.y; t[2] = rgb.z; t[3] = a;
        return;
    }
    uint16_t* t = img + ((size_t)y * W + x) * 4;
    t[0] = f2h(rgb.x); t[1] = f2h(rgb.y); t[2] = f2h(rgb.z); t[3] = f2h(a);
}

}  // namespace

extern "C" {

// ==========================================================================================
// Hi-Z pyramid — shaders/hiz_generate.comp:23-41, one call per mip (PIN-3/4/5).
// src is mip L-1 (sw x sh), dst is mip L (dw x dh), dw = max(1, sw/2), dh = max(1, sh/2).
// textureGather at uv=(2*pos+1)/srcSize selects texels (2x, 2x+1) x (2y, 2y+1), edge-clamped;
// gather component order is (i0,j1), (i1,j1), (i1,j0), (i0,j0) (hiz_generate.comp:33-37).
// ==========================================================================================
void ash_oracle_hiz_level(const float* src, int sw, int sh, float* dst, int dw, int dh) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < dh; ++y) {
        for (int x = 0; x < dw; ++x) {
            int i0 = clampi(2 * x, 0, sw - 1), i1 = clampi(2 * x + 1, 0, sw - 1);
            int j0 = clampi(2 * y, 0, sh - 1), j1 = clampi(2 * y + 1, 0, sh - 1);
            float gx = src[(size_t)j1 * sw + i0], gy = src[(size_t)j1 * sw + i1];
            float gz = src[(size_t)j0 * sw + i1], gw = src[(size_t)j0 * sw + i0];
            dst[(size_t)y * dw + x] = gmax(gmax(gx, gy), gmax(gz, gw));   // :38
        }
    }
}

// levels_built = floor(log2(max(W,H))) + 1 (PIN-4)
int ash_oracle_hiz_level_count(int W, int H) {
    int m = W > H ? W : H, n = 1;
    while (m > 1) { m >>= 1; ++n; }
    return n;
}

// Builds mips 1..levels-1 into `mips`, concatenated in level order; returns texel count written.
size_t ash_oracle_hiz_build(const float* depth, int W, int H, float* mips) {
    int levels = ash_oracle_hiz_level_count(W, H);
    const float* src = depth; int sw = W, sh = H; size_t off = 0;
    for (int l = 1; l < levels; ++l) {
        int dw = sw / 2 > 0 ? sw / 2 : 1, dh = sh / 2 > 0 ? sh / 2 : 1;
        ash_oracle_hiz_level(src, sw, sh, mips + off, dw, dh);
        src = mips + off; off += (size_t)dw * dh; sw = dw; sh = dh;
    }
    return off;
}

// ==========================================================================================
// Occlusion cull — shaders/occlusion_cull.comp:82-233.
// `pyramid` = mip 0 (the depth buffer) followed by mips 1.. exactly as ash_oracle_hiz_build
// lays them out after `depth` (caller passes depth and mips separately).
// Outputs follow PIN-7 (ascending object index) and PIN-8 (only the valid prefix is written).
// ==========================================================================================
static void extract_frustum_planes(const mat4& vp, vec4 planes[6]) {   // :82-100
    const vec4 &c0 = vp.c[0], &c1 = vp.c[1], &c2 = vp.c[2], &c3 = vp.c[3];
    planes[0] = {c3.x + c0.x, c3.y + c0.y, c3.z + c0.z, c3.w + c0.w};
    planes[1] = {c3.x - c0.x, c3.y - c0.y, c3.z - c0.z, c3.w - c0.w};
    planes[2] = {c3.x + c1.x, c3.y + c1.y, c3.z + c1.z, c3.w + c1.w};
    planes[3] = {c3.x - c1.x, c3.y - c1.y, c3.z - c1.z, c3.w - c1.w};
    planes[4] = {c3.x + c2.x, c3.y + c2.y, c3.z + c2.z, c3.w + c2.w};
    planes[5] = {c3.x - c2.x, c3.y - c2.y, c3.z - c2.z, c3.w - c2.w};
    for (int i = 0; i < 6; ++i) {
        float l = length3(xyz(planes[i]));
        planes[i] = {planes[i].x / l, planes[i].y / l, planes[i].z / l, planes[i].w / l};
    }
}

static bool frustum_cull(vec3 center, vec3 extents, const vec4 planes[6]) {   // :103-116
    for (int i = 0; i < 6; ++i) {
        vec3 n = xyz(planes[i]);
        float dist = planes[i].w;
        vec3 sg = {gsign(n.x), gsign(n.y), gsign(n.z)};
        vec3 pv = center + extents * sg;
        if (dot3(n, pv) + dist < 0.0f) return true;
    }
    return false;
}

// exact ceil(log2(s)) clamped to [0, hi] (never call log2f)
static int mip_from_size(float s, int hi) {
    if (!(s > 1.0f)) return 0;                       // log2(s) <= 0 (also s == 0, NaN)
    uint32_t b; memcpy(&b, &s, 4);
    if ((b >> 23) == 0xFFu) return hi;               // +inf
    int e = (int)(b >> 23) - 127;
    int k = (b & 0x7FFFFFu) ? e + 1 : e;
    return k > hi ? hi : k;
}

extern "C" int ash_oracle_hiz_level_count(int W, int H);
struct Pyramid { const float* depth; const float* mips; int W, H; };
static float pyramid_fetch(const Pyramid& p, int level, float u, float v) {
    int w = p.W, h = p.H; const float* base = p.depth; size_t off = 0;
    const int last = ash_oracle_hiz_level_count(p.W, p.H) - 1;     // textureLod clamps to the image's last mip
    if (level > last) level = last;
    for (int l = 1; l <= level; ++l) {
        int nw = w / 2 > 0 ? w / 2 : 1, nh = h / 2 > 0 ? h / 2 : 1;
        if (l > 1) off += (size_t)w * h;
        w = nw; h = nh; base = p.mips + off;
    }
    return base[(size_t)nearest_texel(v, h) * w + nearest_texel(u, w)];
}

static bool occlusion_cull(vec3 center, vec3 e, const mat4& vp, const float screen[4],
                           uint32_t hiz_levels, const Pyramid& pyr) {   // :123-178
    vec3 corners[8] = {
        center + vec3{-e.x, -e.y, -e.z}, center + vec3{e.x, -e.y, -e.z},
        center + vec3{e.x, e.y, -e.z},   center + vec3{-e.x, e.y, -e.z},
        center + vec3{-e.x, -e.y, e.z},  center + vec3{e.x, -e.y, e.z},
        center + vec3{e.x, e.y, e.z},    center + vec3{-e.x, e.y, e.z}};
    float minx = 1.0f, miny = 1.0f, maxx = 0.0f, maxy = 0.0f, minz = 1.0f;
    for (int i = 0; i < 8; ++i) {
        vec4 clip = mul(vp, {corners[i].x, corners[i].y, corners[i].z, 1.0f});
        if (clip.w <= 0.0f) return false;                       // :145-147
        float nx = clip.x / clip.w, ny = clip.y / clip.w, nz = clip.z / clip.w;
        float sx = nx * 0.5f + 0.5f, sy = ny * 0.5f + 0.5f;
        minx = gmin(minx, sx); miny = gmin(miny, sy);
        maxx = gmax(maxx, sx); maxy = gmax(maxy, sy);
        minz = gmin(minz, nz);
    }
    minx = gclamp(minx, 0.0f, 1.0f); miny = gclamp(miny, 0.0f, 1.0f);
    maxx = gclamp(maxx, 0.0f, 1.0f); maxy = gclamp(maxy, 0.0f, 1.0f);
    float ssx = (maxx - minx) * screen[0], ssy = (maxy - miny) * screen[1];
    int level = mip_from_size(gmax(ssx, ssy), (int)(hiz_levels - 1u));   // :162-164
    float u = (minx + maxx) * 0.5f, v = (miny + maxy) * 0.5f;              // :170
    float hiz = pyramid_fetch(pyr, level, u, v);
    return minz > hiz;                                                     // :173
}

// returns visibleCount
// n_draws: size of drawCommandsIn; a drawIndex beyond it reads zeros (Vulkan robustBufferAccess)
uint32_t ash_oracle_occlusion_cull_n(const AshCullObjectData* objs, const AshIndirectDrawCommand* draws_in, uint32_t n_draws,
                                     const AshCullingPushConstants* pc, const float* depth,
                                     const float* mips, int W, int H,
                                     uint32_t* visibility, AshIndirectDrawCommand* draws_out) {
    mat4 vp = load_mat4(pc->view_proj);
    vec4 planes[6];
    extract_frustum_planes(vp, planes);            // same for every invocation (:207-208)
    Pyramid pyr{depth, mips, W, H};
    const int n = (int)pc->object_count;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {                   // main, :184-233
        const AshCullObjectData& o = objs[i];
        if ((o.flags & 1u) == 0u) { visibility[i] = 0u; continue; }
        mat4 model = load_mat4(o.model);
        vec3 wc = xyz(mul(model, {o.bounds.center[0], o.bounds.center[1], o.bounds.center[2], 1.0f}));
        vec3 ext = {o.bounds.extents[0], o.bounds.extents[1], o.bounds.extents[2]};
        vec3 we = ext * vec3{length3(xyz(model.c[0])), length3(xyz(model.c[1])), length3(xyz(model.c[2]))};
        if (frustum_cull(wc, we, planes)) { visibility[i] = 0u; continue; }
        if (occlusion_cull(wc, we, vp, pc->screen_params, pc->hiz_levels, pyr)) { visibility[i] = 0u; continue; }
        visibility[i] = 1u;
    }
    uint32_t count = 0;                              // PIN-7: compaction in ascending index order
    for (int i = 0; i < n; ++i)
        if (visibility[i]) {
            if (draws_out) draws_out[count] = objs[i].draw_index < n_draws ? draws_in[objs[i].draw_index] : AshIndirectDrawCommand{};
            ++count;
        }
    return count;
}
uint32_t ash_oracle_occlusion_cull(const AshCullObjectData* objs, const AshIndirectDrawCommand* draws_in,
                                   const AshCullingPushConstants* pc, const float* depth, const float* mips, int W, int H,
                                   uint32_t* visibility, AshIndirectDrawCommand* draws_out) {
    return ash_oracle_occlusion_cull_n(objs, draws_in, 0xFFFFFFFFu, pc, depth, mips, W, H, visibility, draws_out);
}

// ==========================================================================================
// Tiled light culling — shaders/light_culling.comp:64-203, 16x16 tiles (spec-constant defaults,
// vulkan/light_culling_pipeline.rs:219-223 passes none).  PIN-7: ascending light index, first
// 256 kept on overflow.  PIN-8: only tile_data[off .. off+1+count) is written.
// true_counts (optional, tiles entries) receives the un-clamped sharedLightCount.
// ==========================================================================================
static vec3 screen_to_view(const mat4& inv_proj, float u, float v, float depth) {   // :64-68
    vec4 clip = {u * 2.0f - 1.0f, v * 2.0f - 1.0f, depth, 1.0f};
    vec4 view = mul(inv_proj, clip);
    return {view.x / view.w, view.y / view.w, view.z / view.w};
}

static void tile_frustum(const mat4& inv_proj, uint32_t tx, uint32_t ty, float sw, float sh,
                         float min_depth, float max_depth, vec4 planes[6]) {      // :71-98
    const float ts = 16.0f;
    float min_x = ((float)tx * ts) / sw * 2.0f - 1.0f, min_y = ((float)ty * ts) / sh * 2.0f - 1.0f;
    float max_x = ((float)(tx + 1u) * ts) / sw * 2.0f - 1.0f, max_y = ((float)(ty + 1u) * ts) / sh * 2.0f - 1.0f;
    vec3 c[4];
    c[0] = screen_to_view(inv_proj, min_x * 0.5f + 0.5f, min_y * 0.5f + 0.5f, 0.0f);
    c[1] = screen_to_view(inv_proj, max_x * 0.5f + 0.5f, min_y * 0.5f + 0.5f, 0.0f);
    c[2] = screen_to_view(inv_proj, max_x * 0.5f + 0.5f, max_y * 0.5f + 0.5f, 0.0f);
    c[3] = screen_to_view(inv_proj, min_x * 0.5f + 0.5f, max_y * 0.5f + 0.5f, 0.0f);
    vec3 v[4] = {cross3(c[3], c[0]), cross3(c[1], c[2]), cross3(c[0], c[1]), cross3(c[2], c[3])};
    for (int i = 0; i < 4; ++i) {
        vec3 n = (length3(v[i]) > 0.0001f) ? normalize3(v[i]) : vec3{0.0f, 0.0f, 0.0f};
        planes[i] = {n.x, n.y, n.z, 0.0f};
    }
    planes[4] = {0.0f, 0.0f, -1.0f, -min_depth};    // raw [0,1] depth used as a view distance (:96)
    planes[5] = {0.0f, 0.0f, 1.0f, max_depth};      // (:97)
}

static bool sphere_in_frustum(vec3 c, float radius, const vec4 planes[6]) {        // :101-109
    for (int i = 0; i < 6; ++i) {
        float dist = dot3(xyz(planes[i]), c) + planes[i].w;
        if (dist < -radius) return false;
    }
    return true;
}

void ash_oracle_light_cull(const AshGpuLight* lights, const AshLightCullingPushConstants* pc,
                           const AshCullingCameraData* cam, const float* depth,
                           uint32_t* tile_data, uint32_t* true_counts, uint32_t* tile_minmax) {
    const uint32_t W = pc->screen_size[0], H = pc->screen_size[1], L = pc->light_count;
    const uint32_t tiles_x = (W + 15u) / 16u, tiles_y = (H + 15u) / 16u;         // :117
    const mat4 view = load_mat4(cam->view), inv_proj = load_mat4(cam->inv_projection);
    // (camera.view * vec4(pos,1)).xyz is tile-independent: evaluate once per light (:181)
    std::vector<vec4> lv(L);
    for (uint32_t i = 0; i < L; ++i) {
        vec4 p = mul(view, {lights[i].position[0], lights[i].position[1], lights[i].position[2], 1.0f});
        lv[i] = {p.x, p.y, p.z, lights[i].position[3]};
    }
    const int ntiles = (int)(tiles_x * tiles_y);
#pragma omp parallel for schedule(dynamic, 16)
    for (int t = 0; t < ntiles; ++t) {
        const uint32_t tx = (uint32_t)t % tiles_x, ty = (uint32_t)t / tiles_x;
        uint32_t mn = 0xFFFFFFFFu, mx = 0u;                                       // :124-125
        for (uint32_t py = ty * 16u; py < ty * 16u + 16u && py < H; ++py)
            for (uint32_t px = tx * 16u; px < tx * 16u + 16u && px < W; ++px) {
                uint32_t b; memcpy(&b, &depth[(size_t)py * W + px], 4);           // :133-137
                mn = b < mn ? b : mn; mx = b > mx ? b : mx;
            }
        if (tile_minmax) { tile_minmax[2 * t] = mn; tile_minmax[2 * t + 1] = mx; }
        float min_depth, max_depth; memcpy(&min_depth, &mn, 4); memcpy(&max_depth, &mx, 4);
        uint32_t* out = tile_data + (size_t)t * ASH_TILE_STRIDE;                  // :119
        if (L == 0u) { out[0] = 0u; if (true_counts) true_counts[t] = 0u; continue; }   // :160-165
        vec4 planes[6];
        tile_frustum(inv_proj, tx, ty, (float)W, (float)H, min_depth, max_depth, planes);
        uint32_t count = 0;
        for (uint32_t i = 0; i < L; ++i) {                                        // :174-191
            if (lights[i].params[3] < 0.5f) continue;
            if (sphere_in_frustum({lv[i].x, lv[i].y, lv[i].z}, lv[i].w, planes)) {
                if (count < ASH_MAX_LIGHTS_PER_TILE) out[1 + count] = i;
                ++count;
            }
        }
        if (true_counts) true_counts[t] = count;
        out[0] = count < ASH_MAX_LIGHTS_PER_TILE ? count : ASH_MAX_LIGHTS_PER_TILE;   // :196-197
    }
}

// ==========================================================================================
// TAA resolve — shaders/taa_resolve.comp:34-167.  All images RGBA16F (W x H); velocity is
// RG16F (2 halfs per texel) or NULL (PIN-12: absent == all-zero == jitter-delta branch).
// ==========================================================================================
static inline vec3 rgb_to_ycocg(vec3 c) {                        // :34-40
    return {(0.25f * c.x + 0.5f * c.y) + 0.25f * c.z,
            0.5f * c.x - 0.5f * c.z,
            (-0.25f * c.x + 0.5f * c.y) - 0.25f * c.z};
}
static inline vec3 ycocg_to_rgb(vec3 c) {                        // :43-52
    return {(c.x + c.y) - c.z, c.x + c.z, (c.x - c.y) - c.z};
}
static inline vec3 taa_tonemap(vec3 c) { return c / (vec3{1.0f, 1.0f, 1.0f} + c); }          // :55-57
static inline vec3 taa_untonemap(vec3 c) {                                                  // :59-61
    return c / ((vec3{1.0f, 1.0f, 1.0f} - c) + 0.0001f);
}
static inline vec3 clip_aabb(vec3 mn, vec3 mx, vec3 hist) {      // :64-77 (ignores `color`)
    vec3 center = 0.5f * (mx + mn);
    vec3 half = 0.5f * (mx - mn) + 0.001f;
    vec3 delta = hist - center;
    vec3 v = delta / half;
    float maxv = gmax(gmax(fabsf(v.x), fabsf(v.y)), fabsf(v.z));
    if (maxv > 1.0f) return center + delta / maxv;
    return hist;
}

// fmt bits (fp32_intermediates mode, PIN-6): 1 = `current` is RGBA32F, 2 = `history` is, 4 = `output` is
void ash_oracle_taa_resolve_fmt(const uint16_t* current, const uint16_t* history, const uint16_t* velocity,
                                const AshTaaPushConstants* pc, uint16_t* output, int fmt) {
    const float sw = pc->screen_params[0], sh = pc->screen_params[1];
    const int W = (int)sw, H = (int)sh;
    const Img16 cur{current, W, H, fmt & 1}, hist{history, W, H, (fmt >> 1) & 1};
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            float u = ((float)x + 0.5f) / sw, v = ((float)y + 0.5f) / sh;         // :92
            float vx = 0.0f, vy = 0.0f;
            if (velocity) {   // RG16F, LINEAR sampler at a texel centre == that texel (PIN-6)
                int x0, x1, y0, y1; float fx, fy;
                bilinear_setup(u, W, x0, x1, fx); bilinear_setup(v, H, y0, y1, fy);
                auto V = [&](int xi, int yi, int c) { return h2f(velocity[((size_t)yi * W + xi) * 2 + c]); };
                vx = lerp1(lerp1(V(x0, y0, 0), V(x1, y0, 0), fx), lerp1(V(x0, y1, 0), V(x1, y1, 0), fx), fy);
                vy = lerp1(lerp1(V(x0, y0, 1), V(x1, y0, 1), fx), lerp1(V(x0, y1, 1), V(x1, y1, 1), fx), fy);
            }
            if (length2(vx, vy) < 0.0001f) {                                      // :97-99
                vx = (pc->jitter[0] - pc->prev_jitter[0]) / sw;
                vy = (pc->jitter[1] - pc->prev_jitter[1]) / sh;
            }
            float hu = gclamp(u - vx, 0.0f, 1.0f), hv = gclamp(v - vy, 0.0f, 1.0f);   // :102
            vec3 c = xyz(texture16(cur, u, v));                                   // :105
            vec3 h = xyz(texture16(hist, hu, hv));                                // :108
            if (pc->params[1] > 0.5f) {                                           // :112-142
                vec3 mn{}, mx{};
                for (int dy = -1, idx = 0; dy <= 1; ++dy)
                    for (int dx = -1; dx <= 1; ++dx, ++idx) {
                        float su = u + (float)dx * pc->screen_params[2];
                        float sv = v + (float)dy * pc->screen_params[3];
                        vec3 s = rgb_to_ycocg(taa_tonemap(xyz(texture16(cur, su, sv))));
                        if (idx == 0) { mn = s; mx = s; } else { mn = vmin(mn, s); mx = vmax(mx, s); }
                    }
                vec3 ctr = (mn + mx) * 0.5f, ext = (mx - mn) * 0.5f;              // :132-135
                mn = ctr - ext * 1.25f; mx = ctr + ext * 1.25f;
                vec3 hy = rgb_to_ycocg(taa_tonemap(h));
                h = taa_untonemap(ycocg_to_rgb(clip_aabb(mn, mx, hy)));           // :140-141
            }
            float blend = pc->params[0];                                          // :145-150
            if (pc->params[2] > 0.5f) {
                float vl = length2(vx * sw, vy * sh);
                blend *= gclamp(1.0f - vl * 0.1f, 0.0f, 1.0f);
            }
            if (hu < 0.0f || hu > 1.0f || hv < 0.0f || hv > 1.0f) blend = 0.0f;   // :153-156 (dead)
            vec3 r = c * (1.0f - blend) + h * blend;                              // :159 mix()
            store16(output, W, x, y, r, 1.0f, (fmt >> 2) & 1);                    // :162
        }
    }
}
void ash_oracle_taa_resolve(const uint16_t* current, const uint16_t* history, const uint16_t* velocity,
                            const AshTaaPushConstants* pc, uint16_t* output) {
    ash_oracle_taa_resolve_fmt(current, history, velocity, pc, output, 0);
}

// HaltonSequence::halton — src/renderer/temporal_aa.rs:60-69
float ash_oracle_halton(uint32_t index, uint32_t base) {
    float f = 1.0f, r = 0.0f;
    while (index > 0) { f /= (float)base; r += f * (float)(index % base); index /= base; }
    return r;
}

// ==========================================================================================
// Bloom — shaders/bloom_threshold.frag:18-35, bloom_downsample.frag:16-108,
// bloom_upsample.frag:16-41; fragTexCoord = (p+0.5)/dstSize (postprocess.vert:8-14). PIN-9.
// ==========================================================================================
void ash_oracle_bloom_threshold_fmt(const uint16_t* hdr, int W, int H, float threshold, uint16_t* out, int fmt) {   // 1 = src f32, 4 = dst f32
    const Img16 src{hdr, W, H, fmt & 1};
    const float knee = 0.5f;                                                      // :32
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            vec3 c = xyz(texture16(src, ((float)x + 0.5f) / (float)W, ((float)y + 0.5f) / (float)H));
            float br = gmax(c.x, gmax(c.y, c.z));                                 // :19
            float soft = br - threshold + knee;
            soft = gclamp(soft, 0.0f, 2.0f * knee);
            soft = soft * soft / (4.0f * knee + 0.00001f);
            float contrib = gmax(soft, br - threshold);
            contrib /= gmax(br, 0.00001f);
            store16(out, W, x, y, c * contrib, 1.0f, (fmt >> 2) & 1);
        }
}
void ash_oracle_bloom_threshold(const uint16_t* hdr, int W, int H, float threshold, uint16_t* out) {
    ash_oracle_bloom_threshold_fmt(hdr, W, H, threshold, out, 0);
}

void ash_oracle_bloom_downsample_fmt(const uint16_t* srcp, int sw, int sh, uint16_t* dst, int dw, int dh, int f32) {
    const Img16 src{srcp, sw, sh, f32};
    const float dx = 1.0f / (float)sw, dy = 1.0f / (float)sh;                     // texelSize (:11)
    const vec3 LUMA = {0.2126f, 0.7152f, 0.0722f};
#pragma omp parallel for schedule(static)
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            float u = ((float)x + 0.5f) / (float)dw, v = ((float)y + 0.5f) / (float)dh;
            // order: center, a, b, c, d, e, f, g, h (:22-34) with weights (:87-101)
            const float ou[9] = {0.0f, -dx, dx, -dx, dx, -dx, dx, 0.0f, 0.0f};
            const float ov[9] = {0.0f, -dy, -dy, dy, dy, 0.0f, 0.0f, -dy, dy};
            const float wt[9] = {0.25f, 0.0625f, 0.0625f, 0.0625f, 0.0625f, 0.125f, 0.125f, 0.125f, 0.125f};
            vec3 res = {0.0f, 0.0f, 0.0f}; float sum = 0.0f;
            for (int k = 0; k < 9; ++k) {
                vec3 s = xyz(texture16(src, u + ou[k], v + ov[k]));
                float w = 1.0f / (1.0f + dot3(s, LUMA));
                res = res + (s * w) * wt[k];
                sum += w * wt[k];
            }
            store16(dst, dw, x, y, res / sum, 1.0f, f32);                         // :107
        }
}
void ash_oracle_bloom_downsample(const uint16_t* srcp, int sw, int sh, uint16_t* dst, int dw, int dh) {
    ash_oracle_bloom_downsample_fmt(srcp, sw, sh, dst, dw, dh, 0);
}

// dst (dw x dh) = RN16( float(dst) + upsample(src) )   (additive blend, PIN-9)
void ash_oracle_bloom_upsample_add_fmt(const uint16_t* srcp, int sw, int sh, uint16_t* dst, int dw, int dh,
                                       float intensity, int f32) {
    const Img16 src{srcp, sw, sh, f32};
    const Img16 dsti{dst, dw, dh, f32};
    const float dx = (1.0f / (float)sw) * 0.5f, dy = (1.0f / (float)sh) * 0.5f;   // :18
#pragma omp parallel for schedule(static)
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            float u = ((float)x + 0.5f) / (float)dw, v = ((float)y + 0.5f) / (float)dh;
            vec3 s0 = xyz(texture16(src, u + (-dx * 2.0f), v + 0.0f));            // :21-29
            vec3 s1 = xyz(texture16(src, u + (-dx), v + (-dy)));
            vec3 s2 = xyz(texture16(src, u + 0.0f, v + (-dy * 2.0f)));
            vec3 s3 = xyz(texture16(src, u + dx, v + (-dy)));
            vec3 s4 = xyz(texture16(src, u + dx * 2.0f, v + 0.0f));
            vec3 s5 = xyz(texture16(src, u + dx, v + dy));
            vec3 s6 = xyz(texture16(src, u + 0.0f, v + dy * 2.0f));
            vec3 s7 = xyz(texture16(src, u + (-dx), v + dy));
            vec3 s8 = xyz(texture16(src, u, v));
            vec3 r = s8 * 4.0f;                                                   // :32-38
            r = r + (((s1 + s3) + s5) + s7) * 2.0f;
            r = r + (((s0 + s2) + s4) + s6);
            r = r / 16.0f;
            r = r * intensity;
            const vec3 d = xyz(fetch16(dsti, x, y));
            store16(dst, dw, x, y, r + d, 1.0f, f32);
        }
}
void ash_oracle_bloom_upsample_add(const uint16_t* srcp, int sw, int sh, uint16_t* dst, int dw, int dh, float intensity) {
    ash_oracle_bloom_upsample_add_fmt(srcp, sw, sh, dst, dw, dh, intensity, 0);
}

// ==========================================================================================
// Tonemap — shaders/tonemapping.frag:20-49, 8-bit store per PIN-11.
// ==========================================================================================
// TonemapOperator (src/renderer/features/tonemapping.rs:9-20): Aces = the shader's curve (:20-27); Reinhard, Uncharted2 and
// None have no shader in the reference and follow PIN-13 (include/ashcuda.h): the published closed forms.
static float hable(float x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}
static float tonemap_curve(float xx, uint32_t op) {
    switch (op) {
        case ASH_TONEMAP_REINHARD:   return gclamp(xx / (1.0f + xx), 0.0f, 1.0f);
        case ASH_TONEMAP_UNCHARTED2: return gclamp(hable(2.0f * xx) / hable(11.2f), 0.0f, 1.0f);
        case ASH_TONEMAP_NONE:       return gclamp(xx, 0.0f, 1.0f);
        default:                     return gclamp((xx * (2.51f * xx + 0.03f)) / (xx * (2.43f * xx + 0.59f) + 0.14f), 0.0f, 1.0f);   // :20-27
    }
}

void ash_oracle_tonemap_fmt(const uint16_t* hdrp, const uint16_t* bloomp, int W, int H,
                            const AshPostProcessPushConstants* pc, uint32_t op, uint8_t* ldr8, int fmt) {   // 1 = hdr f32, 2 = bloom f32
    const Img16 hdr{hdrp, W, H, fmt & 1}, bloom{bloomp, W, H, (fmt >> 1) & 1};
    const float inv_gamma = 1.0f / pc->gamma;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float u = ((float)x + 0.5f) / (float)W, v = ((float)y + 0.5f) / (float)H;
            vec3 c = xyz(texture16(hdr, u, v));
            vec3 b = bloomp ? xyz(texture16(bloom, u, v)) : vec3{0.0f, 0.0f, 0.0f};
            c = c + b * pc->bloom_intensity;                                      // :38
            c = c * pc->exposure;                                                 // :41
            float in[3] = {c.x, c.y, c.z};
            uint8_t* o = ldr8 + ((size_t)y * W + x) * 4;
            for (int k = 0; k < 3; ++k) {
                float l = tonemap_curve(in[k], op);
                l = powf(l, inv_gamma);                                           // :47
                o[k] = (uint8_t)floorf(gclamp(l, 0.0f, 1.0f) * 255.0f + 0.5f);   // PIN-11
            }
            o[3] = 255;
        }
}

// PIN-6 weight model of the LINEAR sampler: 0 = 8-bit sub-texel weights (what the library implements),
// 1 = exact fp32 weights (SURVEY.md PIN-6 as first written); tests/test_oracle_ref.py bounds the difference.
void ash_oracle_set_weight_mode(int exact) { g_exact_weights = exact ? 1 : 0; }

void ash_oracle_tonemap_op(const uint16_t* hdrp, const uint16_t* bloomp, int W, int H,
                           const AshPostProcessPushConstants* pc, uint32_t op, uint8_t* ldr8) {
    ash_oracle_tonemap_fmt(hdrp, bloomp, W, H, pc, op, ldr8, 0);
}
void ash_oracle_tonemap(const uint16_t* hdrp, const uint16_t* bloomp, int W, int H,
                        const AshPostProcessPushConstants* pc, uint8_t* ldr8) {
    ash_oracle_tonemap_op(hdrp, bloomp, W, H, pc, ASH_TONEMAP_ACES, ldr8);
}

// fp16 helpers exported for the tests (checked against numpy.float16)
uint16_t ash_oracle_f2h(float f) { return f2h(f); }
float    ash_oracle_h2f(uint16_t h) { return h2f(h); }
void ash_oracle_f2h_array(const float* in, uint16_t* out, size_t n) { for (size_t i = 0; i < n; ++i) out[i] = f2h(in[i]); }
void ash_oracle_h2f_array(const uint16_t* in, float* out, size_t n) { for (size_t i = 0; i < n; ++i) out[i] = h2f(in[i]); }
int ash_oracle_num_threads(void);

}  // extern "C"

#ifdef _OPENMP
#include <omp.h>
extern "C" int ash_oracle_num_threads(void) { return omp_get_max_threads(); }
#else
extern "C" int ash_oracle_num_threads(void) { return 1; }
#endif
