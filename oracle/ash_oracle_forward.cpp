// ash_oracle_forward.cpp — CPU ORACLE (test infrastructure, NOT product code) for SURVEY §8.6 row f-1
// (depth / HDR producers) and the tiled-forward light loop of row f-2.
//
// PARITY UNPINNED for the geometry and the light loop: the reference rasterises triangle meshes through
// Vulkan (renderer.rs:1804-1992, shaders/vert.vert) and never reads the tile light lists, so the occluder
// spheres, the ray cast and the point-light attenuation below are this repository's own definition
// (DESIGN.md §8c).  What IS a restatement of reference code is the surface shading: shaders/frag.frag's
// Cook-Torrance terms, line by line —
//   distribution_ggx :89-94, geometry_schlick_ggx_fast :97-99, geometry_smith :101-105,
//   fresnel_schlick_fast :108-114, main :167-213 (NdotL, F0, half vector, D/G/F, specular cap, kD, diffuse,
//   Lo) and :216-226 (ambient, emissive, Reinhard) — with textures absent (bindless index -1 => factor
//   only), shadow = 0 and the geometric normal.  Evaluation order is left to right as written (PIN-1),
//   in view space (the camera at the origin) instead of world space.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../include/ashcuda.h"

namespace {

struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
inline float gmin(float a, float b) { return (b < a) ? b : a; }
inline float gmax(float a, float b) { return (a < b) ? b : a; }
inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot3(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length3(vec3 v) { return sqrtf(dot3(v, v)); }
inline vec3 normalize3(vec3 v) { return v / length3(v); }
// m * v, columns summed left to right (m[c] = column c)
inline vec4 mul(const float m[4][4], vec4 v) {
    vec4 r;
    r.x = ((m[0][0] * v.x + m[1][0] * v.y) + m[2][0] * v.z) + m[3][0] * v.w;
    r.y = ((m[0][1] * v.x + m[1][1] * v.y) + m[2][1] * v.z) + m[3][1] * v.w;
    r.z = ((m[0][2] * v.x + m[1][2] * v.y) + m[2][2] * v.z) + m[3][2] * v.w;
    r.w = ((m[0][3] * v.x + m[1][3] * v.y) + m[2][3] * v.z) + m[3][3] * v.w;
    return r;
}
inline uint32_t fbits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float bitsf(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

uint16_t f2h(float f) {                                   // fp32 -> fp16, round to nearest even
    uint32_t x = fbits(f);
    uint32_t s = (x >> 16) & 0x8000u, a = x & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? 0x200u : 0u));
    if (a >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);
    if (a < 0x33000001u) return (uint16_t)s;
    int e = (int)(a >> 23) - 127; uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;
    uint32_t hm = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (hm & 1u))) ++hm;
    uint32_t out = (e < -14) ? hm : ((((uint32_t)(e + 15) - 1u) << 10) + hm);
    return (uint16_t)(s | out);
}

// ---- the occluder of one cull object: a sphere inside its world AABB, in view space
struct Sphere { vec3 c; float r; };
Sphere sphere_of(const AshCullObjectData& o, const AshCullingCameraData& cam, float radius_scale) {
    if (!(o.flags & 1u)) return {{0, 0, 0}, 0.0f};
    const vec4 wc = mul(o.model, {o.bounds.center[0], o.bounds.center[1], o.bounds.center[2], 1.0f});
    float half[3];
    for (int k = 0; k < 3; ++k)
        half[k] = o.bounds.extents[k] * length3({o.model[k][0], o.model[k][1], o.model[k][2]});
    const float r = radius_scale * gmin(gmin(half[0], half[1]), half[2]);
    const vec4 cv = mul(cam.view, {wc.x, wc.y, wc.z, 1.0f});
    return {{cv.x, cv.y, cv.z}, r};
}

// per-axis pixel bound of the part of the sphere in front of the near plane, padded by a pixel:
// tangent rays through the eye when the eye plane does not cut the sphere (cz > r), else the box bound
// of { x in [c-r, c+r], z in [near, cz+r] }
void axis_bound(float c, float cz, float r, bool tangent, float near_plane, float P, int S, int* i0, int* i1) {
    float s0, s1;
    if (tangent) {
        const float den = (cz - r) * (cz + r);
        const float D = (c * c + cz * cz) - r * r;
        const float sq = sqrtf(gmax(D, 0.0f));
        const float m = c * cz, k = r * sq;
        s0 = (m - k) / den; s1 = (m + k) / den;
    } else {
        const float zmax = cz + r, lo_c = c - r, hi_c = c + r;
        s0 = lo_c < 0.0f ? lo_c / near_plane : lo_c / zmax;
        s1 = hi_c > 0.0f ? hi_c / near_plane : hi_c / zmax;
    }
    const float n0 = P * s0, n1 = P * s1;
    const float p0 = (n0 * 0.5f + 0.5f) * (float)S, p1 = (n1 * 0.5f + 0.5f) * (float)S;
    const float lo = gmin(p0, p1), hi = gmax(p0, p1);
    if (!(lo <= hi)) { *i0 = 0; *i1 = S; return; }
    *i0 = (int)gclamp(floorf(lo) - 1.0f, 0.0f, (float)S);
    *i1 = (int)gclamp(ceilf(hi) + 1.0f, 0.0f, (float)S);
}
struct Rect { int x0, y0, x1, y1; };
Rect rect_of(const Sphere& s, const AshCullingCameraData& cam, float near_plane, int W, int H) {
    const float cz = -s.c.z;
    if (!(s.r > 0.0f) || cz + s.r < near_plane) return {0, 0, 0, 0};
    const bool tangent = cz > s.r;
    Rect q;
    axis_bound(s.c.x, cz, s.r, tangent, near_plane, cam.projection[0][0], W, &q.x0, &q.x1);
    axis_bound(s.c.y, cz, s.r, tangent, near_plane, cam.projection[1][1], H, &q.y0, &q.y1);
    return q;
}

// view-space ray through a pixel centre: d = (dx, dy, -1)
inline float ray_coord(int i, int S, float P) { return ((((float)i + 0.5f) / (float)S) * 2.0f - 1.0f) / P; }

// nearest intersection of the ray with the sphere: distance along the view axis, or < 0 for a miss
inline float hit_t(const Sphere& s, float dx, float dy, float near_plane) {
    const float a = (dx * dx + dy * dy) + 1.0f;
    const float b = (dx * s.c.x + dy * s.c.y) - s.c.z;
    const float cc = dot3(s.c, s.c) - s.r * s.r;
    const float disc = b * b - a * cc;
    if (!(disc >= 0.0f)) return -1.0f;
    const float t = (b - sqrtf(disc)) / a;
    return (t >= near_plane) ? t : -1.0f;
}
inline bool depth_of(float t, const AshCullingCameraData& cam, float* depth) {   // ndc z of view z = -t (w_clip = t)
    const float d = ((cam.projection[3][2] - cam.projection[2][2] * t) / t) + 0.0f;   // + 0: -0 -> +0
    if (!(d >= 0.0f && d <= 1.0f)) return false;
    *depth = d;
    return true;
}

const float PI_F = 3.14159265359f;                                              // frag.frag:45

vec3 brdf(vec3 N, vec3 V, vec3 L, vec3 base, float metallic, float roughness, vec3 lightColor) {
    const float NdotL = gmax(dot3(N, L), 0.0f);                                 // :167
    const float im = 1.0f - metallic;
    const vec3 F0 = {0.04f * im + base.x * metallic, 0.04f * im + base.y * metallic, 0.04f * im + base.z * metallic};   // :186 mix
    const vec3 Hh = normalize3(V + L);                                          // :188
    const float NdotV = gmax(dot3(N, V), 0.001f);
    const float NdotH = gmax(dot3(N, Hh), 0.0f);
    const float VdotH = gmax(dot3(V, Hh), 0.0f);
    const float a = roughness * roughness, a2 = a * a;                          // :89-94
    const float dd = (NdotH * NdotH) * (a2 - 1.0f) + 1.0f;
    const float D = a2 / ((PI_F * dd) * dd);
    const float r1 = roughness + 1.0f, k = (r1 * r1) * 0.125f;                  // :101-105
    const float G = (NdotV / (NdotV * (1.0f - k) + k)) * (NdotL / (NdotL * (1.0f - k) + k));
    const float t = gclamp(1.0f - VdotH, 0.0f, 1.0f), t2 = t * t, t5 = (t2 * t2) * t;   // :108-114
    const vec3 F = {F0.x + (1.0f - F0.x) * t5, F0.y + (1.0f - F0.y) * t5, F0.z + (1.0f - F0.z) * t5};
    const float DG = D * G;
    const float denom = ((4.0f * NdotV) * NdotL) + 0.001f;                      // :198
    vec3 spec = {(DG * F.x) / denom, (DG * F.y) / denom, (DG * F.z) / denom};
    spec = {gmin(spec.x, 10.0f / gmax(0.04f, F0.x)), gmin(spec.y, 10.0f / gmax(0.04f, F0.y)), gmin(spec.z, 10.0f / gmax(0.04f, F0.z))};   // :202
    const vec3 kD = {(1.0f - F.x) * im, (1.0f - F.y) * im, (1.0f - F.z) * im};  // :204
    const vec3 diffuse = (kD * base) / PI_F;                                    // :205
    return ((diffuse + spec) * lightColor) * NdotL;                             // :213, shadow = 0
}

}  // namespace

extern "C" {

// keys[y*W + x] = (depth bits << 32) | object index of the nearest hit (lowest index on equal depth),
// (bits(1.0f) << 32) | ASH_FWD_NO_OBJECT where nothing is hit; depth[] = the high halves as floats.
// brute != 0 tests every pixel against every sphere instead of the silhouette rectangle (tests only).
// mask (or NULL): only objects with mask[i] != 0 are drawn; accumulate != 0: keys[] already holds earlier
// draws of the same view (two-phase culling draws the late objects on top of the early ones).
void ash_oracle_forward_depth(const AshCullObjectData* objs, uint32_t n, const AshCullingCameraData* cam,
                              const AshForwardParams* fp, uint32_t W, uint32_t H, int brute,
                              const uint32_t* mask, int accumulate, uint64_t* keys, float* depth) {
    const size_t P = (size_t)W * H;
    const uint64_t clear = ((uint64_t)fbits(1.0f) << 32) | ASH_FWD_NO_OBJECT;
    if (!accumulate) for (size_t i = 0; i < P; ++i) keys[i] = clear;
    std::vector<float> rx(W), ry(H);
    for (uint32_t x = 0; x < W; ++x) rx[x] = ray_coord((int)x, (int)W, cam->projection[0][0]);
    for (uint32_t y = 0; y < H; ++y) ry[y] = ray_coord((int)y, (int)H, cam->projection[1][1]);
    for (uint32_t i = 0; i < n; ++i) {
        if (mask && !mask[i]) continue;
        const Sphere s = sphere_of(objs[i], *cam, fp->radius_scale);
        Rect q = rect_of(s, *cam, fp->near_plane, (int)W, (int)H);
        if (brute && s.r > 0.0f) q = {0, 0, (int)W, (int)H};
        for (int y = q.y0; y < q.y1; ++y)
            for (int x = q.x0; x < q.x1; ++x) {
                const float t = hit_t(s, rx[x], ry[y], fp->near_plane);
                float d;
                if (t < 0.0f || !depth_of(t, *cam, &d)) continue;
                const uint64_t key = ((uint64_t)fbits(d) << 32) | i;
                if (key < keys[(size_t)y * W + x]) keys[(size_t)y * W + x] = key;
            }
    }
    for (size_t i = 0; i < P; ++i) depth[i] = bitsf((uint32_t)(keys[i] >> 32));
}

// RG16F velocity of the forward pass (csrc/forward.cu: k_forward_velocity): velocity = uv - prevUV as
// taa_resolve.comp:94-102 consumes it (historyUV = uv - velocity); prevUV from reproj * (view-space hit, 1),
// reproj = prevViewProj * inverse(view) in columns.  No hit, or a point behind the previous camera: 0.
void ash_oracle_forward_velocity(const uint64_t* keys, const AshCullObjectData* objs, uint32_t n,
                                 const AshCullingCameraData* cam, const AshForwardParams* fp, const float reproj[4][4],
                                 uint32_t W, uint32_t H, uint16_t* vel) {
    std::vector<Sphere> spheres(n);
    for (uint32_t i = 0; i < n; ++i) spheres[i] = sphere_of(objs[i], *cam, fp->radius_scale);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)H; ++y) {
        const float dy = ray_coord(y, (int)H, cam->projection[1][1]);
        for (int x = 0; x < (int)W; ++x) {
            const size_t p = (size_t)y * W + x;
            const uint32_t id = (uint32_t)(keys[p] & 0xFFFFFFFFu);
            float vx = 0.0f, vy = 0.0f, t = -1.0f;
            const float dx = ray_coord(x, (int)W, cam->projection[0][0]);
            if (id < n) t = hit_t(spheres[id], dx, dy, fp->near_plane);
            if (t >= 0.0f) {
                const vec4 clip = mul(reproj, {dx * t, dy * t, -t, 1.0f});
                if (clip.w > 0.0f) {
                    const float pu = (clip.x / clip.w) * 0.5f + 0.5f, pv = (clip.y / clip.w) * 0.5f + 0.5f;
                    vx = ((float)x + 0.5f) / (float)W - pu;
                    vy = ((float)y + 0.5f) / (float)H - pv;
                }
            }
            vel[p * 2] = f2h(vx); vel[p * 2 + 1] = f2h(vy);
        }
    }
}

void ash_oracle_forward_shade(const uint64_t* keys, const AshCullObjectData* objs, uint32_t n,
                              const AshGpuLight* lights, uint32_t n_lights, const uint32_t* tile_data,
                              const AshInstanceData* inst_or_null, uint32_t n_inst,
                              const AshCullingCameraData* cam, const AshForwardParams* fp,
                              uint32_t W, uint32_t H, uint16_t* hdr) {
    const uint32_t tiles_x = (W + 15) / 16;
    std::vector<Sphere> spheres(n);
    for (uint32_t i = 0; i < n; ++i) spheres[i] = sphere_of(objs[i], *cam, fp->radius_scale);
    std::vector<vec4> lv(n_lights);                       // point: view position, radius; directional: unit L
    for (uint32_t i = 0; i < n_lights; ++i) {
        const AshGpuLight& l = lights[i];
        if (l.direction[3] == 2.0f) {
            const vec4 d = mul(cam->view, {l.direction[0], l.direction[1], l.direction[2], 0.0f});
            const vec3 L = normalize3(-vec3{d.x, d.y, d.z});
            lv[i] = {L.x, L.y, L.z, 0.0f};
        } else {
            const vec4 p = mul(cam->view, {l.position[0], l.position[1], l.position[2], 1.0f});
            lv[i] = {p.x, p.y, p.z, l.position[3]};
        }
    }
    const vec4 sd = mul(cam->view, {fp->light_direction[0], fp->light_direction[1], fp->light_direction[2], 0.0f});
    const vec3 sunL = normalize3(-vec3{sd.x, sd.y, sd.z});                       // frag.frag:121
    const vec3 sunC = {fp->light_color[0], fp->light_color[1], fp->light_color[2]};
    const float metallic = fp->parameters[0], roughness = gmax(fp->parameters[1], 0.04f);   // :170-171
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < (int)H; ++y) {
        const float dy = ray_coord(y, (int)H, cam->projection[1][1]);
        for (int x = 0; x < (int)W; ++x) {
            const size_t p = (size_t)y * W + x;
            uint16_t* o = hdr + p * 4;
            o[3] = f2h(1.0f);
            const uint32_t id = (uint32_t)(keys[p] & 0xFFFFFFFFu);
            float t = -1.0f;
            const float dx = ray_coord(x, (int)W, cam->projection[0][0]);
            if (id < n) t = hit_t(spheres[id], dx, dy, fp->near_plane);
            if (t < 0.0f) { o[0] = f2h(fp->background[0]); o[1] = f2h(fp->background[1]); o[2] = f2h(fp->background[2]); continue; }
            const vec3 pos = {dx * t, dy * t, -t};
            const vec3 N = normalize3(pos - spheres[id].c);
            const vec3 V = normalize3(-pos);                                     // :120, camera at the origin
            vec3 base = {fp->base_color_factor[0], fp->base_color_factor[1], fp->base_color_factor[2]};   // :128
            if ((fp->flags & ASH_FWD_INSTANCE_COLOR) && inst_or_null && id < n_inst)
                base = base * vec3{inst_or_null[id].color[0], inst_or_null[id].color[1], inst_or_null[id].color[2]};
            const vec3 ambient = vec3{fp->ambient_color[0], fp->ambient_color[1], fp->ambient_color[2]} * base;      // :216
            const vec3 Lo = brdf(N, V, sunL, base, metallic, roughness, sunC);
            vec3 color = (ambient + Lo) + vec3{fp->emissive_factor[0], fp->emissive_factor[1], fp->emissive_factor[2]};   // :224
            if (fp->flags & ASH_FWD_TILE_LIGHTS) {
                const uint32_t* td = tile_data + ((size_t)(y >> 4) * tiles_x + (x >> 4)) * 257;
                const uint32_t cnt = td[0] < 256u ? td[0] : 256u;
                for (uint32_t k = 0; k < cnt; ++k) {
                    const uint32_t li = td[1 + k];
                    if (li >= n_lights) continue;
                    const AshGpuLight& l = lights[li];
                    vec3 rad = vec3{l.color[0], l.color[1], l.color[2]} * l.color[3];
                    vec3 L;
                    if (l.direction[3] == 2.0f) L = {lv[li].x, lv[li].y, lv[li].z};
                    else {
                        const vec3 toL = vec3{lv[li].x, lv[li].y, lv[li].z} - pos;
                        const float d2 = dot3(toL, toL);
                        if (!(d2 > 0.0f)) continue;
                        const float dist = sqrtf(d2);
                        L = toL / dist;
                        const float q = dist / lv[li].w, q2 = q * q;              // windowed inverse square (ours)
                        const float w = gclamp(1.0f - q2 * q2, 0.0f, 1.0f);
                        rad = rad * ((w * w) / (d2 + 1.0f));
                    }
                    color = color + brdf(N, V, L, base, metallic, roughness, rad);
                }
            }
            if (fp->flags & ASH_FWD_REINHARD) color = color / (color + vec3{1.0f, 1.0f, 1.0f});   // :226
            o[0] = f2h(color.x); o[1] = f2h(color.y); o[2] = f2h(color.z);
        }
    }
}

}  // extern "C"
