// ash_oracle_utils.cpp — CPU ORACLE (test infrastructure, NOT product code) for shaders/compute_utils.glsl
// (SURVEY §8.6 row f-4).  A literal, serial emulation of the GLSL: the shared arrays are plain arrays, each
// barrier-separated step is a loop over localId in ascending order (every step reads only elements no
// other invocation writes in the same step, so the order inside a step is irrelevant).
// PARITY UNPINNED: no shader of the reference includes the library and it has no tests.
#include <cmath>
#include <cstdint>
#include <cstring>
#include "../include/ashcuda.h"

namespace {
inline float gmin(float a, float b) { return (b < a) ? b : a; }
inline float gmax(float a, float b) { return (a < b) ? b : a; }
struct vec3 { float x, y, z; };
struct Plane { vec3 normal; float distance; };
inline float dot3(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }

float reduce256(const float* in, int op) {                         // parallelMin / Max / Sum, :22-65
    float s[256];
    for (int i = 0; i < 256; ++i) s[i] = in[i];
    for (unsigned stride = 128; stride > 0; stride >>= 1)
        for (unsigned i = 0; i < stride; ++i)
            s[i] = op == ASH_CU_MIN ? gmin(s[i], s[i + stride]) : op == ASH_CU_MAX ? gmax(s[i], s[i + stride]) : s[i] + s[i + stride];
    return s[0];
}

void scan256(const uint32_t* in, uint32_t* out) {                  // exclusivePrefixSum, :80-111
    uint32_t s[256];
    for (int i = 0; i < 256; ++i) s[i] = in[i];
    for (unsigned stride = 1; stride < 256; stride <<= 1)          // up-sweep
        for (unsigned lid = 0; lid < 256; ++lid) {
            const unsigned index = (lid + 1) * stride * 2 - 1;
            if (index < 256) s[index] += s[index - stride];
        }
    s[255] = 0;
    for (unsigned stride = 128; stride > 0; stride >>= 1)          // down-sweep
        for (unsigned lid = 0; lid < 256; ++lid) {
            const unsigned index = (lid + 1) * stride * 2 - 1;
            if (index < 256) { const uint32_t t = s[index - stride]; s[index - stride] = s[index]; s[index] += t; }
        }
    for (int i = 0; i < 256; ++i) out[i] = s[i];
}

Plane extract_plane(const float m[16], int face) {                 // :126-160, m[c*4 + r] = column c
    const float *c0 = m, *c1 = m + 4, *c2 = m + 8, *c3 = m + 12;
    float row[4];
    for (int k = 0; k < 4; ++k)
        switch (face) {
            case 0: row[k] = c3[k] + c0[k]; break;
            case 1: row[k] = c3[k] - c0[k]; break;
            case 2: row[k] = c3[k] + c1[k]; break;
            case 3: row[k] = c3[k] - c1[k]; break;
            case 4: row[k] = c3[k] + c2[k]; break;
            case 5: row[k] = c3[k] - c2[k]; break;
            default: row[k] = k == 3 ? 1.0f : 0.0f;
        }
    const float len = sqrtf(dot3({row[0], row[1], row[2]}, {row[0], row[1], row[2]}));
    return {{row[0] / len, row[1] / len, row[2] / len}, row[3] / len};
}
bool sphere_in_frustum(vec3 c, float radius, const Plane pl[6]) {  // :164-178
    for (int i = 0; i < 6; ++i)
        if (dot3(pl[i].normal, c) + pl[i].distance < -radius) return false;
    return true;
}
inline float mix_step(float lo, float hi, float n) { const float t = n < 0.0f ? 0.0f : 1.0f; return lo * (1.0f - t) + hi * t; }   // mix(min, max, step(0, n))
bool aabb_in_frustum(vec3 mn, vec3 mx, const Plane pl[6]) {        // :181-195
    for (int i = 0; i < 6; ++i) {
        const vec3 pv = {mix_step(mn.x, mx.x, pl[i].normal.x), mix_step(mn.y, mx.y, pl[i].normal.y), mix_step(mn.z, mx.z, pl[i].normal.z)};
        if (dot3(pl[i].normal, pv) + pl[i].distance < 0.0f) return false;
    }
    return true;
}
}  // namespace

extern "C" int ash_oracle_compute_utils(int op, const void* in, uint32_t n, const float* p, void* out) {
    const float* f = static_cast<const float*>(in);
    Plane pl[6];
    if (op == ASH_CU_SPHERE_IN_FRUSTUM || op == ASH_CU_AABB_IN_FRUSTUM || op == ASH_CU_PLANES)
        for (int i = 0; i < 6; ++i) pl[i] = extract_plane(p, i);
    switch (op) {
        case ASH_CU_MIN: case ASH_CU_MAX: case ASH_CU_SUM:
            for (uint32_t g = 0; g < n / 256; ++g) static_cast<float*>(out)[g] = reduce256(f + g * 256, op);
            return 0;
        case ASH_CU_SCAN:
            for (uint32_t g = 0; g < n / 256; ++g) scan256(static_cast<const uint32_t*>(in) + g * 256, static_cast<uint32_t*>(out) + g * 256);
            return 0;
        case ASH_CU_SPHERE_IN_FRUSTUM:
            for (uint32_t i = 0; i < n; ++i) static_cast<uint32_t*>(out)[i] = sphere_in_frustum({f[i * 4], f[i * 4 + 1], f[i * 4 + 2]}, f[i * 4 + 3], pl);
            return 0;
        case ASH_CU_AABB_IN_FRUSTUM:
            for (uint32_t i = 0; i < n; ++i) static_cast<uint32_t*>(out)[i] = aabb_in_frustum({f[i * 6], f[i * 6 + 1], f[i * 6 + 2]}, {f[i * 6 + 3], f[i * 6 + 4], f[i * 6 + 5]}, pl);
            return 0;
        case ASH_CU_DEPTH:                                           // :205-219; p = inv_proj columns, near, far
            for (uint32_t i = 0; i < n; ++i) {
                const float u = f[i * 3], v = f[i * 3 + 1], d = f[i * 3 + 2], near_ = p[16], far_ = p[17];
                float* o = static_cast<float*>(out) + i * 4;
                o[0] = near_ * far_ / (far_ + d * (near_ - far_));
                const float c[4] = {u * 2.0f - 1.0f, v * 2.0f - 1.0f, d, 1.0f};
                float q[4];
                for (int r = 0; r < 4; ++r) q[r] = ((p[r] * c[0] + p[4 + r] * c[1]) + p[8 + r] * c[2]) + p[12 + r] * c[3];
                o[1] = q[0] / q[3]; o[2] = q[1] / q[3]; o[3] = q[2] / q[3];
            }
            return 0;
        case ASH_CU_PLANES:
            for (int i = 0; i < 6; ++i) {
                float* o = static_cast<float*>(out) + i * 4;
                o[0] = pl[i].normal.x; o[1] = pl[i].normal.y; o[2] = pl[i].normal.z; o[3] = pl[i].distance;
            }
            return 0;
    }
    return -1;
}
