// glsl_compat.hpp — TEST INFRASTRUCTURE (oracle/_ref): the minimum of GLSL 4.50 needed to compile the
// reference's compute-path shaders, UNMODIFIED in their logic, as C++17 with g++ (see strip_glsl.py for the
// five lexical rewrites).  Nothing under ash_renderer_b200/ may include this file.
//
// What comes from the reference: every statement of the eight shaders (included from oracle/_ref/gen/*.inc).
// What this header defines, i.e. what the GLSL spec leaves to the implementation and SURVEY.md §8.0 pins:
//   PIN-1  arithmetic: column-major mat4, m*v summed left to right, dot left to right, no FMA contraction
//          (the Makefile passes -ffp-contract=off), min/max/clamp/sign/mix as the spec writes them;
//   PIN-5  depth / Hi-Z sampler: NEAREST, CLAMP_TO_EDGE, nearest mip;
//   PIN-6  colour sampler: LINEAR, CLAMP_TO_EDGE, RGBA16F / RG16F texels, bilinear in fp32, horizontal then
//          vertical, a*(1-f)+b*f; the weight f is either exact or quantised to 8 fractional bits
//          (ash_ref_set_weight_mode; the library under test uses the 8-bit form);
//   images rgba16f / r32f: imageStore rounds to nearest even (here: gcc's _Float16 conversion, an implementation
//          independent of the bit-twiddling one in oracle/ash_oracle.cpp).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace glsl {

typedef uint32_t uint;

struct vec2; struct vec3; struct vec4; struct ivec2; struct uvec2; struct uvec3;

// ---- swizzle proxies: plain arrays inside the vectors' anonymous unions, convertible to the vector type
template <class V, class T, int N, int A, int B> struct swz2 {
    T d[N];
    operator V() const { return V(d[A], d[B]); }
};
template <class V, class T, int N, int A, int B, int C> struct swz3 {
    T d[N];
    operator V() const { return V(d[A], d[B], d[C]); }
};

struct vec2 {
    union { struct { float x, y; }; struct { float r, g; }; };
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
    explicit vec2(const ivec2& v);
    explicit vec2(const uvec2& v);
    vec2& operator+=(vec2 o) { x += o.x; y += o.y; return *this; }
};
struct vec3 {
    union {
        struct { float x, y, z; }; struct { float r, g, b; };
        swz2<vec2, float, 3, 0, 1> xy;
    };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3& operator+=(vec3 o) { x += o.x; y += o.y; z += o.z; return *this; }
    vec3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
    vec3& operator/=(float s) { x /= s; y /= s; z /= s; return *this; }
};
struct vec4 {
    union {
        struct { float x, y, z, w; }; struct { float r, g, b, a; };
        swz2<vec2, float, 4, 0, 1> xy; swz2<vec2, float, 4, 2, 3> zw;
        swz3<vec3, float, 4, 0, 1, 2> xyz; swz3<vec3, float, 4, 0, 1, 2> rgb;
    };
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    vec4(vec2 v, float c, float d) : x(v.x), y(v.y), z(c), w(d) {}
    vec4& operator/=(float s) { x /= s; y /= s; z /= s; w /= s; return *this; }
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(const uvec2& v);
};
struct uvec2 {
    uint x, y;
    uvec2() : x(0), y(0) {}
    uvec2(uint a, uint b) : x(a), y(b) {}
};
struct uvec3 {
    union { struct { uint x, y, z; }; swz2<uvec2, uint, 3, 0, 1> xy; };
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
};
inline vec2::vec2(const ivec2& v) : x((float)v.x), y((float)v.y) {}
inline vec2::vec2(const uvec2& v) : x((float)v.x), y((float)v.y) {}
inline ivec2::ivec2(const uvec2& v) : x((int)v.x), y((int)v.y) {}

struct mat4 {                       // column-major: m[c] is column c (PIN-1)
    vec4 c[4];
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
static_assert(sizeof(vec2) == 8 && sizeof(vec3) == 12 && sizeof(vec4) == 16 && sizeof(mat4) == 64, "std430 sizes");
static_assert(sizeof(uvec2) == 8 && sizeof(uvec3) == 12, "std430 sizes");

// ---- component-wise arithmetic
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }
inline vec2 operator+(vec2 a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator-(vec2 a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator/(vec2 a, float s) { return vec2(a.x / s, a.y / s); }
inline vec2 operator*(float s, vec2 a) { return vec2(s * a.x, s * a.y); }

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(vec3 a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator+(float s, vec3 a) { return vec3(s + a.x, s + a.y, s + a.z); }
inline vec3 operator-(float s, vec3 a) { return vec3(s - a.x, s - a.y, s - a.z); }

inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }

inline uvec2 operator+(uvec2 a, uvec2 b) { return uvec2(a.x + b.x, a.y + b.y); }
inline uvec2 operator*(uvec2 a, uvec2 b) { return uvec2(a.x * b.x, a.y * b.y); }
inline uvec2 operator/(uvec2 a, uvec2 b) { return uvec2(a.x / b.x, a.y / b.y); }
inline uvec2 operator+(uvec2 a, uint s) { return uvec2(a.x + s, a.y + s); }
inline uvec2 operator-(uvec2 a, uint s) { return uvec2(a.x - s, a.y - s); }

// m * v = m[0]*v.x + m[1]*v.y + m[2]*v.z + m[3]*v.w, left to right (PIN-1)
inline vec4 operator*(const mat4& m, vec4 v) {
    return vec4(((m.c[0].x * v.x + m.c[1].x * v.y) + m.c[2].x * v.z) + m.c[3].x * v.w,
                ((m.c[0].y * v.x + m.c[1].y * v.y) + m.c[2].y * v.z) + m.c[3].y * v.w,
                ((m.c[0].z * v.x + m.c[1].z * v.y) + m.c[2].z * v.z) + m.c[3].z * v.w,
                ((m.c[0].w * v.x + m.c[1].w * v.y) + m.c[2].w * v.z) + m.c[3].w * v.w);
}

// ---- built-ins (GLSL 4.50 §8), scalar then component-wise
inline float min(float a, float b) { return (b < a) ? b : a; }
inline float max(float a, float b) { return (a < b) ? b : a; }
inline uint min(uint a, uint b) { return (b < a) ? b : a; }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline float abs(float x) { return fabsf(x); }
inline float sign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline float floor(float x) { return floorf(x); }
inline float ceil(float x) { return ceilf(x); }
inline float sqrt(float x) { return sqrtf(x); }
inline float log2(float x) { return log2f(x); }     // the library pins the exact integer form (SURVEY §8.1 a-2 caveat)
inline float pow(float x, float y) { return powf(x, y); }
inline vec2 min(vec2 a, vec2 b) { return vec2(min(a.x, b.x), min(a.y, b.y)); }
inline vec2 max(vec2 a, vec2 b) { return vec2(max(a.x, b.x), max(a.y, b.y)); }
inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline vec2 clamp(vec2 v, vec2 lo, vec2 hi) { return min(max(v, lo), hi); }
inline vec2 clamp(vec2 v, float lo, float hi) { return vec2(clamp(v.x, lo, hi), clamp(v.y, lo, hi)); }
inline vec3 clamp(vec3 v, float lo, float hi) { return vec3(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi)); }
inline vec3 sign(vec3 v) { return vec3(sign(v.x), sign(v.y), sign(v.z)); }
inline vec3 pow(vec3 a, vec3 b) { return vec3(pow(a.x, b.x), pow(a.y, b.y), pow(a.z, b.z)); }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(vec2 v) { return sqrtf(dot(v, v)); }
inline float length(vec3 v) { return sqrtf(dot(v, v)); }
inline vec3 normalize(vec3 v) { float l = length(v); return vec3(v.x / l, v.y / l, v.z / l); }
inline vec3 cross(vec3 a, vec3 b) {   // spec: (x1*y2 - y1*x2, x2*y0 - y2*x0, x0*y1 - y0*x1)
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline uint floatBitsToUint(float f) { uint u; memcpy(&u, &f, 4); return u; }
inline float uintBitsToFloat(uint u) { float f; memcpy(&f, &u, 4); return f; }

// ---- atomics / barriers: invocations of one work group are run one at a time (fibres for shaders with
// barrier(), a plain loop otherwise), in ascending gl_LocalInvocationIndex, so the "atomic" is a plain
// read-modify-write and the resulting order is the serial-execution order (PIN-7).
inline uint atomicAdd(uint& m, uint v) { uint o = m; m = o + v; return o; }
inline uint atomicMin(uint& m, uint v) { uint o = m; m = v < o ? v : o; return o; }
inline uint atomicMax(uint& m, uint v) { uint o = m; m = v > o ? v : o; return o; }
inline void (*barrier_hook)() = nullptr;
inline void barrier() { if (barrier_hook) barrier_hook(); }
inline void memoryBarrierImage() {}

inline thread_local uvec3 gl_GlobalInvocationID, gl_WorkGroupID, gl_LocalInvocationID;

// ---- half floats through gcc's _Float16 (IEEE binary16, round to nearest even on conversion)
inline float half_to_float(uint16_t h) { _Float16 v; memcpy(&v, &h, 2); return (float)v; }
inline uint16_t float_to_half(float f) { _Float16 v = (_Float16)f; uint16_t h; memcpy(&h, &v, 2); return h; }

// ---- textures and images
enum Format { FMT_R32F = 0, FMT_RGBA16F = 1, FMT_RG16F = 2, FMT_RGBA8 = 3, FMT_RGBA32F = 4 };   // RGBA32F: fp32_intermediates mode (PIN-6)
inline int g_exact_weights = 0;  // 1 = exact fp32 bilinear weights (SURVEY PIN-6), 0 = 8-bit sub-texel weights

struct sampler2D {
    const void* level[16];
    int w[16], h[16];
    int levels, format, linear;
};
struct image2D {
    void* p; int w, h, format;
};
inline sampler2D make_sampler(const void* p, int w, int h, int format, int linear) {
    sampler2D s{}; s.level[0] = p; s.w[0] = w; s.h[0] = h; s.levels = 1; s.format = format; s.linear = linear; return s;
}
inline ivec2 textureSize(const sampler2D& s, int lod) { return ivec2(s.w[lod], s.h[lod]); }
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

inline vec4 texel_fetch(const sampler2D& s, int l, int x, int y) {
    size_t i = (size_t)y * s.w[l] + x;
    switch (s.format) {
    case FMT_R32F:    return vec4(((const float*)s.level[l])[i], 0.0f, 0.0f, 1.0f);
    case FMT_RGBA32F: { const float* t = (const float*)s.level[l] + i * 4; return vec4(t[0], t[1], t[2], t[3]); }
    case FMT_RGBA16F: { const uint16_t* t = (const uint16_t*)s.level[l] + i * 4;
                        return vec4(half_to_float(t[0]), half_to_float(t[1]), half_to_float(t[2]), half_to_float(t[3])); }
    default:          { const uint16_t* t = (const uint16_t*)s.level[l] + i * 2;
                        return vec4(half_to_float(t[0]), half_to_float(t[1]), 0.0f, 1.0f); }
    }
}
// Vulkan 1.3 §16.5-16.8: unnormalised coordinate u*W; NEAREST picks floor(u*W); LINEAR picks floor(u*W - 0.5) and
// its right / lower neighbour with the fractional part as weight; CLAMP_TO_EDGE clamps the texel indices.
inline int nearest_index(float u, int W) {
    float t = floorf(u * (float)W);
    if (!(t > -1.0f)) return 0;       // also NaN
    if (t >= (float)W) return W - 1;
    return (int)t;
}
inline void linear_index(float u, int W, int& i0, int& i1, float& f) {
    float x = u * (float)W - 0.5f;
    float fl = floorf(x);
    f = x - fl;
    if (!g_exact_weights) f = floorf(f * 256.0f + 0.5f) / 256.0f;
    int i = (int)fl;
    i0 = clampi(i, 0, W - 1);
    i1 = clampi(i + 1, 0, W - 1);
}
inline float lerp(float a, float b, float f) { return a * (1.0f - f) + b * f; }
inline vec4 lerp(vec4 a, vec4 b, float f) { return vec4(lerp(a.x, b.x, f), lerp(a.y, b.y, f), lerp(a.z, b.z, f), lerp(a.w, b.w, f)); }

inline vec4 textureLod(const sampler2D& s, vec2 uv, float lod) {
    int l = clampi((int)floorf(lod + 0.5f), 0, s.levels - 1);     // mipmapMode NEAREST
    if (!s.linear)
        return texel_fetch(s, l, nearest_index(uv.x, s.w[l]), nearest_index(uv.y, s.h[l]));
    int x0, x1, y0, y1; float fx, fy;
    linear_index(uv.x, s.w[l], x0, x1, fx);
    linear_index(uv.y, s.h[l], y0, y1, fy);
    vec4 top = lerp(texel_fetch(s, l, x0, y0), texel_fetch(s, l, x1, y0), fx);
    vec4 bot = lerp(texel_fetch(s, l, x0, y1), texel_fetch(s, l, x1, y1), fx);
    return lerp(top, bot, fy);
}
inline vec4 texture(const sampler2D& s, vec2 uv) { return textureLod(s, uv, 0.0f); }
// textureGather (§16.8.4 of the Vulkan spec): the 2x2 footprint LINEAR filtering would use, returned as
// (i0,j1), (i1,j1), (i1,j0), (i0,j0).
inline vec4 textureGather(const sampler2D& s, vec2 uv, int comp) {
    float x = uv.x * (float)s.w[0] - 0.5f, y = uv.y * (float)s.h[0] - 0.5f;
    int i = (int)floorf(x), j = (int)floorf(y);
    int i0 = clampi(i, 0, s.w[0] - 1), i1 = clampi(i + 1, 0, s.w[0] - 1);
    int j0 = clampi(j, 0, s.h[0] - 1), j1 = clampi(j + 1, 0, s.h[0] - 1);
    auto c = [&](int xx, int yy) { vec4 t = texel_fetch(s, 0, xx, yy); return comp == 0 ? t.x : comp == 1 ? t.y : comp == 2 ? t.z : t.w; };
    return vec4(c(i0, j1), c(i1, j1), c(i1, j0), c(i0, j0));
}
inline void imageStore(const image2D& im, ivec2 p, vec4 v) {
    if (p.x < 0 || p.y < 0 || p.x >= im.w || p.y >= im.h) return;
    size_t i = (size_t)p.y * im.w + p.x;
    if (im.format == FMT_R32F) ((float*)im.p)[i] = v.x;
    else if (im.format == FMT_RGBA32F) { float* t = (float*)im.p + i * 4; t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w; }
    else { uint16_t* t = (uint16_t*)im.p + i * 4;
           t[0] = float_to_half(v.x); t[1] = float_to_half(v.y); t[2] = float_to_half(v.z); t[3] = float_to_half(v.w); }
}

}  // namespace glsl
