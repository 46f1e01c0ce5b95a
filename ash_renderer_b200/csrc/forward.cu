// forward.cu — SURVEY §8.6 row f-1 (depth / HDR producers) + the tiled-forward light loop of row f-2.
//
// A compute forward pass that makes the two big inputs of the path on the device, so a view needs no
// 99.5 MB host upload: the reference produces them by rasterising meshes through Vulkan
// (renderer.rs:1804-1992, shaders/vert.vert + frag.frag), which is unavailable here, so the geometry is
// this repository's own: one occluder SPHERE per enabled cull object, inside the object's world AABB
// (consistent with what occlusion_cull.comp tests), ray-cast through the pixel centres.
//
//   k_forward_setup   thread i: object i -> view-space sphere + conservative pixel rectangle of its
//                     silhouette (tangent rays), classed small / big; light i -> view space;
//                     column i / row i -> ray table; thread 0 -> per-view constants
//   k_forward_clear   keys = (bits(1.0) << 32) | NO_OBJECT
//   k_splat_big       rectangles above kBigArea pixels, launched FIRST (big = near): every CTA walks the
//                     big list (staged through shared memory) and takes the 32x32-pixel chunks congruent
//                     to its index (a near sphere can cover a third of the screen)
//   k_splat_small     one warp per sphere, lanes 8x4 over the rectangle, two launches (mid, then small
//                     areas); per pixel: skip if the key already there is nearer than this sphere can
//                     be, else a ray / sphere test and a 64-bit atomicMin of (depth bits << 32 | object
//                     index): nearest hit, lowest index on ties, whatever the order — bit-exact against
//                     the serial oracle
//   k_forward_depth   depth buffer = high halves of the keys (what set_depth would have uploaded)
//   k_forward_shade   one CTA per 16x16 light tile: tile list -> shared memory, then per pixel
//                     frag.frag's Cook-Torrance BRDF (:89-114, :167-213) for the MVP light and every
//                     light of the list; RGBA16F out (what set_color would have uploaded)
//
// Arithmetic: an explicit IEEE sequence with a fixed evaluation order (PIN-1), which the CPU checker restates.
// Roofline: the splat is bound by the per-pixel test rate (~35 fp32 instructions, one IEEE sqrt and
// two IEEE divides per covered pixel and sphere) and the shade kernel by the BRDF (~170 instructions
// per pixel and light); neither is HBM-bound (8 B key traffic per test hits in L2 for overlapping
// spheres).  Algorithmic bytes: depth 8 B clear + 8 B resolve read + 4 B write, shade 8 B key + 8 B out.
#include <cstring>
#include "ash_internal.cuh"
#include "ash_math.cuh"

namespace ash {

constexpr int kBigArea = 16384;         // pixels: above this a rectangle goes to the CTA-cooperative kernel
constexpr int kMidArea = 2048;          // warp-per-sphere launches: (kMidArea, kBigArea] first, then the rest (near before far)
constexpr int kChunk = 32;              // k_splat_big works in 32x32-pixel chunks

struct FwdConsts {                      // written by k_forward_setup thread 0, read by the other kernels
    float P22, P32, near_plane; uint32_t W, H, object_count;
    float sun[3];                       // unit vector towards the MVP light, view space
};

__device__ __forceinline__ float ray_coord(int i, int S, float P) {
    return fdiv(fsub(fmul(fdiv(fadd((float)i, 0.5f), (float)S), 2.0f), 1.0f), P);
}

// per-axis pixel bound of the part of the sphere in front of the near plane, padded by a pixel.
// tangent: the eye plane does not cut the sphere (cz > r): slopes x/z of the two tangent rays.
// otherwise the box bound of { x in [c-r, c+r], z in [near, cz+r] }.
__device__ __forceinline__ void axis_bound(float c, float cz, float r, bool tangent, float near_plane, float P, int S, int* i0, int* i1) {
    float s0, s1;
    if (tangent) {
        const float den = fmul(fsub(cz, r), fadd(cz, r));
        const float D = fsub(fadd(fmul(c, c), fmul(cz, cz)), fmul(r, r));
        const float sq = fsqrt(gmax(D, 0.0f));
        const float m = fmul(c, cz), k = fmul(r, sq);
        s0 = fdiv(fsub(m, k), den); s1 = fdiv(fadd(m, k), den);
    } else {
        const float zmax = fadd(cz, r), lo_c = fsub(c, r), hi_c = fadd(c, r);
        s0 = lo_c < 0.0f ? fdiv(lo_c, near_plane) : fdiv(lo_c, zmax);
        s1 = hi_c > 0.0f ? fdiv(hi_c, near_plane) : fdiv(hi_c, zmax);
    }
    const float n0 = fmul(P, s0), n1 = fmul(P, s1);
    const float p0 = fmul(fadd(fmul(n0, 0.5f), 0.5f), (float)S), p1 = fmul(fadd(fmul(n1, 0.5f), 0.5f), (float)S);
    const float lo = gmin(p0, p1), hi = gmax(p0, p1);
    if (!(lo <= hi)) { *i0 = 0; *i1 = S; return; }
    *i0 = (int)gclamp(fsub(floorf(lo), 1.0f), 0.0f, (float)S);
    *i1 = (int)gclamp(fadd(ceilf(hi), 1.0f), 0.0f, (float)S);
}

// nearest intersection of the ray d = (dx, dy, -1) with the sphere (c, r); < 0: miss
__device__ __forceinline__ float hit_t(float4 s, float cc, float dx, float dy, float near_plane) {
    const float a = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), 1.0f);
    const float b = fsub(fadd(fmul(dx, s.x), fmul(dy, s.y)), s.z);
    const float disc = fsub(fmul(b, b), fmul(a, cc));
    if (!(disc >= 0.0f)) return -1.0f;
    const float t = fdiv(fsub(b, fsqrt(disc)), a);
    return (t >= near_plane) ? t : -1.0f;
}
__device__ __forceinline__ float sphere_cc(float4 s) { return fsub(dot3(v3(s.x, s.y, s.z), v3(s.x, s.y, s.z)), fmul(s.w, s.w)); }

__global__ void __launch_bounds__(256)
k_forward_setup(AshCullingCameraData cam, AshForwardParams fp, const AshCullObjectData* __restrict__ objs, uint32_t n,
                const AshGpuLight* __restrict__ lights, uint32_t n_lights, uint32_t W, uint32_t H,
                float4* __restrict__ spheres, int4* __restrict__ rects, uint32_t* __restrict__ big_list,
                uint32_t* __restrict__ big_count, float4* __restrict__ lights_view,
                float* __restrict__ ray_x, float* __restrict__ ray_y, FwdConsts* __restrict__ consts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const M4 view = load_m4(cam.view);
    if (i < n) {
        const AshCullObjectData& o = objs[i];
        float4 s = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        int4 q = make_int4(0, 0, 0, 0);
        if (o.flags & 1u) {
            const M4 model = load_m4(o.model);
            const V4 wc = mul4(model, o.bounds.center[0], o.bounds.center[1], o.bounds.center[2], 1.0f);
            const float h0 = fmul(o.bounds.extents[0], length3(v3(model.c[0].x, model.c[0].y, model.c[0].z)));
            const float h1 = fmul(o.bounds.extents[1], length3(v3(model.c[1].x, model.c[1].y, model.c[1].z)));
            const float h2 = fmul(o.bounds.extents[2], length3(v3(model.c[2].x, model.c[2].y, model.c[2].z)));
            const float r = fmul(fp.radius_scale, gmin(gmin(h0, h1), h2));
            const V4 cv = mul4(view, wc.x, wc.y, wc.z, 1.0f);
            s = make_float4(cv.x, cv.y, cv.z, r);
            const float cz = -cv.z;
            if (r > 0.0f && !(fadd(cz, r) < fp.near_plane)) {
                const bool tangent = cz > r;
                int x0, x1, y0, y1;
                axis_bound(cv.x, cz, r, tangent, fp.near_plane, cam.projection[0][0], (int)W, &x0, &x1);
                axis_bound(cv.y, cz, r, tangent, fp.near_plane, cam.projection[1][1], (int)H, &y0, &y1);
                // a lower bound of every depth this sphere can produce (hits are at view distance >= cz - r; 0.2 % and
                // 1e-6 cover the rounding of the quadratic and of the depth formula): pixels already nearer skip the test
                float lb = 0.0f;
                const float t_lb = fmul(fsub(cz, r), 0.998f);
                if (t_lb > fp.near_plane && cam.projection[3][2] < 0.0f)
                    lb = gmax(fsub(fdiv(fsub(cam.projection[3][2], fmul(cam.projection[2][2], t_lb)), t_lb), 1e-6f), 0.0f);
                q = make_int4((int)((uint32_t)x0 | ((uint32_t)x1 << 16)), (int)((uint32_t)y0 | ((uint32_t)y1 << 16)), (int)__float_as_uint(lb), max(x1 - x0, 0) * max(y1 - y0, 0));
            }
        }
        spheres[i] = s;
        rects[i] = q;                        // {x0 | x1 << 16, y0 | y1 << 16, depth lower bound bits, area}
        if (q.w > kBigArea) big_list[atomicAdd(big_count, 1u)] = i;   // order is irrelevant (atomicMin)
    }
    if (i < n_lights) {
        const AshGpuLight& l = lights[i];
        if (l.direction[3] == 2.0f) {
            const V4 d = mul4(view, l.direction[0], l.direction[1], l.direction[2], 0.0f);
            const V3 L = normalize3(v3(-d.x, -d.y, -d.z));
            lights_view[i] = make_float4(L.x, L.y, L.z, 0.0f);
        } else {
            const V4 p = mul4(view, l.position[0], l.position[1], l.position[2], 1.0f);
            lights_view[i] = make_float4(p.x, p.y, p.z, l.position[3]);
        }
    }
    if (i < W) ray_x[i] = ray_coord((int)i, (int)W, cam.projection[0][0]);
    if (i < H) ray_y[i] = ray_coord((int)i, (int)H, cam.projection[1][1]);
    if (i == 0) {
        const V4 sd = mul4(view, fp.light_direction[0], fp.light_direction[1], fp.light_direction[2], 0.0f);
        const V3 L = normalize3(v3(-sd.x, -sd.y, -sd.z));
        FwdConsts k;
        k.P22 = cam.projection[2][2]; k.P32 = cam.projection[3][2]; k.near_plane = fp.near_plane;
        k.W = W; k.H = H; k.object_count = n;
        k.sun[0] = L.x; k.sun[1] = L.y; k.sun[2] = L.z;
        *consts = k;
    }
}

__global__ void __launch_bounds__(256) k_forward_clear(ulonglong2* __restrict__ keys2, size_t pairs, unsigned long long* __restrict__ keys, size_t n) {
    const unsigned long long clear = ((unsigned long long)0x3F800000u << 32) | ASH_FWD_NO_OBJECT;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < pairs; i += (size_t)gridDim.x * blockDim.x)
        keys2[i] = make_ulonglong2(clear, clear);
    if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) keys[n - 1] = clear;
}

__device__ __forceinline__ void splat_pixel(float4 s, float cc, uint32_t lb_bits, uint32_t id, int x, int y, const FwdConsts& k,
                                            const float* __restrict__ ray_x, const float* __restrict__ ray_y,
                                            unsigned long long* __restrict__ keys) {
    unsigned long long* p = keys + (size_t)y * k.W + x;
    const unsigned long long cur = *p;                  // plain read: keys only decrease, a stale value only costs work
    if ((uint32_t)(cur >> 32) < lb_bits) return;        // something nearer than this sphere can be is already there
    const float t = hit_t(s, cc, __ldg(ray_x + x), __ldg(ray_y + y), k.near_plane);
    if (t < 0.0f) return;
    const float d = fadd(fdiv(fsub(k.P32, fmul(k.P22, t)), t), 0.0f);              // ndc z of view z = -t; -0 -> +0
    if (!(d >= 0.0f && d <= 1.0f)) return;
    const unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | id;
    if (key < cur) atomicMin(p, key);
}

// one warp per sphere whose rectangle area is in (area_lo, area_hi]
__global__ void __launch_bounds__(256)
k_splat_small(const float4* __restrict__ spheres, const int4* __restrict__ rects, const FwdConsts* __restrict__ consts,
              const float* __restrict__ ray_x, const float* __restrict__ ray_y, unsigned long long* __restrict__ keys,
              int area_lo, int area_hi, const uint32_t* __restrict__ mask) {
    const FwdConsts k = *consts;
    const uint32_t id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (id >= k.object_count) return;
    if (mask && !__ldg(mask + id)) return;                 // two-phase culling: only the objects of this phase
    const int4 q = __ldg(rects + id);
    if (q.w <= area_lo || q.w > area_hi) return;
    const int x0 = q.x & 0xFFFF, x1 = (int)((uint32_t)q.x >> 16), y0 = q.y & 0xFFFF, y1 = (int)((uint32_t)q.y >> 16);
    const float4 s = __ldg(spheres + id);
    const float cc = sphere_cc(s);
    const int lane = threadIdx.x & 31, lx = lane & 7, ly = lane >> 3;
    for (int y = y0 + ly; y < y1; y += 4)
        for (int x = x0 + lx; x < x1; x += 8) splat_pixel(s, cc, (uint32_t)q.z, id, x, y, k, ray_x, ray_y, keys);
}

// rectangles above kBigArea: the list is staged through shared memory 256 entries at a time; chunk c of
// entry b belongs to CTA (c + 7 b) mod grid, so that spheres of a few chunks spread over the whole grid
__global__ void __launch_bounds__(256)
k_splat_big(const float4* __restrict__ spheres, const int4* __restrict__ rects, const uint32_t* __restrict__ big_list,
            const uint32_t* __restrict__ big_count, const FwdConsts* __restrict__ consts,
            const float* __restrict__ ray_x, const float* __restrict__ ray_y, unsigned long long* __restrict__ keys,
            const uint32_t* __restrict__ mask) {
    __shared__ int4 s_q[256];
    __shared__ float4 s_s[256];
    __shared__ uint32_t s_id[256];
    const FwdConsts k = *consts;
    const uint32_t nb = *big_count;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (uint32_t b0 = 0; b0 < nb; b0 += 256) {
        __syncthreads();
        if (b0 + threadIdx.x < nb) {
            const uint32_t id = big_list[b0 + threadIdx.x];
            int4 q = __ldg(rects + id);
            if (mask && !__ldg(mask + id)) q = make_int4(0, 0, 0, 0);          // not in this phase: no chunks
            s_id[threadIdx.x] = id; s_q[threadIdx.x] = q; s_s[threadIdx.x] = __ldg(spheres + id);
        }
        __syncthreads();
        const uint32_t nloc = min(256u, nb - b0);
        for (uint32_t j = 0; j < nloc; ++j) {
            const int4 q = s_q[j];
            const int x0 = q.x & 0xFFFF, x1 = (int)((uint32_t)q.x >> 16), y0 = q.y & 0xFFFF, y1 = (int)((uint32_t)q.y >> 16);
            const int ncx = (x1 - x0 + kChunk - 1) / kChunk, ncy = (y1 - y0 + kChunk - 1) / kChunk;
            const uint32_t first = (blockIdx.x + gridDim.x - (7u * (b0 + j)) % gridDim.x) % gridDim.x;
            if (first >= (uint32_t)(ncx * ncy)) continue;
            const float4 s = s_s[j];
            const float cc = sphere_cc(s);
            for (uint32_t c = first; c < (uint32_t)(ncx * ncy); c += gridDim.x) {
                const int x = x0 + (int)(c % ncx) * kChunk + tx, yy = y0 + (int)(c / ncx) * kChunk + ty;
                if (x >= x1) continue;
#pragma unroll
                for (int r = 0; r < kChunk; r += 8)
                    if (yy + r < y1) splat_pixel(s, cc, (uint32_t)q.z, s_id[j], x, yy + r, k, ray_x, ray_y, keys);
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_forward_depth(const uint2* __restrict__ keys, float* __restrict__ depth, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        depth[i] = __uint_as_float(keys[i].y);      // little endian: .y = high half = depth bits
}

// ---------------------------------------------------------------------------------- shading (frag.frag)
struct Material { V3 base; float metallic, roughness; };

__device__ __forceinline__ V3 brdf(V3 N, V3 V, V3 L, const Material& m, V3 lightColor) {
    const float PI_F = 3.14159265359f;                                                  // frag.frag:45
    const float NdotL = gmax(dot3(N, L), 0.0f);                                         // :167
    const float im = fsub(1.0f, m.metallic);
    const float f4 = fmul(0.04f, im);
    const V3 F0 = {fadd(f4, fmul(m.base.x, m.metallic)), fadd(f4, fmul(m.base.y, m.metallic)), fadd(f4, fmul(m.base.z, m.metallic))};   // :186
    const V3 Hh = normalize3(add3(V, L));                                               // :188
    const float NdotV = gmax(dot3(N, V), 0.001f);
    const float NdotH = gmax(dot3(N, Hh), 0.0f);
    const float VdotH = gmax(dot3(V, Hh), 0.0f);
    const float a = fmul(m.roughness, m.roughness), a2 = fmul(a, a);                    // :89-94
    const float dd = fadd(fmul(fmul(NdotH, NdotH), fsub(a2, 1.0f)), 1.0f);
    const float D = fdiv(a2, fmul(fmul(PI_F, dd), dd));
    const float r1 = fadd(m.roughness, 1.0f), k = fmul(fmul(r1, r1), 0.125f);           // :101-105
    const float ik = fsub(1.0f, k);
    const float G = fmul(fdiv(NdotV, fadd(fmul(NdotV, ik), k)), fdiv(NdotL, fadd(fmul(NdotL, ik), k)));
    const float t = gclamp(fsub(1.0f, VdotH), 0.0f, 1.0f), t2 = fmul(t, t), t5 = fmul(fmul(t2, t2), t);   // :108-114
    const V3 F = {fadd(F0.x, fmul(fsub(1.0f, F0.x), t5)), fadd(F0.y, fmul(fsub(1.0f, F0.y), t5)), fadd(F0.z, fmul(fsub(1.0f, F0.z), t5))};
    const float DG = fmul(D, G);
    const float denom = fadd(fmul(fmul(4.0f, NdotV), NdotL), 0.001f);                   // :198
    V3 spec = {fdiv(fmul(DG, F.x), denom), fdiv(fmul(DG, F.y), denom), fdiv(fmul(DG, F.z), denom)};
    spec = {gmin(spec.x, fdiv(10.0f, gmax(0.04f, F0.x))), gmin(spec.y, fdiv(10.0f, gmax(0.04f, F0.y))),
            gmin(spec.z, fdiv(10.0f, gmax(0.04f, F0.z)))};                              // :202
    const V3 kD = {fmul(fsub(1.0f, F.x), im), fmul(fsub(1.0f, F.y), im), fmul(fsub(1.0f, F.z), im)};   // :204
    const V3 diffuse = divs3(mul3(kD, m.base), PI_F);                                   // :205
    return scale3(mul3(add3(diffuse, spec), lightColor), NdotL);                        // :213, shadow = 0
}

struct TileLight { float4 v; float4 rad; };       // view position + radius (or unit L), rgb * intensity + type

__global__ void __launch_bounds__(256)
k_forward_shade(const unsigned long long* __restrict__ keys, const float4* __restrict__ spheres,
                const AshGpuLight* __restrict__ lights, uint32_t n_lights, const float4* __restrict__ lights_view,
                const uint32_t* __restrict__ tile_data, const AshInstanceData* __restrict__ inst, uint32_t n_inst,
                const FwdConsts* __restrict__ consts, const float* __restrict__ ray_x, const float* __restrict__ ray_y,
                AshForwardParams fp, uint2* __restrict__ hdr) {
    __shared__ TileLight s_l[256];
    __shared__ uint32_t s_cnt;
    const FwdConsts k = *consts;
    const bool use_lights = (fp.flags & ASH_FWD_TILE_LIGHTS) != 0;
    uint32_t cnt = 0;
    if (use_lights) {
        const uint32_t* td = tile_data + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 257;
        cnt = min(__ldg(td), 256u);
        if (threadIdx.x < cnt) {
            const uint32_t li = __ldg(td + 1 + threadIdx.x);
            TileLight tl;
            tl.v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            tl.rad = make_float4(0.0f, 0.0f, 0.0f, -1.0f);                           // type -1: skip (index out of range)
            if (li < n_lights) {
                const float4 c = __ldg(reinterpret_cast<const float4*>(lights[li].color));
                tl.v = __ldg(lights_view + li);
                tl.rad = make_float4(fmul(c.x, c.w), fmul(c.y, c.w), fmul(c.z, c.w), __ldg(&lights[li].direction[3]));
            }
            s_l[threadIdx.x] = tl;
        }
        if (threadIdx.x == 0) s_cnt = cnt;
        __syncthreads();
        cnt = s_cnt;
    }
    const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 16 + (threadIdx.x >> 4);
    if (x >= (int)k.W || y >= (int)k.H) return;
    const size_t p = (size_t)y * k.W + x;
    const uint32_t id = (uint32_t)(keys[p] & 0xFFFFFFFFull);
    const float dx = __ldg(ray_x + x), dy = __ldg(ray_y + y);
    float t = -1.0f;
    float4 s = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (id < k.object_count) { s = __ldg(spheres + id); t = hit_t(s, sphere_cc(s), dx, dy, k.near_plane); }
    if (t < 0.0f) { hdr[p] = pack_rgba16f(fp.background[0], fp.background[1], fp.background[2], 1.0f); return; }
    const V3 pos = {fmul(dx, t), fmul(dy, t), -t};
    const V3 N = normalize3(sub3(pos, v3(s.x, s.y, s.z)));
    const V3 V = normalize3(v3(-pos.x, -pos.y, -pos.z));                             // :120, camera at the origin
    Material m;
    m.base = v3(fp.base_color_factor[0], fp.base_color_factor[1], fp.base_color_factor[2]);   // :128
    if ((fp.flags & ASH_FWD_INSTANCE_COLOR) && inst && id < n_inst) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(inst[id].color));
        m.base = mul3(m.base, v3(c.x, c.y, c.z));
    }
    m.metallic = fp.parameters[0];
    m.roughness = gmax(fp.parameters[1], 0.04f);                                     // :171
    const V3 ambient = mul3(v3(fp.ambient_color[0], fp.ambient_color[1], fp.ambient_color[2]), m.base);   // :216
    const V3 Lo = brdf(N, V, v3(k.sun[0], k.sun[1], k.sun[2]), m, v3(fp.light_color[0], fp.light_color[1], fp.light_color[2]));
    V3 color = add3(add3(ambient, Lo), v3(fp.emissive_factor[0], fp.emissive_factor[1], fp.emissive_factor[2]));   // :224
    for (uint32_t i = 0; i < cnt; ++i) {
        const TileLight tl = s_l[i];
        V3 rad = v3(tl.rad.x, tl.rad.y, tl.rad.z), L;
        if (tl.rad.w == 2.0f) L = v3(tl.v.x, tl.v.y, tl.v.z);
        else if (tl.rad.w < 0.0f) continue;
        else {
            const V3 toL = sub3(v3(tl.v.x, tl.v.y, tl.v.z), pos);
            const float d2 = dot3(toL, toL);
            if (!(d2 > 0.0f)) continue;
            const float dist = fsqrt(d2);
            L = divs3(toL, dist);
            const float q = fdiv(dist, tl.v.w), q2 = fmul(q, q);                      // windowed inverse square (ours)
            const float w = gclamp(fsub(1.0f, fmul(q2, q2)), 0.0f, 1.0f);
            rad = scale3(rad, fdiv(fmul(w, w), fadd(d2, 1.0f)));
        }
        color = add3(color, brdf(N, V, L, m, rad));
    }
    if (fp.flags & ASH_FWD_REINHARD) color = div3(color, adds3(color, 1.0f));        // :226
    hdr[p] = pack_rgba16f(color.x, color.y, color.z, 1.0f);
}

// ---------------------------------------------------------------------------------- velocity (RG16F)
// The motion vector taa_resolve.comp:94-102 consumes: velocity = uv - prevUV, so that historyUV = uv - velocity is where
// this pixel's surface point was on screen in the previous frame.  The scene is static (cull objects carry no previous
// model matrix), so the point's previous clip position is reproj * (pos_view, 1) with reproj = prevViewProj * view^-1,
// built by the caller like viewProj (float64, rounded once).  Pixels without a hit, or whose point was behind the previous
// camera, get velocity 0 — the shader then falls back to the jitter delta (:97-99).  One thread per pixel; the hit is the
// shade kernel's (same sphere, same ray tables, same t bits).
struct Reproj { float m[4][4]; };
__global__ void __launch_bounds__(256)
k_forward_velocity(const unsigned long long* __restrict__ keys, const float4* __restrict__ spheres,
                   const FwdConsts* __restrict__ consts, const float* __restrict__ ray_x, const float* __restrict__ ray_y,
                   Reproj rp, __half2* __restrict__ vel) {
    const FwdConsts k = *consts;
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= (int)k.W || y >= (int)k.H) return;
    const size_t p = (size_t)y * k.W + x;
    const uint32_t id = (uint32_t)(keys[p] & 0xFFFFFFFFull);
    float vx = 0.0f, vy = 0.0f;
    if (id < k.object_count) {
        const float dx = __ldg(ray_x + x), dy = __ldg(ray_y + y);
        const float4 s = __ldg(spheres + id);
        const float t = hit_t(s, sphere_cc(s), dx, dy, k.near_plane);
        if (t >= 0.0f) {
            const M4 m = load_m4(rp.m);
            const V4 clip = mul4(m, fmul(dx, t), fmul(dy, t), -t, 1.0f);
            if (clip.w > 0.0f) {
                const float pu = fadd(fmul(fdiv(clip.x, clip.w), 0.5f), 0.5f), pv = fadd(fmul(fdiv(clip.y, clip.w), 0.5f), 0.5f);
                vx = fsub(fdiv(fadd((float)x, 0.5f), (float)k.W), pu);
                vy = fsub(fdiv(fadd((float)y, 0.5f), (float)k.H), pv);
            }
        }
    }
    vel[p] = __floats2half2_rn(vx, vy);
}

void forward_velocity_patch(const float reproj[4][4], PatchArg* out) {             // graph replay: argument 5 is the matrix
    out->func = reinterpret_cast<const void*>(&k_forward_velocity);
    out->arg_index = 5; out->n_args = 7;
    std::memcpy(out->value, reproj, sizeof(Reproj));
}

cudaError_t launch_forward_velocity(const unsigned long long* keys, const float4* spheres, const void* consts, const float* ray_x,
                                    const float* ray_y, const float reproj[4][4], uint32_t W, uint32_t H, void* velocity, const Launch& L) {
    Reproj rp;
    std::memcpy(rp.m, reproj, sizeof(rp.m));
    Timed timed(L, ASH_KERNEL_FORWARD_SHADE);
    k_forward_velocity<<<dim3((W + 31) / 32, (H + 7) / 8), 256, 0, L.s>>>(keys, spheres, static_cast<const FwdConsts*>(consts), ray_x, ray_y,
                                                                            rp, static_cast<__half2*>(velocity));
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------- launchers
size_t forward_consts_bytes() { return sizeof(FwdConsts); }

void forward_setup_patch(const AshCullingCameraData& cam, PatchArg* out) {      // graph replay: argument 0 is the camera
    static_assert(sizeof(AshCullingCameraData) <= sizeof(out->value), "PatchArg too small");
    out->func = reinterpret_cast<const void*>(&k_forward_setup);
    out->arg_index = 0; out->n_args = 16;
    std::memcpy(out->value, &cam, sizeof(cam));
}

// setup + key clear | splat (optionally only the objects whose mask is set; may be called again to add
// objects to the same keys) | resolve keys -> depth.  One ASH_KERNEL_FORWARD_DEPTH span each.
cudaError_t launch_forward_setup(const AshCullingCameraData& cam, const AshForwardParams& fp, const AshCullObjectData* objs, uint32_t n,
                                 const AshGpuLight* lights, uint32_t n_lights, uint32_t W, uint32_t H, const ForwardBuffers& B, const Launch& L) {
    Timed timed(L, ASH_KERNEL_FORWARD_DEPTH);
    cudaError_t e = cudaMemsetAsync(B.big_count, 0, 4, L.s);
    if (e != cudaSuccess) return e;
    const uint32_t span = max(max(n, n_lights), max(W, H));
    const size_t P = (size_t)W * H;
    k_forward_setup<<<(span + 255) / 256, 256, 0, L.s>>>(cam, fp, objs, n, lights, n_lights, W, H, B.spheres, B.rects, B.big_list, B.big_count,
                                                         B.lights_view, B.ray_x, B.ray_y, static_cast<FwdConsts*>(B.consts));
    k_forward_clear<<<148 * 8, 256, 0, L.s>>>(reinterpret_cast<ulonglong2*>(B.keys), P / 2, B.keys, P);
    *L.launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_forward_splat(uint32_t n, uint32_t W, uint32_t H, const ForwardBuffers& B, const uint32_t* mask, const Launch& L) {
    if (!n) return cudaSuccess;
    Timed timed(L, ASH_KERNEL_FORWARD_DEPTH);
    // big (near) spheres first: what they cover lets the far ones skip their tests (same result in any order)
    const FwdConsts* kc = static_cast<const FwdConsts*>(B.consts);
    k_splat_big<<<148 * 4, 256, 0, L.s>>>(B.spheres, B.rects, B.big_list, B.big_count, kc, B.ray_x, B.ray_y, B.keys, mask);
    k_splat_small<<<(n + 7) / 8, 256, 0, L.s>>>(B.spheres, B.rects, kc, B.ray_x, B.ray_y, B.keys, kMidArea, kBigArea, mask);
    k_splat_small<<<(n + 7) / 8, 256, 0, L.s>>>(B.spheres, B.rects, kc, B.ray_x, B.ray_y, B.keys, 0, kMidArea, mask);
    *L.launches += 2;
    (void)W; (void)H;
    return cudaGetLastError();
}

cudaError_t launch_forward_resolve(uint32_t W, uint32_t H, const ForwardBuffers& B, const Launch& L) {
    Timed timed(L, ASH_KERNEL_FORWARD_DEPTH);
    k_forward_depth<<<148 * 8, 256, 0, L.s>>>(reinterpret_cast<const uint2*>(B.keys), B.depth, (size_t)W * H);
    return cudaGetLastError();
}

cudaError_t launch_forward_shade(const AshForwardParams& fp, const unsigned long long* keys, const float4* spheres,
                                 const AshGpuLight* lights, uint32_t n_lights, const float4* lights_view, const uint32_t* tile_data,
                                 const AshInstanceData* inst, uint32_t n_inst, const void* consts, const float* ray_x,
                                 const float* ray_y, uint32_t W, uint32_t H, void* hdr, const Launch& L) {
    Timed timed(L, ASH_KERNEL_FORWARD_SHADE);
    const dim3 grid((W + 15) / 16, (H + 15) / 16);
    k_forward_shade<<<grid, 256, 0, L.s>>>(keys, spheres, lights, n_lights, lights_view, tile_data, inst, n_inst,
                                           static_cast<const FwdConsts*>(consts), ray_x, ray_y, fp, static_cast<uint2*>(hdr));
    return cudaGetLastError();
}

}  // namespace ash
