// ref_shaders.cpp — TEST INFRASTRUCTURE: oracle/_ref/libashref.so = the REFERENCE'S OWN shader sources
// (/root/reference/shaders/*.comp, *.frag, read where they lie, passed through oracle/ref/strip_glsl.py into
// oracle/_ref/gen/*.inc, never committed) compiled as C++ behind oracle/ref/glsl_compat.hpp, plus this harness,
// which plays the part of the Vulkan host code: binds buffers / samplers / images, fills push constants,
// iterates the dispatch grid or the fullscreen triangle's pixels, applies the fixed-function output stage
// (RGBA16F store, additive blend of PIN-9, UNORM8 conversion of PIN-11).
//
// No statement of shader logic lives in this file.  Entry points mirror oracle/ash_oracle.cpp (ash_oracle_* ->
// ash_ref_*) so tests/test_oracle_ref.py can compare the two bit for bit; only tests/ and the golden-vector
// generator load this library.
#include <ucontext.h>
#include <vector>
#include "glsl_compat.hpp"
#include "../../include/ashcuda.h"

// ---- GLSL storage / interface qualifiers on the unmodified token stream
#define uniform struct
#define readonly
#define writeonly
#define shared static
#define main shader_main

namespace glsl {

// compute shaders: `layout(local_size...) in;` leaves `in;`, `out` only qualifies array parameters
#define in
#define out
namespace sh_hiz {
#include "gen/hiz_generate.comp.inc"
}
namespace sh_cull {
#include "gen/occlusion_cull.comp.inc"
}
namespace sh_light {
#include "gen/light_culling.comp.inc"
}
namespace sh_taa {
#include "gen/taa_resolve.comp.inc"
}
#undef in
#undef out

// fragment shaders: `in vec2 fragTexCoord;` / `out vec4 outColor;` are per-invocation variables
#define in thread_local
#define out thread_local
namespace sh_threshold {
#include "gen/bloom_threshold.frag.inc"
}
namespace sh_down {
#include "gen/bloom_downsample.frag.inc"
}
namespace sh_up {
#include "gen/bloom_upsample.frag.inc"
}
namespace sh_tonemap {
#include "gen/tonemapping.frag.inc"
}
#undef in
#undef out

}  // namespace glsl
#undef uniform
#undef readonly
#undef writeonly
#undef shared
#undef main

using namespace glsl;

static_assert(sizeof(sh_cull::ObjectData) == sizeof(AshCullObjectData), "ObjectData layout (occlusion_cull.comp:19-26)");
static_assert(sizeof(sh_cull::DrawCommand) == sizeof(AshIndirectDrawCommand), "DrawCommand layout (:28-34)");
static_assert(sizeof(sh_cull::pc) == sizeof(AshCullingPushConstants), "cull push constants (:69-75)");
static_assert(sizeof(sh_light::Light) == sizeof(AshGpuLight), "Light layout (light_culling.comp:18-23)");
static_assert(sizeof(sh_light::pc) == sizeof(AshLightCullingPushConstants), "light push constants (:26-30)");
static_assert(sizeof(sh_light::camera) == sizeof(AshCullingCameraData), "CameraData (:48-53)");
static_assert(sizeof(sh_taa::pc) == sizeof(AshTaaPushConstants), "TAA push constants (taa_resolve.comp:22-27)");
static_assert(sizeof(sh_tonemap::pc) == sizeof(AshPostProcessPushConstants), "post push constants (tonemapping.frag:12-17)");

namespace {

inline int div_up(int a, int b) { return (a + b - 1) / b; }

// one fullscreen pass: fragTexCoord at the pixel centre (postprocess.vert:8-14 interpolated), RGBA16F target
template <class Set, class Main, class Get>
void fullscreen_rgba16f(int W, int H, uint16_t* dst, bool additive, Set set_uv, Main shader, Get get_color, int f32 = 0) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            set_uv(vec2(((float)x + 0.5f) / (float)W, ((float)y + 0.5f) / (float)H));
            shader();
            vec4 c = get_color();
            if (f32) {                                     // RGBA32F target (fp32_intermediates): the same blend, no rounding
                float* t = (float*)dst + ((size_t)y * W + x) * 4;
                if (additive) { c.x = c.x + t[0]; c.y = c.y + t[1]; c.z = c.z + t[2]; }
                t[0] = c.x; t[1] = c.y; t[2] = c.z; t[3] = c.w;
                continue;
            }
            uint16_t* t = dst + ((size_t)y * W + x) * 4;
            if (additive) {   // PIN-9: VK_BLEND_OP_ADD, ONE / ONE on colour; alpha written from the source
                c.x = c.x + half_to_float(t[0]); c.y = c.y + half_to_float(t[1]); c.z = c.z + half_to_float(t[2]);
            }
            t[0] = float_to_half(c.x); t[1] = float_to_half(c.y); t[2] = float_to_half(c.z); t[3] = float_to_half(c.w);
        }
}

const uint16_t kZeroTexel[4] = {0, 0, 0, 0};

// ---- fibres: the 256 invocations of one light-culling work group, resumed round-robin in ascending
// gl_LocalInvocationIndex between barrier()s
constexpr int kMaxFibres = 256;
constexpr size_t kStack = 64 * 1024;
ucontext_t g_sched, g_fibre[kMaxFibres];
bool g_done[kMaxFibres];
int g_cur = 0;
void (*g_entry)() = nullptr;

void fibre_yield() { swapcontext(&g_fibre[g_cur], &g_sched); }
void fibre_main() {
    g_entry();
    g_done[g_cur] = true;
    swapcontext(&g_fibre[g_cur], &g_sched);
}
// runs `entry` once per invocation of a (lx x ly) work group; set_ids(i) installs the gl_* built-ins
template <class SetIds>
void run_group(int n, std::vector<char>& stacks, void (*entry)(), SetIds set_ids) {
    g_entry = entry;
    for (int i = 0; i < n; ++i) {
        getcontext(&g_fibre[i]);
        g_fibre[i].uc_stack.ss_sp = stacks.data() + (size_t)i * kStack;
        g_fibre[i].uc_stack.ss_size = kStack;
        g_fibre[i].uc_link = nullptr;
        makecontext(&g_fibre[i], fibre_main, 0);
        g_done[i] = false;
    }
    for (bool any = true; any;) {
        any = false;
        for (int i = 0; i < n; ++i) {
            if (g_done[i]) continue;
            g_cur = i;
            set_ids(i);
            swapcontext(&g_sched, &g_fibre[i]);
            any = any || !g_done[i];
        }
    }
}

}  // namespace

extern "C" {

void ash_ref_set_weight_mode(int exact) { g_exact_weights = exact ? 1 : 0; }

// shaders/hiz_generate.comp, one dispatch = one mip (PIN-3/4/5); local size 8x8
void ash_ref_hiz_level(const float* src, int sw, int sh, float* dst, int dw, int dh, int mip_level) {
    sh_hiz::inputDepth = make_sampler(src, sw, sh, FMT_R32F, 0);
    sh_hiz::outputMip = image2D{dst, dw, dh, FMT_R32F};
    sh_hiz::pc.outputSize = uvec2((uint)dw, (uint)dh);
    sh_hiz::pc.mipLevel = (uint)mip_level;
    const int gy = div_up(dh, 8) * 8, gx = div_up(dw, 8) * 8;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < gy; ++y)
        for (int x = 0; x < gx; ++x) {
            gl_GlobalInvocationID = uvec3((uint)x, (uint)y, 0u);
            sh_hiz::shader_main();
        }
}

size_t ash_ref_hiz_build(const float* depth, int W, int H, float* mips) {
    int m = W > H ? W : H, levels = 1;
    while (m > 1) { m >>= 1; ++levels; }                       // PIN-4 (host side, occlusion_culling.rs:21,144)
    const float* src = depth; int sw = W, sh = H; size_t off = 0;
    for (int l = 1; l < levels; ++l) {
        int dw = sw / 2 > 0 ? sw / 2 : 1, dh = sh / 2 > 0 ? sh / 2 : 1;   // PIN-3: Vulkan mip extents
        ash_ref_hiz_level(src, sw, sh, mips + off, dw, dh, l);
        src = mips + off; off += (size_t)dw * dh; sw = dw; sh = dh;
    }
    return off;
}

// shaders/occlusion_cull.comp; local size 64; invocations run serially in ascending index, which makes the
// atomicAdd slot order the canonical one (PIN-7)
uint32_t ash_ref_occlusion_cull(const AshCullObjectData* objs, const AshIndirectDrawCommand* draws_in,
                                const AshCullingPushConstants* pc, const float* depth, const float* mips,
                                int W, int H, uint32_t* visibility, AshIndirectDrawCommand* draws_out) {
    using namespace sh_cull;
    objects = (ObjectData*)objs;
    drawCommandsIn = (DrawCommand*)draws_in;
    sh_cull::visibility = visibility;
    std::vector<DrawCommand> scratch;
    if (!draws_out) { scratch.resize(pc->object_count + 1); draws_out = (AshIndirectDrawCommand*)scratch.data(); }
    drawCommandsOut = (DrawCommand*)draws_out;
    visibleCount = 0u;                                          // PIN-8: zeroed by the host before the dispatch
    memcpy(&sh_cull::pc, pc, sizeof(*pc));
    sampler2D s = make_sampler(depth, W, H, FMT_R32F, 0);       // PIN-3: one mipmapped R32F image, mip 0 = depth
    int m = W > H ? W : H, levels = 1;
    while (m > 1) { m >>= 1; ++levels; }
    size_t off = 0;
    for (int l = 1; l < levels; ++l) {
        s.w[l] = s.w[l - 1] / 2 > 0 ? s.w[l - 1] / 2 : 1;
        s.h[l] = s.h[l - 1] / 2 > 0 ? s.h[l - 1] / 2 : 1;
        s.level[l] = mips + off;
        off += (size_t)s.w[l] * s.h[l];
    }
    s.levels = levels;
    hizPyramid = s;
    const uint32_t n = div_up((int)pc->object_count, 64) * 64;
    for (uint32_t i = 0; i < n; ++i) {
        gl_GlobalInvocationID = uvec3(i, 0u, 0u);
        shader_main();
    }
    return visibleCount;
}

// shaders/light_culling.comp; 16x16 work groups (the specialisation-constant defaults), one per tile
void ash_ref_light_cull(const AshGpuLight* lights_in, const AshLightCullingPushConstants* pc,
                        const AshCullingCameraData* cam, const float* depth,
                        uint32_t* tile_data, uint32_t* true_counts) {
    using namespace sh_light;
    const int W = (int)pc->screen_size[0], H = (int)pc->screen_size[1];
    lights = (Light*)lights_in;
    depthBuffer = make_sampler(depth, W, H, FMT_R32F, 0);
    tileData = tile_data;
    memcpy(&sh_light::pc, pc, sizeof(*pc));
    memcpy(&camera, cam, sizeof(*cam));
    const int tx = div_up(W, 16), ty = div_up(H, 16);
    std::vector<char> stacks((size_t)kMaxFibres * kStack);
    barrier_hook = fibre_yield;
    for (int gy = 0; gy < ty; ++gy)
        for (int gx = 0; gx < tx; ++gx) {
            gl_WorkGroupID = uvec3((uint)gx, (uint)gy, 0u);
            run_group(256, stacks, shader_main, [&](int i) {
                gl_LocalInvocationID = uvec3((uint)(i % 16), (uint)(i / 16), 0u);
                gl_GlobalInvocationID = uvec3((uint)(gx * 16 + i % 16), (uint)(gy * 16 + i / 16), 0u);
            });
            if (true_counts) true_counts[gy * tx + gx] = pc->light_count ? sharedLightCount : 0u;
        }
    barrier_hook = nullptr;
}

// shaders/taa_resolve.comp; local size 8x8; velocity == NULL binds a 1x1 zero texture (PIN-12)
void ash_ref_taa_resolve_fmt(const uint16_t* current, const uint16_t* history, const uint16_t* velocity,
                             const AshTaaPushConstants* pc, uint16_t* output, int fmt);
void ash_ref_taa_resolve(const uint16_t* current, const uint16_t* history, const uint16_t* velocity,
                         const AshTaaPushConstants* pc, uint16_t* output) {
    ash_ref_taa_resolve_fmt(current, history, velocity, pc, output, 0);
}
// fmt bits: 1 = `current` is RGBA32F, 2 = `history` is, 4 = `outputColor` is an rgba32f image
void ash_ref_taa_resolve_fmt(const uint16_t* current, const uint16_t* history, const uint16_t* velocity,
                             const AshTaaPushConstants* pc, uint16_t* output, int fmt) {
    using namespace sh_taa;
    const int W = (int)pc->screen_params[0], H = (int)pc->screen_params[1];
    currentColor = make_sampler(current, W, H, (fmt & 1) ? FMT_RGBA32F : FMT_RGBA16F, 1);
    historyColor = make_sampler(history, W, H, (fmt & 2) ? FMT_RGBA32F : FMT_RGBA16F, 1);
    velocityBuffer = velocity ? make_sampler(velocity, W, H, FMT_RG16F, 1) : make_sampler(kZeroTexel, 1, 1, FMT_RG16F, 1);
    depthBuffer = make_sampler(kZeroTexel, 1, 1, FMT_RG16F, 0);      // binding 3 is never read
    outputColor = image2D{output, W, H, (fmt & 4) ? FMT_RGBA32F : FMT_RGBA16F};
    memcpy(&sh_taa::pc, pc, sizeof(*pc));
    const int gy = div_up(H, 8) * 8, gx = div_up(W, 8) * 8;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < gy; ++y)
        for (int x = 0; x < gx; ++x) {
            gl_GlobalInvocationID = uvec3((uint)x, (uint)y, 0u);
            shader_main();
        }
}

// shaders/bloom_threshold.frag (push constants: fullscreen_pass.rs PostProcessPushConstants, .threshold in slot 3)
void ash_ref_bloom_threshold_fmt(const uint16_t* hdr, int W, int H, float threshold, uint16_t* dst, int fmt);
void ash_ref_bloom_threshold(const uint16_t* hdr, int W, int H, float threshold, uint16_t* dst) {
    ash_ref_bloom_threshold_fmt(hdr, W, H, threshold, dst, 0);
}
void ash_ref_bloom_threshold_fmt(const uint16_t* hdr, int W, int H, float threshold, uint16_t* dst, int fmt) {   // 1 = src f32, 4 = dst f32
    using namespace sh_threshold;
    hdrBuffer = make_sampler(hdr, W, H, (fmt & 1) ? FMT_RGBA32F : FMT_RGBA16F, 1);
    sh_threshold::pc.exposure = 1.0f; sh_threshold::pc.gamma = 2.2f; sh_threshold::pc.bloomIntensity = 0.5f;
    sh_threshold::pc.threshold = threshold;
    fullscreen_rgba16f(W, H, dst, false, [](vec2 uv) { fragTexCoord = uv; }, shader_main, [] { return outColor; }, (fmt >> 2) & 1);
}

// shaders/bloom_downsample.frag; texelSize = 1 / size(source) (PIN-9)
void ash_ref_bloom_downsample_fmt(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh, int f32);
void ash_ref_bloom_downsample(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh) {
    ash_ref_bloom_downsample_fmt(src, sw, sh, dst, dw, dh, 0);
}
void ash_ref_bloom_downsample_fmt(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh, int f32) {
    using namespace sh_down;
    sourceTexture = make_sampler(src, sw, sh, f32 ? FMT_RGBA32F : FMT_RGBA16F, 1);
    sh_down::pc.texelSize = vec2(1.0f / (float)sw, 1.0f / (float)sh);
    fullscreen_rgba16f(dw, dh, dst, false, [](vec2 uv) { fragTexCoord = uv; }, shader_main, [] { return outColor; }, f32);
}

// shaders/bloom_upsample.frag, additively blended onto dst (PIN-9)
void ash_ref_bloom_upsample_add_fmt(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh, float intensity, int f32);
void ash_ref_bloom_upsample_add(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh, float intensity) {
    ash_ref_bloom_upsample_add_fmt(src, sw, sh, dst, dw, dh, intensity, 0);
}
void ash_ref_bloom_upsample_add_fmt(const uint16_t* src, int sw, int sh, uint16_t* dst, int dw, int dh, float intensity, int f32) {
    using namespace sh_up;
    sourceTexture = make_sampler(src, sw, sh, f32 ? FMT_RGBA32F : FMT_RGBA16F, 1);
    sh_up::pc.texelSize = vec2(1.0f / (float)sw, 1.0f / (float)sh);
    sh_up::pc.intensity = intensity;
    fullscreen_rgba16f(dw, dh, dst, true, [](vec2 uv) { fragTexCoord = uv; }, shader_main, [] { return outColor; }, f32);
}

// shaders/tonemapping.frag into an R8G8B8A8_UNORM target (PIN-11); bloom == NULL binds a 1x1 zero texture
void ash_ref_tonemap_fmt(const uint16_t* hdr, const uint16_t* bloom, int W, int H,
                         const AshPostProcessPushConstants* pc, uint8_t* ldr8, int fmt);
void ash_ref_tonemap(const uint16_t* hdr, const uint16_t* bloom, int W, int H,
                     const AshPostProcessPushConstants* pc, uint8_t* ldr8) {
    ash_ref_tonemap_fmt(hdr, bloom, W, H, pc, ldr8, 0);
}
void ash_ref_tonemap_fmt(const uint16_t* hdr, const uint16_t* bloom, int W, int H,
                         const AshPostProcessPushConstants* pc, uint8_t* ldr8, int fmt) {   // 1 = hdr f32, 2 = bloom f32
    using namespace sh_tonemap;
    hdrBuffer = make_sampler(hdr, W, H, (fmt & 1) ? FMT_RGBA32F : FMT_RGBA16F, 1);
    bloomBuffer = bloom ? make_sampler(bloom, W, H, (fmt & 2) ? FMT_RGBA32F : FMT_RGBA16F, 1) : make_sampler(kZeroTexel, 1, 1, FMT_RGBA16F, 1);
    memcpy(&sh_tonemap::pc, pc, sizeof(*pc));
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            fragTexCoord = vec2(((float)x + 0.5f) / (float)W, ((float)y + 0.5f) / (float)H);
            shader_main();
            const float c[4] = {outColor.x, outColor.y, outColor.z, outColor.w};
            uint8_t* o = ldr8 + ((size_t)y * W + x) * 4;
            for (int k = 0; k < 4; ++k) o[k] = (uint8_t)floorf(clamp(c[k], 0.0f, 1.0f) * 255.0f + 0.5f);
        }
}

uint16_t ash_ref_f2h(float f) { return float_to_half(f); }
float ash_ref_h2f(uint16_t h) { return half_to_float(h); }

}  // extern "C"
