// tools/probe/tma_probe.cu — one cp.async.bulk.tensor.2d box (36 x 18 texels of 8 bytes) of a W x H image; the origin may lie outside the image (zero fill) but its innermost coordinate must be 16-byte aligned (x = -1 traps: "illegal instruction")
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
struct __align__(128) Tile { uint2 t[18][36]; };
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int MODE>
__global__ void k(const __grid_constant__ CUtensorMap m, const CUtensorMap* gm, uint2* out, int x, int y) {
    __shared__ Tile T;
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (MODE == 0) {        // mbarrier only
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(&bar)) : "memory");
        } else {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(&bar)), "r"(36 * 18 * 8) : "memory");
            if (MODE == 1)
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             :: "r"(smem_u32(&T)), "l"(&m), "r"(smem_u32(&bar)), "r"(x), "r"(y) : "memory");
            if (MODE == 2)
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             :: "r"(smem_u32(&T)), "l"(gm), "r"(smem_u32(&bar)), "r"(x), "r"(y) : "memory");
            if (MODE == 3)
                asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             :: "r"(smem_u32(&T)), "l"(&m), "r"(smem_u32(&bar)), "r"(x), "r"(y) : "memory");
        }
    }
    asm volatile("{\n\t.reg .pred P1;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra D;\n\tbra W;\n\tD:\n\t}" :: "r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int e = threadIdx.x; e < 18 * 36; e += blockDim.x) out[e] = T.t[e / 36][e % 36];
}
typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    int dt = argc > 1 ? atoi(argv[1]) : 0; int mode = argc > 2 ? atoi(argv[2]) : 1; int only = argc > 3 ? atoi(argv[3]) : -1; int promo = argc > 4 ? atoi(argv[4]) : 2;      // 0: UINT64 box 34 ; 1: UINT32 box 68
    const int W = 200, H = 120;
    std::vector<uint2> img(W * H);
    for (int i = 0; i < W * H; ++i) img[i] = make_uint2(i, ~i);
    uint2 *d, *o; cudaMalloc(&d, W * H * 8); cudaMalloc(&o, 18 * 36 * 8);
    cudaMemcpy(d, img.data(), W * H * 8, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry: %d %d %p\n", (int)e, (int)q, p);
    CUtensorMap m{};
    cuuint64_t dim[2] = {(cuuint64_t)(dt ? 2 * W : W), (cuuint64_t)H}, pitch[1] = {(cuuint64_t)W * 8};
    cuuint32_t box[2] = {(cuuint32_t)(dt ? 72 : 36), 18}, one[2] = {1, 1};
    CUresult r = ((Fn)p)(&m, dt ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, d, dim, pitch, box, one, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    { const unsigned long long* w = reinterpret_cast<const unsigned long long*>(&m); for (int i = 0; i < 16; ++i) printf("%016llx%c", w[i], i % 4 == 3 ? '\n' : ' '); }
    const int xs[3] = {-2, 30, 190}, ys[3] = {-1, 15, 111};
    for (int c = 0; c < 3; ++c) {
        if (only >= 0 && c != only) continue;
        CUtensorMap* gm; cudaMalloc(&gm, sizeof(m)); cudaMemcpy(gm, &m, sizeof(m), cudaMemcpyHostToDevice);
        const int X = dt ? 2 * xs[c] : xs[c];
        if (mode == 0) k<0><<<1, 128>>>(m, gm, o, X, ys[c]);
        if (mode == 1) k<1><<<1, 128>>>(m, gm, o, X, ys[c]);
        if (mode == 2) k<2><<<1, 128>>>(m, gm, o, X, ys[c]);
        if (mode == 3) k<3><<<1, 128>>>(m, gm, o, X, ys[c]);
        e = cudaDeviceSynchronize();
        printf("launch (%d, %d): %s\n", xs[c], ys[c], cudaGetErrorString(e));
        if (e) return 1;
        std::vector<uint2> got(18 * 36);
        cudaMemcpy(got.data(), o, 18 * 36 * 8, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int ly = 0; ly < 18; ++ly) for (int lx = 0; lx < 36; ++lx) {
            int gx = xs[c] + lx, gy = ys[c] + ly;
            uint2 want = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? make_uint2(0, 0) : img[gy * W + gx];
            uint2 g = got[ly * 36 + lx];
            if (g.x != want.x || g.y != want.y) ++bad;
        }
        printf("  mismatches: %d\n", bad);
    }
    return 0;
}
