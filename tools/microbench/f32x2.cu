// Microbenchmark: is the packed fp32 pipe (FFMA2 / FADD2 / FMUL2, sm_100a) worth using for issue-bound kernels?
// Each variant does the same number of fp32 FMAs per thread; prints GFMA/s.  Build: nvcc -arch=sm_100a -O3 f32x2.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void fma2(float& x0, float& x1, float a0, float a1, float b0, float b1) {
    asm volatile("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%0,%1}; mov.b64 rb, {%2,%3}; mov.b64 rc, {%4,%5};"
                 " fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0,%1}, ra; }"
                 : "+f"(x0), "+f"(x1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}

template <int MODE>   // 0: scalar FFMA, 1: FFMA2, 2: scalar FFMA + 1 integer op per FMA, 3: FFMA2 + 1 integer op per 2 FMA... (same int work as 2)
__global__ void __launch_bounds__(256) k(float* out, int iters, float a, float b) {
    float x[16];
    unsigned u = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = __fmaf_rn(x[i], a, b);
        } else {
#pragma unroll
            for (int i = 0; i < 16; i += 2) fma2(x[i], x[i + 1], a, a, b, b);
        }
        if (MODE >= 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) u = (u ^ (u >> 3)) + 0x9E3779B9u * i;   // ~3 ALU ops each
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)u;
}

template <int MODE>
void run(const char* name, float* out) {
    const int iters = 4096, blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 16, 0.999f, 0.001f);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, iters, 0.999f, 0.001f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = (double)blocks * 256 * iters * 16;
    printf("%-28s %8.3f ms  %8.1f GFMA/s\n", name, ms, fmas / ms * 1e-6);
}

int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("scalar FFMA", out);
    run<1>("packed FFMA2", out);
    run<2>("scalar FFMA + ALU", out);
    run<3>("packed FFMA2 + ALU", out);
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
