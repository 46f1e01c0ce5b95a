// Microbenchmark: what does a packed fp32 instruction (FFMA2) or an fp16->fp32 conversion (HADD2.F32) cost the
// warp scheduler on sm_100a?  Each mode runs the same unrolled body per warp, 8 warps per SM sub-partition, and
// prints SMSP cycles per body (clock read from the device).  Bodies (per iteration, all chains independent):
//   A 16 FFMA            B 16 FFMA2           C 16 FFMA2 + 16 LOP3     D 16 FFMA2 + 32 LOP3    E 16 FFMA + 16 LOP3
//   F 48 LOP3            G 16 HADD2.F32       H 16 HADD2.F32 + 16 LOP3  I 16 FFMA2 + 16 HADD2.F32
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 issue_mix.cu -o issue_mix
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

#define FFMA2(x) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x) : "l"(a2), "l"(b2))
#define FFMA(x)  asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(a), "f"(b))
#define LOP(u)   asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(m0), "r"(m1))
#define CVT(f, h) asm volatile("cvt.f32.f16 %0, %1;" : "=f"(f) : "h"(h))

template <int MODE>
__global__ void __launch_bounds__(1024) k(float* out, long long* cyc, int iters, float a, float b, unsigned m0, unsigned m1, unsigned short hseed) {
    unsigned long long x[16], a2, b2;
    float s[16], c[16];
    unsigned u[32];
    unsigned short h[16];
    asm("mov.b64 %0, {%1,%1};" : "=l"(a2) : "f"(a));
    asm("mov.b64 %0, {%1,%1};" : "=l"(b2) : "f"(b));
#pragma unroll
    for (int i = 0; i < 16; ++i) { float v = (float)(threadIdx.x + i); asm("mov.b64 %0, {%1,%1};" : "=l"(x[i]) : "f"(v)); s[i] = v; c[i] = 0.f; h[i] = hseed + i; }
#pragma unroll
    for (int i = 0; i < 32; ++i) u[i] = threadIdx.x * 31 + i;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (MODE == 0 || MODE == 4) FFMA(s[i]);
            if (MODE == 1 || MODE == 2 || MODE == 3 || MODE == 8) FFMA2(x[i]);
            if (MODE == 2 || MODE == 3 || MODE == 4 || MODE == 5 || MODE == 7) LOP(u[i]);
            if (MODE == 3 || MODE == 5) LOP(u[16 + i]);
            if (MODE == 5) LOP(u[(i + 7) & 31]);
            if (MODE == 6 || MODE == 7 || MODE == 8) { CVT(c[i], h[i]); }
        }
        if (MODE >= 6) {
#pragma unroll
            for (int i = 0; i < 16; ++i) h[i] ^= (unsigned short)(__float_as_uint(c[i]) >> 13);   // keeps the conversions live and dependent (1 SHF + 1 LOP each)
        }
    }
    const long long t1 = clock64();
    float r = 0; unsigned w = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[i])); r += lo + hi + s[i] + c[i]; w += h[i]; }
#pragma unroll
    for (int i = 0; i < 32; ++i) w += u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + (float)w;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, float* out, long long* cyc) {
    const int iters = 2048, blocks = 148;
    k<MODE><<<blocks, 1024>>>(out, cyc, 8, 0.999f, 0.001f, 0x1234567u, 0x89abcdeu, 0x3c00);
    k<MODE><<<blocks, 1024>>>(out, cyc, iters, 0.999f, 0.001f, 0x1234567u, 0x89abcdeu, 0x3c00);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)h[i]; avg /= 148;
    // 1024 threads = 32 warps = 8 warps per SMSP; one body per warp per iteration
    printf("%-34s %8.1f SMSP cycles per body (per warp: %.1f)\n", name, avg / iters / 8.0, avg / iters);
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    run<0>("A 16 FFMA", out, cyc);
    run<1>("B 16 FFMA2", out, cyc);
    run<2>("C 16 FFMA2 + 16 LOP3", out, cyc);
    run<3>("D 16 FFMA2 + 32 LOP3", out, cyc);
    run<4>("E 16 FFMA + 16 LOP3", out, cyc);
    run<5>("F 48 LOP3", out, cyc);
    run<6>("G 16 HADD2.F32 (+16 SHF +16 LOP)", out, cyc);
    run<7>("H G + 16 LOP3", out, cyc);
    run<8>("I 16 FFMA2 + G", out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
