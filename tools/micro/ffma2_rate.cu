// Micro-benchmark: issue / pipe rate of scalar FFMA against packed FFMA2 (fma.rn.f32x2) on sm_100, and of a 50:50 mix.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu && ./ffma2_rate
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ float fma1(float a, float b, float c) {
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
template <int MODE>   // 0 scalar, 1 packed, 2 mixed (one packed + two scalar per group)
__global__ void k(float* out, int iters, float s) {
    float a[8];
    unsigned long long p[8];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x + i; p[i] = (unsigned long long)(threadIdx.x + i) * 0x3f8000013f800001ull; }
    const unsigned long long sp = 0x3f7fff003f7fff00ull;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) a[i] = fma1(a[i], s, s);
            if (MODE == 1) p[i] = fma2(p[i], sp, sp);
            if (MODE == 2) { if (i & 1) p[i] = fma2(p[i], sp, sp); else { a[i] = fma1(a[i], s, s); a[(i + 1) & 7] = fma1(a[(i + 1) & 7], s, s); } }
        }
    }
    float r = 0;
    for (int i = 0; i < 8; i++) r += a[i] + (float)(p[i] & 0xff);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE>
void run(const char* name, int insts_per_iter, int flops_lane_per_iter) {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(out, 100, 0.999f);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, iters, 0.999f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps = 148.0 * 8 * 8;
    const double winst = warps * iters * insts_per_iter;
    const double clk = 1.965e9;
    printf("%-8s %.3f ms  %.2f warp-instr/clk/SM  %.1f FMA-lanes/clk/SM\n", name, ms, winst / (ms * 1e-3) / clk / 148.0,
           warps * iters * (double)flops_lane_per_iter * 32 / (ms * 1e-3) / clk / 148.0);
    cudaFree(out);
}
int main() {
    run<0>("scalar", 8, 8);
    run<1>("packed", 8, 16);
    run<2>("mixed", 12, 16);
    return 0;
}
