// Micro-benchmark: FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
template <int MODE>
__global__ void k(float* out, int iters, float w) {
    float a[8];
    unsigned long long p[8];
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x + i; p[i] = (unsigned long long)(threadIdx.x + i) * 0x100000001ull; }
    unsigned long long ww;
    asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "f"(w));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (MODE == 0) {
#pragma unroll
                for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(w), "f"(a[(i + 1) & 7]));
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(ww, p[(i + 1) & 7], p[i]);
            }
        }
    }
    float s = 0;
    for (int i = 0; i < 8; ++i) s += a[i] + (float)(p[i] & 0xff);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float* out;
    cudaMalloc(&out, 148 * 8 * 1024 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096;
    for (int mode = 0; mode < 2; ++mode) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(out, iters, 1.0001f); else k<1><<<148 * 8, 256>>>(out, iters, 1.0001f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double inst = 148.0 * 8 * 256 * (double)iters * 64;  // thread-instructions
            printf("%s: %.3f ms, %.2f T thread-inst/s, %.2f TFMA/s\n", mode ? "FFMA2" : "FFMA ", ms, inst / ms / 1e9, inst * (mode ? 2 : 1) / ms / 1e9);
        }
    }
    return 0;
}
