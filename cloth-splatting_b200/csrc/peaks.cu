// Micro-benchmarks of the pipes that bound the compositing kernels, run on the device the caller is on so that bench.py can
// state its FP32 / SFU roofline against a peak measured in the same run ("peaks to be measured on the box
// with a micro-benchmark and stated").  Nominal: 148 SMs x 128 FMA lanes x 1.965 GHz = 37.2 TFMA/s = 74.4 TFLOP/s;
// MUFU 16 lanes / clk / SM = 4.65 T/s.
#include "common.cuh"

namespace csb {

template <int MODE>  // 0: fma.rn.f32, 1: fma.rn.f32x2, 2: ex2.approx
__global__ void __launch_bounds__(256) peak_kernel(float* out, int iters, float w) {
    float a[8];
    unsigned long long p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = threadIdx.x * 1e-3f + i;
        p[i] = (unsigned long long)(threadIdx.x + i) * 0x100000001ull;
    }
    unsigned long long ww;
    asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "f"(w));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(w), "f"(a[(i + 1) & 7]));
                if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p[i]) : "l"(ww), "l"(p[(i + 1) & 7]));
                if (MODE == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i] + (float)(p[i] & 0xff);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static int time_peak(float* scratch, int blocks, int iters, cudaStream_t stream, double* per_s) {
    cudaEvent_t e0, e1;
    CSB_CUDA_TRY(cudaEventCreate(&e0));
    CSB_CUDA_TRY(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0, stream);
        peak_kernel<MODE><<<blocks, 256, 0, stream>>>(scratch, iters, 1.0001f);
        cudaEventRecord(e1, stream);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) return (int)e;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double inst = (double)blocks * 256 * (double)iters * 64;  // thread-instructions
        best = std::max(best, inst / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *per_s = best;
    return 0;
}

}  // namespace csb

extern "C" int csb_measure_fp32_peaks(void* scratch, size_t scratch_bytes, double* ffma_per_s, double* ffma2_per_s,
                                      double* mufu_per_s, void* stream_) {
    using namespace csb;
    int dev = 0, sms = 148;
    CSB_CUDA_TRY(cudaGetDevice(&dev));
    CSB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = sms * 8;
    if (!scratch || scratch_bytes < (size_t)blocks * 256 * sizeof(float)) return CSB_ERR_WORKSPACE;
    cudaStream_t stream = (cudaStream_t)stream_;
    double v = 0.0;
    int rc;
    if ((rc = time_peak<0>((float*)scratch, blocks, 2048, stream, &v))) return rc;
    if (ffma_per_s) *ffma_per_s = v;
    if ((rc = time_peak<1>((float*)scratch, blocks, 2048, stream, &v))) return rc;
    if (ffma2_per_s) *ffma2_per_s = 2.0 * v;  // two FMAs per instruction
    if ((rc = time_peak<2>((float*)scratch, blocks, 512, stream, &v))) return rc;
    if (mufu_per_s) *mufu_per_s = v;
    return 0;
}
