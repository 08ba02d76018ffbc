// Multi-tensor Adam step for the Gaussian parameter groups (and any other fp32 parameter list).
//
// Replaces torch.optim.Adam(l, lr=0.0, eps=1e-15).step() as set up by MultiGaussianMesh.training_setup
// (scene_reconstruction/gaussian_mesh.py:121-135) and called at scene_reconstruction/train_utils.py:312-317:
// per parameter tensor torch issues lerp_, mul_, addcmul_, sqrt, div, add_, addcdiv_ (7 groups -> ~50 launches,
// 262 meshnet tensors -> ~1800).  Here one launch covers up to 48 tensors; a block owns 1024 consecutive elements of
// one tensor.  HBM-bound: 16 B read (p, g, m, v) + 12 B written (+4 when the gradient is zeroed) per element.
// Arithmetic follows torch/optim/adam.py:_single_tensor_adam operation by operation (fp32, scalars derived in
// double on the host and rounded once).
#include "common.cuh"

namespace csb {

constexpr int ADAM_MAX_TENSORS = 48;
constexpr int ADAM_BLOCK_ELEMS = 1024;

struct AdamTable {
    float* p[ADAM_MAX_TENSORS];
    float* g[ADAM_MAX_TENSORS];
    float* m[ADAM_MAX_TENSORS];
    float* v[ADAM_MAX_TENSORS];
    long long numel[ADAM_MAX_TENSORS];
    unsigned block_end[ADAM_MAX_TENSORS];  // exclusive prefix end of the blocks of tensor i
    float neg_step_size[ADAM_MAX_TENSORS]; // -(lr / (1 - beta1^t))
    double lr[ADAM_MAX_TENSORS];           // capturable mode: the step size is formed on the device
    unsigned char vec[ADAM_MAX_TENSORS];   // all four pointers 16-byte aligned
    int n;
};

struct AdamScalars {
    float w1;        // 1 - beta1
    float beta2;
    float w2;        // 1 - beta2
    float bc2_sqrt;  // sqrt(1 - beta2^t)
    float eps;
    int zero_grads;
    // capturable mode (CUDA graphs): the step count lives on the device; adam_advance_kernel derives the two bias
    // corrections from it in double, like torch does on the host, and the step kernel reads them from here
    const double* dev_bias;  // [0] = 1 - beta1^t, [1] = sqrt(1 - beta2^t); nullptr: host-side scalars above
};

__global__ void adam_advance_kernel(long long* step, double beta1, double beta2, double* bias) {
    const long long t = *step + 1;
    *step = t;
    bias[0] = 1.0 - pow(beta1, (double)t);
    bias[1] = sqrt(1.0 - pow(beta2, (double)t));
}

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float nss, const AdamScalars& s) {
    m = __fadd_rn(m, __fmul_rn(s.w1, __fsub_rn(g, m)));                    // exp_avg.lerp_(grad, 1 - beta1)
    v = __fadd_rn(__fmul_rn(v, s.beta2), __fmul_rn(__fmul_rn(s.w2, g), g)); // mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc2_sqrt), s.eps);
    p = __fadd_rn(p, __fmul_rn(nss, __fdiv_rn(m, denom)));                  // addcdiv_(exp_avg, denom, -step_size)
}

__global__ void __launch_bounds__(256)
adam_step_kernel(const __grid_constant__ AdamTable t, const AdamScalars s_in) {
    AdamScalars s = s_in;
    // which tensor owns this block
    int lo = 0, hi = t.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (blockIdx.x < t.block_end[mid]) hi = mid; else lo = mid + 1;
    }
    const int ti = lo;
    const unsigned first = ti ? t.block_end[ti - 1] : 0u;
    const long long base = (long long)(blockIdx.x - first) * ADAM_BLOCK_ELEMS;
    const long long n = t.numel[ti];
    float* __restrict__ P = t.p[ti];
    float* __restrict__ G = t.g[ti];
    float* __restrict__ M = t.m[ti];
    float* __restrict__ V = t.v[ti];
    float nss = t.neg_step_size[ti];
    if (s.dev_bias) {
        nss = (float)(-(t.lr[ti] / s.dev_bias[0]));
        s.bc2_sqrt = (float)s.dev_bias[1];
    }
    if (t.vec[ti] && base + ADAM_BLOCK_ELEMS <= n) {
#pragma unroll
        for (int k = 0; k < ADAM_BLOCK_ELEMS / (256 * 4); ++k) {
            const long long i = base + (long long)(k * 256 + threadIdx.x) * 4;
            float4 p = *reinterpret_cast<float4*>(P + i);
            const float4 g = *reinterpret_cast<const float4*>(G + i);
            float4 m = *reinterpret_cast<float4*>(M + i);
            float4 v = *reinterpret_cast<float4*>(V + i);
            adam_update(p.x, g.x, m.x, v.x, nss, s);
            adam_update(p.y, g.y, m.y, v.y, nss, s);
            adam_update(p.z, g.z, m.z, v.z, nss, s);
            adam_update(p.w, g.w, m.w, v.w, nss, s);
            *reinterpret_cast<float4*>(P + i) = p;
            *reinterpret_cast<float4*>(M + i) = m;
            *reinterpret_cast<float4*>(V + i) = v;
            if (s.zero_grads) *reinterpret_cast<float4*>(G + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    } else {
        for (long long i = base + threadIdx.x; i < n && i < base + ADAM_BLOCK_ELEMS; i += 256) {
            float p = P[i], m = M[i], v = V[i];
            adam_update(p, G[i], m, v, nss, s);
            P[i] = p;
            M[i] = m;
            V[i] = v;
            if (s.zero_grads) G[i] = 0.f;
        }
    }
}

}  // namespace csb

using namespace csb;

static int adam_step_impl(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                          float* const* exp_avg_sq, const int64_t* numel, const double* lr, double beta1,
                          double beta2, double eps, int step, long long* step_dev, double* bias_dev, int zero_grads,
                          void* stream_) {
    if (n_tensors < 0 || (step < 1 && !step_dev)) return CSB_ERR_BAD_ARG;
    if (n_tensors == 0) return 0;
    if (!params || !grads || !exp_avg || !exp_avg_sq || !numel || !lr) return CSB_ERR_BAD_ARG;
    if (step_dev && !bias_dev) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (step_dev) {
        step = 1;  // placeholders below; the kernel takes the device values
        CSB_LAUNCH("adam_advance", stream, adam_advance_kernel<<<1, 1, 0, stream>>>(step_dev, beta1, beta2, bias_dev));
    }
    // torch/optim/adam.py: bias_correction1 = 1 - beta1 ** step; step_size = lr / bias_correction1;
    // bias_correction2_sqrt = sqrt(1 - beta2 ** step)   (Python floats = doubles)
    const double bc1 = 1.0 - __builtin_pow(beta1, (double)step);
    const double bc2_sqrt = __builtin_sqrt(1.0 - __builtin_pow(beta2, (double)step));
    AdamScalars s;
    s.w1 = (float)(1.0 - beta1);
    s.beta2 = (float)beta2;
    s.w2 = (float)(1.0 - beta2);
    s.bc2_sqrt = (float)bc2_sqrt;
    s.eps = (float)eps;
    s.zero_grads = zero_grads;
    s.dev_bias = step_dev ? bias_dev : nullptr;

    AdamTable t;
    t.n = 0;
    unsigned blocks = 0;
    auto flush = [&]() -> int {
        if (t.n == 0) return 0;
        CSB_LAUNCH("adam_step", stream, adam_step_kernel<<<blocks, 256, 0, stream>>>(t, s));
        t.n = 0;
        blocks = 0;
        return 0;
    };
    for (int i = 0; i < n_tensors; ++i) {
        if (numel[i] < 0) return CSB_ERR_BAD_ARG;
        if (numel[i] == 0 || !grads[i]) continue;
        if (!params[i] || !exp_avg[i] || !exp_avg_sq[i]) return CSB_ERR_BAD_ARG;
        const long long nb = (numel[i] + ADAM_BLOCK_ELEMS - 1) / ADAM_BLOCK_ELEMS;
        if (nb > 0x7fffffffLL) return CSB_ERR_UNSUPPORTED;
        if (t.n == ADAM_MAX_TENSORS || (unsigned long long)blocks + nb > 0x7fffffffULL) {
            const int rc = flush();
            if (rc) return rc;
        }
        const int k = t.n++;
        t.p[k] = params[i];
        t.g[k] = grads[i];
        t.m[k] = exp_avg[i];
        t.v[k] = exp_avg_sq[i];
        t.numel[k] = numel[i];
        blocks += (unsigned)nb;
        t.block_end[k] = blocks;
        t.neg_step_size[k] = (float)(-(lr[i] / bc1));
        t.lr[k] = lr[i];
        t.vec[k] = ((((uintptr_t)params[i]) | ((uintptr_t)grads[i]) | ((uintptr_t)exp_avg[i]) |
                     ((uintptr_t)exp_avg_sq[i])) & 15u) == 0;
    }
    return flush();
}

extern "C" int csb_adam_step(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                             float* const* exp_avg_sq, const int64_t* numel, const double* lr, double beta1,
                             double beta2, double eps, int step, int zero_grads, void* stream_) {
    return adam_step_impl(n_tensors, params, grads, exp_avg, exp_avg_sq, numel, lr, beta1, beta2, eps, step, nullptr,
                          nullptr, zero_grads, stream_);
}

extern "C" int csb_adam_step_capturable(int n_tensors, float* const* params, float* const* grads,
                                        float* const* exp_avg, float* const* exp_avg_sq, const int64_t* numel,
                                        const double* lr, double beta1, double beta2, double eps, long long* step_dev,
                                        double* bias_dev, int zero_grads, void* stream_) {
    if (!step_dev || !bias_dev) return CSB_ERR_BAD_ARG;
    return adam_step_impl(n_tensors, params, grads, exp_avg, exp_avg_sq, numel, lr, beta1, beta2, eps, 0, step_dev,
                          bias_dev, zero_grads, stream_);
}
