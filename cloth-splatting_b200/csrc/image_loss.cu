// Photometric loss of the refinement step, forward and backward: L1 + lambda * (1 - SSIM).
//
// Replaces utils/loss_utils.py:20-23 (l1_loss), :28-70 (ssim / _ssim: five depthwise 11x11 conv2d with zero
// padding, ~15 elementwise kernels, the window rebuilt and uploaded on every call) and their autograd graph, as
// called by scene_reconstruction/train_utils.py:50-74 (image_losses).
//
//   mu1 = w*x, mu2 = w*y, s1 = w*(x x) - mu1^2, s2 = w*(y y) - mu2^2, s12 = w*(x y) - mu1 mu2
//   S   = (2 mu1 mu2 + C1)(2 s12 + C2) / ((mu1^2 + mu2^2 + C1)(s1 + s2 + C2))
//
// The 11x11 window is the outer product of an 11-tap Gaussian (sigma 1.5), so every convolution is run as a
// horizontal then a vertical 11-tap pass over a shared-memory tile.  Two kernels:
//   image_loss_forward_kernel   tile of 32x32 pixels of one image plane; emits per pixel S (optional) and the three
//       partial derivatives  A = dS/dmu1 (total, through s1 and s12 too), B = dS/ds1, C = dS/ds12,  and per block the
//       sums of |x - y| m and (1 - S) m  (m = mask, broadcast over channels).
//   image_loss_backward_kernel  dL/dx(p) = sum_q w(q-p) g(q) [A(q) + 2 x(p) B(q) + y(p) C(q)] + l1w * m(p) sign(x - y)
//       with g(q) = dL/dS(q): again one separable 11-tap convolution, of the three maps g A, g B, g C.
// Block sums are written to a partials array and folded in a fixed order by image_loss_finalize_kernel (fp64), so the
// loss value is deterministic.  Algorithmic HBM traffic: forward 8 B read + 12 B written, backward 20 B read + 4 B
// written per pixel and channel (44 B; + the mask).
#include "common.cuh"

namespace csb {

constexpr int LT = 32;             // tile edge (outputs)
constexpr int LR = 5;              // window radius (11 taps)
constexpr int LH = LT + 2 * LR;    // 42: tile + halo
constexpr int LP = 43;             // pitch of the halo tiles (conflict-free for the 4-wide horizontal groups)
constexpr int LG = 4;              // outputs per thread along the pass direction

struct LossWindow {
    float w[11];
};

struct LossShape {
    int planes;      // n_img * C
    int C, H, W;
    int mask_c;      // 0 = no mask, 1 = one mask plane per image, C = one per image plane
    int tiles_x, tiles_y;
};

__device__ __forceinline__ float mask_at(const float* __restrict__ mask, const LossShape& s, int plane, int y, int x) {
    if (!mask) return 1.f;
    const int img = plane / s.C, c = plane - img * s.C;
    const size_t mp = s.mask_c == 1 ? (size_t)img : (size_t)img * s.C + c;
    return mask[(mp * s.H + y) * s.W + x];
}

// grid: x = tiles_x * tiles_y, y = planes.  256 threads.
__global__ void __launch_bounds__(256)
image_loss_forward_kernel(LossShape s, LossWindow win, const float* __restrict__ img, const float* __restrict__ gt,
                          const float* __restrict__ mask, float* __restrict__ ssim_map, float* __restrict__ dmaps,
                          float* __restrict__ partials) {
    __shared__ float sx[LH * LP], sy[LH * LP];
    __shared__ __align__(16) float hs[5][LH][LT];
    __shared__ float red[2][8];

    const int plane = blockIdx.y;
    const int ty = blockIdx.x / s.tiles_x, tx = blockIdx.x - ty * s.tiles_x;
    const int x0 = tx * LT - LR, y0 = ty * LT - LR;
    const size_t pbase = (size_t)plane * s.H * s.W;
    const int tid = threadIdx.x;

    // The moments are taken about the tile's centre values (cx, cy): variances and the covariance do not depend on
    // the origin, and E[x^2] - mu^2 about a nearby origin loses |x - cx|^2 / x^2 fewer bits to cancellation than
    // about 0 (exact zeros in flat regions, where the fp32 reference carries ~1e-7 / C2 = 1e-4 of noise).  Zero
    // padding is kept: a padded sample is the value 0, i.e. -cx after the shift.
    const int ccy = min(ty * LT + LT / 2, s.H - 1), ccx = min(tx * LT + LT / 2, s.W - 1);
    const float cx = img[pbase + (size_t)ccy * s.W + ccx], cy = gt[pbase + (size_t)ccy * s.W + ccx];
    for (int i = tid; i < LH * LH; i += 256) {
        const int r = i / LH, c = i - r * LH;
        const int gy = y0 + r, gx = x0 + c;
        float a = 0.f, b = 0.f;
        if (gy >= 0 && gy < s.H && gx >= 0 && gx < s.W) {
            a = img[pbase + (size_t)gy * s.W + gx];
            b = gt[pbase + (size_t)gy * s.W + gx];
        }
        sx[r * LP + c] = a - cx;
        sy[r * LP + c] = b - cy;
    }
    __syncthreads();

    // horizontal pass: item = (row, group of 4 columns)
    for (int i = tid; i < LH * (LT / LG); i += 256) {
        const int r = i >> 3, cg = (i & 7) * LG;
        float vx[LG + 10], vy[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) {
            vx[k] = sx[r * LP + cg + k];
            vy[k] = sy[r * LP + cg + k];
        }
        float acc[5][LG];
#pragma unroll
        for (int o = 0; o < LG; ++o) {
#pragma unroll
            for (int q = 0; q < 5; ++q) acc[q][o] = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                const float a = vx[o + k], b = vy[o + k], w = win.w[k];
                acc[0][o] = fmaf(w, a, acc[0][o]);
                acc[1][o] = fmaf(w, b, acc[1][o]);
                acc[2][o] = fmaf(w, a * a, acc[2][o]);
                acc[3][o] = fmaf(w, b * b, acc[3][o]);
                acc[4][o] = fmaf(w, a * b, acc[4][o]);
            }
        }
#pragma unroll
        for (int q = 0; q < 5; ++q)
            *reinterpret_cast<float4*>(&hs[q][r][cg]) = make_float4(acc[q][0], acc[q][1], acc[q][2], acc[q][3]);
    }
    __syncthreads();

    // vertical pass: thread = (column, group of 4 rows)
    const int col = tid & 31, rg = (tid >> 5) * LG;
    float out[5][LG];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        float v[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) v[k] = hs[q][rg + k][col];
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) a = fmaf(win.w[k], v[o + k], a);
            out[q][o] = a;
        }
    }

    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const size_t n = (size_t)s.planes * s.H * s.W;
    float sum_l1 = 0.f, sum_ds = 0.f;
    const int gx = tx * LT + col;
#pragma unroll
    for (int o = 0; o < LG; ++o) {
        const int gy = ty * LT + rg + o;
        if (gx < s.W && gy < s.H) {
            const float a0 = out[0][o], a1 = out[1][o];
            const float mu1 = cx + a0, mu2 = cy + a1;
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = out[2][o] - a0 * a0, s2 = out[3][o] - a1 * a1, s12 = out[4][o] - a0 * a1;
            const float N1 = 2.f * mu12 + C1, N2 = 2.f * s12 + C2;
            const float D1 = mu1_sq + mu2_sq + C1, D2 = s1 + s2 + C2;
            const float S = (N1 * N2) / (D1 * D2);
            const size_t pix = pbase + (size_t)gy * s.W + gx;
            const float m = mask_at(mask, s, plane, gy, gx);
            const float xv = img[pix], yv = gt[pix];  // L1/L2 hits: the tile was just staged from these lines
            sum_l1 += fabsf((xv - yv) * m);
            sum_ds += (1.f - S) * m;
            if (ssim_map) ssim_map[pix] = S;
            if (dmaps) {
                const float inv = 1.f / (D1 * D2);
                const float dS_dmu1 = 2.f * N2 * inv * (mu2 - mu1 * N1 / D1);
                const float dS_ds1 = -S / D2;
                const float dS_ds12 = 2.f * N1 * inv;
                dmaps[pix] = dS_dmu1 - 2.f * mu1 * dS_ds1 - mu2 * dS_ds12;
                dmaps[n + pix] = dS_ds1;
                dmaps[2 * n + pix] = dS_ds12;
            }
        }
    }
    // block sums
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum_l1 += __shfl_xor_sync(0xffffffffu, sum_l1, d);
        sum_ds += __shfl_xor_sync(0xffffffffu, sum_ds, d);
    }
    if ((tid & 31) == 0) {
        red[0][tid >> 5] = sum_l1;
        red[1][tid >> 5] = sum_ds;
    }
    __syncthreads();
    if (tid == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a += red[0][k];
            b += red[1][k];
        }
        const size_t blk = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
        partials[2 * blk] = a;
        partials[2 * blk + 1] = b;
    }
}

// One block folds the per-tile sums in a fixed order: losses = {l1 + lambda * dssim, l1, dssim}.
__global__ void __launch_bounds__(256)
image_loss_finalize_kernel(size_t blocks, double inv_count, float lambda, const float* __restrict__ partials,
                           float* __restrict__ losses) {
    __shared__ double red[2][256];
    double a = 0.0, b = 0.0;
    for (size_t i = threadIdx.x; i < blocks; i += 256) {
        a += (double)partials[2 * i];
        b += (double)partials[2 * i + 1];
    }
    red[0][threadIdx.x] = a;
    red[1][threadIdx.x] = b;
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) {
            red[0][threadIdx.x] += red[0][threadIdx.x + d];
            red[1][threadIdx.x] += red[1][threadIdx.x + d];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float l1 = (float)(red[0][0] * inv_count), ds = (float)(red[1][0] * inv_count);
        losses[0] = l1 + lambda * ds;
        losses[1] = l1;
        losses[2] = ds;
    }
}

// dL/dimg.  g(q) = grad_map ? grad_map[q] : w_ssim * m(q) * scale,  L1 term l1w * m(p) * sign(x - y) * scale,
// scale = *grad_scale (device scalar; NULL = 1).
__global__ void __launch_bounds__(256)
image_loss_backward_kernel(LossShape s, LossWindow win, const float* __restrict__ img, const float* __restrict__ gt,
                           const float* __restrict__ mask, const float* __restrict__ dmaps,
                           const float* __restrict__ grad_map, const float* __restrict__ grad_scale, float w_ssim,
                           float w_l1, float* __restrict__ dimg) {
    __shared__ float sa[3][LH * LP];
    __shared__ __align__(16) float hs[3][LH][LT];

    const int plane = blockIdx.y;
    const int ty = blockIdx.x / s.tiles_x, tx = blockIdx.x - ty * s.tiles_x;
    const int x0 = tx * LT - LR, y0 = ty * LT - LR;
    const size_t pbase = (size_t)plane * s.H * s.W;
    const size_t n = (size_t)s.planes * s.H * s.W;
    const int tid = threadIdx.x;
    const float scale = grad_scale ? *grad_scale : 1.f;
    const float ws = w_ssim * scale;

    for (int i = tid; i < LH * LH; i += 256) {
        const int r = i / LH, c = i - r * LH;
        const int gy = y0 + r, gx = x0 + c;
        float a = 0.f, b = 0.f, cc = 0.f;
        if (gy >= 0 && gy < s.H && gx >= 0 && gx < s.W) {
            const size_t pix = pbase + (size_t)gy * s.W + gx;
            const float g = grad_map ? grad_map[pix] : ws * mask_at(mask, s, plane, gy, gx);
            a = g * dmaps[pix];
            b = g * dmaps[n + pix];
            cc = g * dmaps[2 * n + pix];
        }
        sa[0][r * LP + c] = a;
        sa[1][r * LP + c] = b;
        sa[2][r * LP + c] = cc;
    }
    __syncthreads();

    for (int i = tid; i < LH * (LT / LG); i += 256) {
        const int r = i >> 3, cg = (i & 7) * LG;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            float v[LG + 10];
#pragma unroll
            for (int k = 0; k < LG + 10; ++k) v[k] = sa[q][r * LP + cg + k];
            float acc[LG];
#pragma unroll
            for (int o = 0; o < LG; ++o) {
                acc[o] = 0.f;
#pragma unroll
                for (int k = 0; k < 11; ++k) acc[o] = fmaf(win.w[k], v[o + k], acc[o]);
            }
            *reinterpret_cast<float4*>(&hs[q][r][cg]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        }
    }
    __syncthreads();

    const int col = tid & 31, rg = (tid >> 5) * LG;
    float out[3][LG];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        float v[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) v[k] = hs[q][rg + k][col];
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) a = fmaf(win.w[k], v[o + k], a);
            out[q][o] = a;
        }
    }
    const int gx = tx * LT + col;
    const float l1w = w_l1 * scale;
#pragma unroll
    for (int o = 0; o < LG; ++o) {
        const int gy = ty * LT + rg + o;
        if (gx < s.W && gy < s.H) {
            const size_t pix = pbase + (size_t)gy * s.W + gx;
            const float xv = img[pix], yv = gt[pix];
            float g = out[0][o] + 2.f * xv * out[1][o] + yv * out[2][o];
            if (l1w != 0.f) {
                const float m = mask_at(mask, s, plane, gy, gx);
                const float d = (xv - yv) * m;
                g += l1w * m * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
            }
            dimg[pix] = g;
        }
    }
}

static LossWindow make_window() {
    // utils/loss_utils.py:28-30: torch.Tensor([exp(-(x - 5)**2 / (2 sigma^2))]) (python doubles rounded to
    // fp32), divided by its fp32 sum
    LossWindow w;
    float g[11];
    float sum = 0.f;
    for (int x = 0; x < 11; ++x) {
        const double d = (double)(x - 5);
        g[x] = (float)__builtin_exp(-(d * d) / (2.0 * 1.5 * 1.5));
    }
    // torch.sum over 11 floats: sequential accumulation is what the vectorised kernel reduces to at this size
    for (int x = 0; x < 11; ++x) sum += g[x];
    for (int x = 0; x < 11; ++x) w.w[x] = g[x] / sum;
    return w;
}

static bool make_shape(LossShape& s, int n_img, int C, int H, int W, int mask_channels, bool has_mask) {
    if (n_img < 0 || C <= 0 || H <= 0 || W <= 0) return false;
    if (has_mask && mask_channels != 1 && mask_channels != C) return false;
    s.planes = n_img * C;
    s.C = C;
    s.H = H;
    s.W = W;
    s.mask_c = has_mask ? mask_channels : 0;
    s.tiles_x = (W + LT - 1) / LT;
    s.tiles_y = (H + LT - 1) / LT;
    return true;
}

}  // namespace csb

using namespace csb;

extern "C" {

size_t csb_image_loss_workspace_bytes(int n_img, int C, int H, int W) {
    if (n_img <= 0 || C <= 0 || H <= 0 || W <= 0) return 256;
    const size_t blocks = (size_t)n_img * C * ((W + LT - 1) / LT) * ((H + LT - 1) / LT);
    return align_up(blocks * 2 * sizeof(float), 256);
}

int csb_image_loss_forward(int n_img, int C, int H, int W, const float* img, const float* gt, const float* mask,
                           int mask_channels, float lambda_dssim, float* ssim_map, float* dmaps, float* losses,
                           void* workspace, size_t workspace_bytes, void* stream_) {
    LossShape s;
    if (!make_shape(s, n_img, C, H, W, mask_channels, mask != nullptr)) return CSB_ERR_BAD_ARG;
    if (n_img == 0) return CSB_ERR_BAD_ARG;  // mean over an empty tensor: the caller's nan, not ours
    if (!img || !gt || !losses || !workspace) return CSB_ERR_BAD_ARG;
    if (s.planes > 65535) return CSB_ERR_UNSUPPORTED;
    if (workspace_bytes < csb_image_loss_workspace_bytes(n_img, C, H, W)) return CSB_ERR_WORKSPACE;
    cudaStream_t stream = (cudaStream_t)stream_;
    const dim3 grid(s.tiles_x * s.tiles_y, s.planes);
    const LossWindow win = make_window();
    CSB_LAUNCH("image_loss_forward", stream, image_loss_forward_kernel<<<grid, 256, 0, stream>>>(
        s, win, img, gt, mask, ssim_map, dmaps, (float*)workspace));
    const size_t blocks = (size_t)grid.x * grid.y;
    const double inv_count = 1.0 / ((double)s.planes * H * W);
    CSB_LAUNCH("image_loss_finalize", stream, image_loss_finalize_kernel<<<1, 256, 0, stream>>>(
        blocks, inv_count, lambda_dssim, (const float*)workspace, losses));
    return 0;
}

int csb_image_loss_backward(int n_img, int C, int H, int W, const float* img, const float* gt, const float* mask,
                            int mask_channels, const float* dmaps, const float* grad_map, const float* grad_scale,
                            float w_ssim, float w_l1, float* dL_dimg, void* stream_) {
    LossShape s;
    if (!make_shape(s, n_img, C, H, W, mask_channels, mask != nullptr)) return CSB_ERR_BAD_ARG;
    if (n_img == 0) return 0;
    if (!img || !gt || !dmaps || !dL_dimg) return CSB_ERR_BAD_ARG;
    if (s.planes > 65535) return CSB_ERR_UNSUPPORTED;
    cudaStream_t stream = (cudaStream_t)stream_;
    const dim3 grid(s.tiles_x * s.tiles_y, s.planes);
    const LossWindow win = make_window();
    CSB_LAUNCH("image_loss_backward", stream, image_loss_backward_kernel<<<grid, 256, 0, stream>>>(
        s, win, img, gt, mask, dmaps, grad_map, grad_scale, w_ssim, w_l1, dL_dimg));
    return 0;
}

}  // extern "C"
