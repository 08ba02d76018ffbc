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
constexpr int LP = 43;             // pitch of the halo tiles in elements (odd: conflict-free 64-bit accesses, see below)
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

// Packed fp32x2 arithmetic of sm_100 (FFMA2 / FMUL2): one issue slot, two lanes.  The separable convolutions below
// are almost pure FMA streams, so pairing (x, y), (x^2, y^2), (gA, gB) halves their instruction count.
__device__ __forceinline__ float2 ffma2(float w, float2 b, float2 c) {
    unsigned long long rw, rb, rc, rd;
    asm("mov.b64 %0, {%1, %1};" : "=l"(rw) : "f"(w));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(rw), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long ra, rb, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

__device__ __forceinline__ const float* mask_plane(const float* __restrict__ mask, const LossShape& s, int plane) {
    if (!mask) return nullptr;
    const int img = plane / s.C;
    const size_t mp = s.mask_c == 1 ? (size_t)img : (size_t)plane;
    return mask + mp * s.H * s.W;
}

// Horizontal-pass work item of lane-linear index i: 4 rows x 4 column groups per half-warp, so that the 16 64-bit
// shared-memory accesses of a half-warp fall into 16 distinct bank pairs (2 * LP * r mod 8 takes 0, 6, 4, 2).
__device__ __forceinline__ void h_item(int i, int& r, int& cg) {
    cg = ((i & 3) + ((i >> 2) & 4)) * LG;
    r = ((i >> 2) & 3) + ((i >> 5) << 2);
}
constexpr int H_ITEMS = ((LH + 3) / 4) * 32;  // 352 slots, rows >= LH idle

// grid: x = tiles_x * tiles_y, y = planes.  256 threads.  MAP / GRAD: ssim_map / dmaps are written.
template <bool MAP, bool GRAD>
__global__ void __launch_bounds__(256, 3)
image_loss_forward_kernel(LossShape s, LossWindow win, const float* __restrict__ img, const float* __restrict__ gt,
                          const float* __restrict__ mask, float* __restrict__ ssim_map, float* __restrict__ dmaps,
                          float* __restrict__ partials) {
    __shared__ __align__(16) float2 sxy[LH * LP];       // (x - cx, y - cy)
    __shared__ __align__(16) float2 h01[LH][LT];        // horizontal sums of (x, y)
    __shared__ __align__(16) float2 h23[LH][LT];        // ... of (x^2, y^2)
    __shared__ __align__(16) float h4[LH][LT];          // ... of x y
    __shared__ float red[2][8];

    const int plane = blockIdx.y;
    const int ty = blockIdx.x / s.tiles_x, tx = blockIdx.x - ty * s.tiles_x;
    const int x0 = tx * LT - LR, y0 = ty * LT - LR;
    const float* __restrict__ pimg = img + (size_t)plane * s.H * s.W;
    const float* __restrict__ pgt = gt + (size_t)plane * s.H * s.W;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // The moments are taken about the tile's centre values (cx, cy): variances and the covariance do not depend on
    // the origin, and E[x^2] - mu^2 about a nearby origin loses |x - cx|^2 / x^2 fewer bits to cancellation than
    // about 0 (exact zeros in flat regions, where the fp32 reference carries ~1e-7 / C2 = 1e-4 of noise).  Zero
    // padding is kept: a padded sample is the value 0, i.e. -cx after the shift.
    const int ccy = min(ty * LT + LT / 2, s.H - 1), ccx = min(tx * LT + LT / 2, s.W - 1);
    const float cx = pimg[(size_t)ccy * s.W + ccx], cy = pgt[(size_t)ccy * s.W + ccx];
    {
        // every warp stages rows warp, warp + 8, ...: all global loads are issued before the first use so that one
        // DRAM latency covers the whole tile instead of one per row
        constexpr int NR = (LH + 7) / 8;
        float va[NR][2], vb[NR][2];
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const int r = warp + 8 * j, gy = y0 + r;
            const bool rok = r < LH && gy >= 0 && gy < s.H;
            const int rowoff = (rok ? gy : 0) * s.W;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = lane + 32 * h, gx = x0 + c;
                const bool ok = rok && c < LH && gx >= 0 && gx < s.W;
                va[j][h] = ok ? pimg[rowoff + gx] : 0.f;
                vb[j][h] = ok ? pgt[rowoff + gx] : 0.f;
            }
        }
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const int r = warp + 8 * j;
            if (r < LH) {
                sxy[r * LP + lane] = make_float2(va[j][0] - cx, vb[j][0] - cy);
                if (lane < LH - 32) sxy[r * LP + lane + 32] = make_float2(va[j][1] - cx, vb[j][1] - cy);
            }
        }
    }
    __syncthreads();

    // horizontal pass: item = (row, group of 4 columns)
    for (int i = tid; i < H_ITEMS; i += 256) {
        int r, cg;
        h_item(i, r, cg);
        if (r >= LH) continue;
        float2 v[LG + 10], sq[LG + 10];
        float xy[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) {
            v[k] = sxy[r * LP + cg + k];
            sq[k] = fmul2(v[k], v[k]);
            xy[k] = v[k].x * v[k].y;
        }
        float2 a01[LG], a23[LG];
        float a4[LG];
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            a01[o] = make_float2(0.f, 0.f);
            a23[o] = make_float2(0.f, 0.f);
            a4[o] = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                a01[o] = ffma2(win.w[k], v[o + k], a01[o]);
                a23[o] = ffma2(win.w[k], sq[o + k], a23[o]);
                a4[o] = fmaf(win.w[k], xy[o + k], a4[o]);
            }
        }
        float4* d01 = reinterpret_cast<float4*>(&h01[r][cg]);
        float4* d23 = reinterpret_cast<float4*>(&h23[r][cg]);
        d01[0] = make_float4(a01[0].x, a01[0].y, a01[1].x, a01[1].y);
        d01[1] = make_float4(a01[2].x, a01[2].y, a01[3].x, a01[3].y);
        d23[0] = make_float4(a23[0].x, a23[0].y, a23[1].x, a23[1].y);
        d23[1] = make_float4(a23[2].x, a23[2].y, a23[3].x, a23[3].y);
        *reinterpret_cast<float4*>(&h4[r][cg]) = make_float4(a4[0], a4[1], a4[2], a4[3]);
    }
    __syncthreads();

    // vertical pass: thread = (column, group of 4 rows)
    const int col = lane, rg = warp * LG;
    float2 o01[LG], o23[LG];
    float o4[LG];
    {
        float2 v01[LG + 10], v23[LG + 10];
        float v4[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) {
            v01[k] = h01[rg + k][col];
            v23[k] = h23[rg + k][col];
            v4[k] = h4[rg + k][col];
        }
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            o01[o] = make_float2(0.f, 0.f);
            o23[o] = make_float2(0.f, 0.f);
            o4[o] = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                o01[o] = ffma2(win.w[k], v01[o + k], o01[o]);
                o23[o] = ffma2(win.w[k], v23[o + k], o23[o]);
                o4[o] = fmaf(win.w[k], v4[o + k], o4[o]);
            }
        }
    }

    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const size_t n = (size_t)s.planes * s.H * s.W;
    const size_t pbase = (size_t)plane * s.H * s.W;
    const float* __restrict__ pmask = mask_plane(mask, s, plane);
    float sum_l1 = 0.f, sum_ds = 0.f;
    const int gx = tx * LT + col;
    if (gx < s.W) {
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            const int gy = ty * LT + rg + o;
            if (gy < s.H) {
                const float a0 = o01[o].x, a1 = o01[o].y;
                const float mu1 = cx + a0, mu2 = cy + a1;
                const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
                const float s1 = o23[o].x - a0 * a0, s2 = o23[o].y - a1 * a1, s12 = o4[o] - a0 * a1;
                const float N1 = 2.f * mu12 + C1, N2 = 2.f * s12 + C2;
                const float D1 = mu1_sq + mu2_sq + C1, D2 = s1 + s2 + C2;
                const float inv = __fdividef(1.f, D1 * D2);  // D1 D2 >= C1 C2 = 9e-8: no range issue; 1 ulp
                const float S = N1 * N2 * inv;
                const size_t off = (size_t)gy * s.W + gx;
                const float m = pmask ? pmask[off] : 1.f;
                // (x - cx) - (y - cy) + (cx - cy) would round differently from the reference's x - y: re-read
                const float xv = pimg[off], yv = pgt[off];  // L1 hits: the tile was just staged from these lines
                sum_l1 += fabsf((xv - yv) * m);
                sum_ds += (1.f - S) * m;
                if (MAP) ssim_map[pbase + off] = S;
                if (GRAD) {
                    const float invD1 = D2 * inv, invD2 = D1 * inv;
                    const float dS_dmu1 = 2.f * N2 * inv * (mu2 - mu1 * N1 * invD1);
                    const float dS_ds1 = -S * invD2;
                    const float dS_ds12 = 2.f * N1 * inv;
                    dmaps[pbase + off] = dS_dmu1 - 2.f * mu1 * dS_ds1 - mu2 * dS_ds12;
                    dmaps[n + pbase + off] = dS_ds1;
                    dmaps[2 * n + pbase + off] = dS_ds12;
                }
            }
        }
    }
    // block sums
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum_l1 += __shfl_xor_sync(0xffffffffu, sum_l1, d);
        sum_ds += __shfl_xor_sync(0xffffffffu, sum_ds, d);
    }
    if (lane == 0) {
        red[0][warp] = sum_l1;
        red[1][warp] = sum_ds;
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
        const float2 p = reinterpret_cast<const float2*>(partials)[i];
        a += (double)p.x;
        b += (double)p.y;
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
__global__ void __launch_bounds__(256, 4)
image_loss_backward_kernel(LossShape s, LossWindow win, const float* __restrict__ img, const float* __restrict__ gt,
                           const float* __restrict__ mask, const float* __restrict__ dmaps,
                           const float* __restrict__ grad_map, const float* __restrict__ grad_scale, float w_ssim,
                           float w_l1, float* __restrict__ dimg) {
    __shared__ __align__(16) float2 sab[LH * LP];   // (g A, g B)
    __shared__ float sc[LH * LP];                   // g C
    __shared__ __align__(16) float2 hab[LH][LT];
    __shared__ __align__(16) float hc[LH][LT];

    const int plane = blockIdx.y;
    const int ty = blockIdx.x / s.tiles_x, tx = blockIdx.x - ty * s.tiles_x;
    const int x0 = tx * LT - LR, y0 = ty * LT - LR;
    const size_t pbase = (size_t)plane * s.H * s.W;
    const size_t n = (size_t)s.planes * s.H * s.W;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float scale = grad_scale ? *grad_scale : 1.f;
    const float ws = w_ssim * scale;
    const float* __restrict__ pmask = mask_plane(mask, s, plane);
    const float* __restrict__ pA = dmaps + pbase;
    const float* __restrict__ pB = dmaps + n + pbase;
    const float* __restrict__ pC = dmaps + 2 * n + pbase;
    const float* __restrict__ pG = grad_map ? grad_map + pbase : nullptr;

    {
        constexpr int NR = (LH + 7) / 8;
        float va[NR][2], vb[NR][2], vc[NR][2], vg[NR][2];
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const int r = warp + 8 * j, gy = y0 + r;
            const bool rok = r < LH && gy >= 0 && gy < s.H;
            const int rowoff = (rok ? gy : 0) * s.W;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = lane + 32 * h, gx = x0 + c;
                const bool ok = rok && c < LH && gx >= 0 && gx < s.W;
                const int off = ok ? rowoff + gx : 0;
                va[j][h] = ok ? pA[off] : 0.f;
                vb[j][h] = ok ? pB[off] : 0.f;
                vc[j][h] = ok ? pC[off] : 0.f;
                vg[j][h] = !ok ? 0.f : (pG ? pG[off] : (pmask ? ws * pmask[off] : ws));
            }
        }
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const int r = warp + 8 * j;
            if (r < LH) {
                sab[r * LP + lane] = make_float2(vg[j][0] * va[j][0], vg[j][0] * vb[j][0]);
                sc[r * LP + lane] = vg[j][0] * vc[j][0];
                if (lane < LH - 32) {
                    sab[r * LP + lane + 32] = make_float2(vg[j][1] * va[j][1], vg[j][1] * vb[j][1]);
                    sc[r * LP + lane + 32] = vg[j][1] * vc[j][1];
                }
            }
        }
    }
    __syncthreads();

    for (int i = tid; i < H_ITEMS; i += 256) {
        int r, cg;
        h_item(i, r, cg);
        if (r >= LH) continue;
        float2 v[LG + 10];
        float u[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) {
            v[k] = sab[r * LP + cg + k];
            u[k] = sc[r * LP + cg + k];
        }
        float2 aab[LG];
        float ac[LG];
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            aab[o] = make_float2(0.f, 0.f);
            ac[o] = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                aab[o] = ffma2(win.w[k], v[o + k], aab[o]);
                ac[o] = fmaf(win.w[k], u[o + k], ac[o]);
            }
        }
        float4* dab = reinterpret_cast<float4*>(&hab[r][cg]);
        dab[0] = make_float4(aab[0].x, aab[0].y, aab[1].x, aab[1].y);
        dab[1] = make_float4(aab[2].x, aab[2].y, aab[3].x, aab[3].y);
        *reinterpret_cast<float4*>(&hc[r][cg]) = make_float4(ac[0], ac[1], ac[2], ac[3]);
    }
    __syncthreads();

    const int col = lane, rg = warp * LG;
    float2 oab[LG];
    float oc[LG];
    {
        float2 v[LG + 10];
        float u[LG + 10];
#pragma unroll
        for (int k = 0; k < LG + 10; ++k) {
            v[k] = hab[rg + k][col];
            u[k] = hc[rg + k][col];
        }
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            oab[o] = make_float2(0.f, 0.f);
            oc[o] = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                oab[o] = ffma2(win.w[k], v[o + k], oab[o]);
                oc[o] = fmaf(win.w[k], u[o + k], oc[o]);
            }
        }
    }
    const int gx = tx * LT + col;
    const float l1w = w_l1 * scale;
    if (gx < s.W) {
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            const int gy = ty * LT + rg + o;
            if (gy < s.H) {
                const size_t off = (size_t)gy * s.W + gx;
                const float xv = img[pbase + off], yv = gt[pbase + off];
                float g = oab[o].x + 2.f * xv * oab[o].y + yv * oc[o];
                if (l1w != 0.f) {
                    const float m = pmask ? pmask[off] : 1.f;
                    const float d = (xv - yv) * m;
                    g += l1w * m * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
                }
                dimg[pbase + off] = g;
            }
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
    auto kernel = ssim_map ? (dmaps ? image_loss_forward_kernel<true, true> : image_loss_forward_kernel<true, false>)
                           : (dmaps ? image_loss_forward_kernel<false, true> : image_loss_forward_kernel<false, false>);
    CSB_LAUNCH("image_loss_forward", stream, kernel<<<grid, 256, 0, stream>>>(
        s, win, img, gt, mask, ssim_map, dmaps, (float*)workspace));
    const size_t blocks = (size_t)grid.x * grid.y;
    const double inv_count = 1.0 / ((double)s.planes * H * W);
    CSB_LAUNCH("image_loss_finalize", stream, image_loss_finalize_kernel<<<1, 256, 0, stream>>>(
        blocks, inv_count, lambda_dssim, (const float*)workspace, losses));
    return 0;
}

// ---- csb_image_loss_forward_static: the same two kernels, partial sums in a library-owned array ----------------
// For callers that cannot allocate (stream capture, serving loops): no cudaMalloc / cudaFree / allocating routine
// anywhere below.  Every device that loads the library holds CSB_IMAGE_LOSS_STATIC_BYTES of it until exit.
}  // extern "C"

static __device__ __align__(256) unsigned char g_loss_static_partials[CSB_IMAGE_LOSS_STATIC_BYTES];

// In host code the array's name is a host-side stand-in, not a device address; the device address is looked up
// once per device (with that device current) and kept.
static int loss_static_scratch(float** out) {
    constexpr int MAX_DEV = 64;
    static float* cached[MAX_DEV] = {};
    int dev = 0;
    CSB_CUDA_TRY(cudaGetDevice(&dev));
    float* p = (dev >= 0 && dev < MAX_DEV) ? __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE) : nullptr;
    if (!p) {
        void* sym = nullptr;
        CSB_CUDA_TRY(cudaGetSymbolAddress(&sym, g_loss_static_partials));
        if (!sym) return CSB_ERR_UNSUPPORTED;
        p = (float*)sym;
        if (dev >= 0 && dev < MAX_DEV) __atomic_store_n(&cached[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}

extern "C" {

size_t csb_image_loss_static_max_bytes(void) { return CSB_IMAGE_LOSS_STATIC_BYTES; }

int csb_image_loss_forward_static(int n_img, int C, int H, int W, const float* img, const float* gt,
                                  const float* mask, int mask_channels, float lambda_dssim, float* ssim_map,
                                  float* dmaps, float* losses, void* stream_) {
    LossShape s;
    if (!make_shape(s, n_img, C, H, W, mask_channels, mask != nullptr)) return CSB_ERR_BAD_ARG;
    if (n_img == 0) return CSB_ERR_BAD_ARG;
    if (!img || !gt || !losses) return CSB_ERR_BAD_ARG;
    if (s.planes > 65535) return CSB_ERR_UNSUPPORTED;
    const size_t blocks = (size_t)s.tiles_x * s.tiles_y * s.planes;
    if (blocks * 2 * sizeof(float) > CSB_IMAGE_LOSS_STATIC_BYTES) return CSB_ERR_STATIC_LIMIT;
    float* partials = nullptr;
    const int rc = loss_static_scratch(&partials);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    const dim3 grid(s.tiles_x * s.tiles_y, s.planes);
    const LossWindow win = make_window();
    // every block of the forward kernel overwrites its own two partials and the fold reads exactly `blocks` of them:
    // nothing of the previous call's contents is read, so the shared array needs no clearing between calls
    auto kernel = ssim_map ? (dmaps ? image_loss_forward_kernel<true, true> : image_loss_forward_kernel<true, false>)
                           : (dmaps ? image_loss_forward_kernel<false, true> : image_loss_forward_kernel<false, false>);
    CSB_LAUNCH("image_loss_forward", stream, kernel<<<grid, 256, 0, stream>>>(
        s, win, img, gt, mask, ssim_map, dmaps, partials));
    const double inv_count = 1.0 / ((double)s.planes * H * W);
    CSB_LAUNCH("image_loss_finalize", stream, image_loss_finalize_kernel<<<1, 256, 0, stream>>>(
        blocks, inv_count, lambda_dssim, (const float*)partials, losses));
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
