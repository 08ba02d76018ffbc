// Backward pass of the tile rasterizer.
//
// Replaces CudaRasterizer::Rasterizer::backward (RAST/cuda_rasterizer/rasterizer_impl.cu:343-437):
//   render_backward_kernel   <- renderCUDA (backward.cu:399-557)
//   gaussian_backward_kernel <- computeCov2DCUDA + preprocessCUDA (backward.cu:144-274, 346-396)
//                               with computeColorFromSH / computeCov3D backward (backward.cu:20-139, 278-341)
//
// The reference issues 9 global atomics per blended (pixel, Gaussian) pair.  Here the 32 pixels
// of a warp are summed with shuffles first and one lane adds the warp's total, so the atomic
// count drops by up to 32x; the per-pair arithmetic is unchanged.
#include "render_lanes.cuh"

namespace csb {

constexpr int BW_BATCH = 256;
constexpr int ACC_STRIDE = 12;  // dmean2D.xy, dconic.xyz, dopacity, dcolor.rgb, 3 pad

__global__ void __launch_bounds__(CSB_TILE_PIX)
render_backward_kernel(const ViewBatch vb, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                       const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                       const float4* __restrict__ rgb_depth, const float* __restrict__ final_T,
                       const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                       float* __restrict__ grad_acc, const int view0) {
    __shared__ uint32_t s_id[BW_BATCH];
    __shared__ float2 s_xy[BW_BATCH];
    __shared__ float4 s_co[BW_BATCH];
    __shared__ float4 s_cd[BW_BATCH];
    __shared__ uint32_t s_mask[BW_BATCH];  // bit w: the Gaussian can reach the pixel patch of warp w (patch_mask)
    __shared__ uint32_t s_max_last;

    grid_dep_wait();
    grid_dep_launch();
    const int v = view0 + blockIdx.z;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int lx = (warp & 1) * 8 + (lane & 7);
    const int ly = (warp >> 1) * 4 + (lane >> 3);
    const uint32_t px = blockIdx.x * CSB_TILE_X + lx;
    const uint32_t py = blockIdx.y * CSB_TILE_Y + ly;
    const bool inside = px < (uint32_t)vb.W && py < (uint32_t)vb.H;
    const float pxf = (float)px, pyf = (float)py;
    const float tile_x0 = (float)(blockIdx.x * CSB_TILE_X), tile_y0 = (float)(blockIdx.y * CSB_TILE_Y);
    const size_t hw = (size_t)vb.W * vb.H;
    const size_t pix = (size_t)py * vb.W + px;

    const uint2 range = ranges[(size_t)v * vb.tiles + blockIdx.y * vb.gx + blockIdx.x];
    const uint32_t len = range.y - range.x;
    if (len == 0) return;

    const float T_final = inside ? final_T[v * hw + pix] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[v * hw + pix] : 0u;

    // entries behind the last contributor of every pixel of the tile are never touched
    if (tid == 0) s_max_last = 0;
    __syncthreads();
    uint32_t wmax = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xFFFFFFFFu, wmax, o));
    if (lane == 0) atomicMax(&s_max_last, wmax);
    __syncthreads();
    const uint32_t used = min(s_max_last, len);  // list positions [0, used) matter
    if (used == 0) return;

    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f;
    if (inside) {
        dpix0 = dL_dpix[(v * 3 + 0) * hw + pix];
        dpix1 = dL_dpix[(v * 3 + 1) * hw + pix];
        dpix2 = dL_dpix[(v * 3 + 2) * hw + pix];
    }
    const float* __restrict__ bg = vb.background[v];
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];

    float2 accum01 = make_float2(0.f, 0.f), lastc01 = make_float2(0.f, 0.f);
    float accum2 = 0.f, lastc2 = 0.f, last_alpha = 0.f;
    const float2 dpix01 = make_float2(dpix0, dpix1);
    const float2 ddel = make_float2(0.5 * vb.W, 0.5 * vb.H);  // d(pixel coordinate) / d(ndc)
    const float bg_dot_dpixel = bg0 * dpix0 + bg1 * dpix1 + bg2 * dpix2;

    // which of the 9 accumulator components (mx my cx cy cw op c0 c1 c2) this lane ends up holding, see XCH below
    const bool up16 = (lane & 16) != 0, up8 = (lane & 8) != 0, up4 = (lane & 4) != 0, up2 = (lane & 2) != 0;
    int my_comp = -1;
    if (!(lane & 1)) {
        const int base = up16 ? 5 : 0;
        if (!up8) my_comp = up4 ? (up2 ? -1 : base + 2) : base + (up2 ? 1 : 0);
        else if (!up4) my_comp = up16 ? (up2 ? -1 : 8) : 3 + (up2 ? 1 : 0);
    }
#define XCH2(a, b, o, up)                                               \
    {                                                                   \
        const float snd_ = (up) ? (a) : (b);                            \
        const float kp_ = (up) ? (b) : (a);                             \
        (a) = kp_ + __shfl_xor_sync(0xFFFFFFFFu, snd_, (o));            \
    }
#define XCH1(a, o, up)                                                  \
    {                                                                   \
        const float snd_ = (up) ? (a) : 0.f;                            \
        const float kp_ = (up) ? 0.f : (a);                             \
        (a) = kp_ + __shfl_xor_sync(0xFFFFFFFFu, snd_, (o));            \
    }

    // opaque: keeps the shared-window addresses in registers instead of rematerialising them (S2R SR_CgaCtaId + LEA)
    // at each of their uses inside the divergent loop
    uint32_t a_xy = smem_addr(s_xy), a_co = smem_addr(s_co), a_cd = smem_addr(s_cd), a_id = smem_addr(s_id);
    asm volatile("" : "+r"(a_xy), "+r"(a_co), "+r"(a_cd), "+r"(a_id));
    const int rounds = (int)((used + BW_BATCH - 1) / BW_BATCH);
    int todo = (int)used;
    for (int r = 0; r < rounds; ++r, todo -= BW_BATCH) {
        __syncthreads();
        const uint32_t progress = (uint32_t)r * BW_BATCH + tid;
        if (progress < used) {
            const uint32_t id = point_list[range.x + used - progress - 1];
            const float2 xy = means2D[id];
            const float4 co = conic_opacity[id];
            s_id[tid] = id;
            s_xy[tid] = xy;
            s_co[tid] = co;
            s_cd[tid] = rgb_depth[id];
            s_mask[tid] = patch_mask(xy, co, tile_x0, tile_y0);
        }
        __syncthreads();
        const int cnt = min(BW_BATCH, todo);
        // entry j of this batch sits at list position used - 1 - (r * BW_BATCH + j)
        const uint32_t top = used - 1u - (uint32_t)r * BW_BATCH;
        for (int g = 0; g < cnt; g += 32) {
          const uint32_t mk = (g + lane < cnt) ? s_mask[g + lane] : 0u;
          uint32_t bits = __ballot_sync(0xFFFFFFFFu, (mk >> warp) & 1u);
          while (bits) {
            const int j = g + __ffs(bits) - 1;
            bits &= bits - 1;
            const uint32_t contributor = top - (uint32_t)j;
            if (contributor >= wmax) continue;  // warp-uniform
            bool valid = inside && contributor < last_contributor;
            const float2 xy = lds_f2(a_xy + (uint32_t)j * 8u);
            const float dx = xy.x - pxf, dy = xy.y - pyf;
            const float4 con_o = lds_f4(a_co + (uint32_t)j * 16u);
            const float power = -0.5f * (con_o.x * dx * dx + con_o.z * dy * dy) - con_o.y * dx * dy;
            valid = valid && !(power > 0.0f);
            // 2^(power log2 e) on the SFU: |power| <= 5.6 wherever a pair is blended (alpha >= 1/255), there the
            // result is within 4e-7 relative of the forward's expf, far inside the gradient tolerance
            const float G = exp2f_fast(power * 1.4426950408889634f);
            const float alpha = fminf(0.99f, con_o.w * G);
            valid = valid && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(0xFFFFFFFFu, valid)) continue;

            float g_mx = 0.f, g_my = 0.f, g_cx = 0.f, g_cy = 0.f, g_cw = 0.f, g_op = 0.f;
            float g_c0 = 0.f, g_c1 = 0.f, g_c2 = 0.f;
            if (valid) {
                // one SFU reciprocal (1 ulp; 1 - alpha is in [0.01, 1]) serves T / (1 - alpha) and T_final / (1 - alpha)
                const float inv_1ma = __fdividef(1.f, 1.f - alpha);
                T = T * inv_1ma;
                const float dchannel_dcolor = alpha * T;
                const float4 cd = lds_f4(a_cd + (uint32_t)j * 16u);
                // channel pairs on packed fp32x2 instructions (same roundings, half the issue slots)
                const float oma = 1.f - last_alpha;
                const float2 cd01 = make_float2(cd.x, cd.y);
                accum01 = ffma2_bcast(last_alpha, lastc01, fmul2_bcast(oma, accum01));
                accum2 = last_alpha * lastc2 + oma * accum2;
                lastc01 = cd01;
                lastc2 = cd.z;
                const float2 e01 = fmul2(ffma2_bcast(-1.f, accum01, cd01), dpix01);
                float dL_dalpha = e01.x + e01.y + (cd.z - accum2) * dpix2;
                const float2 gc01 = fmul2_bcast(dchannel_dcolor, dpix01);
                g_c0 = gc01.x;
                g_c1 = gc01.y;
                g_c2 = dchannel_dcolor * dpix2;
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;

                const float dL_dG = con_o.w * dL_dalpha;
                const float2 dxy = make_float2(dx, dy);
                const float2 gd = fmul2_bcast(G, dxy);  // (gdx, gdy)
                const float dG_ddelx = -gd.x * con_o.x - gd.y * con_o.y;
                const float dG_ddely = -gd.y * con_o.z - gd.x * con_o.y;
                const float2 gm = fmul2(fmul2_bcast(dL_dG, make_float2(dG_ddelx, dG_ddely)), ddel);
                g_mx = gm.x;
                g_my = gm.y;
                const float h = -0.5f * dL_dG;
                const float2 gcxy = fmul2_bcast(h, fmul2_bcast(gd.x, dxy));  // -0.5 gdx (dx, dy) dL_dG
                g_cx = gcxy.x;
                g_cy = gcxy.y;
                g_cw = h * gd.y * dy;
                g_op = G * dL_dalpha;
            }
            // Transposed butterfly: at every step a lane hands one half of its remaining partial sums to its
            // partner and keeps the other half (9 -> 5 -> 3 -> 2 -> 1 values: 12 shuffles instead of 9 x 5).
            // Afterwards lane l (even) holds the warp total of component comp_of_lane (set up before the loop).
            XCH2(g_mx, g_op, 16, up16); XCH2(g_my, g_c0, 16, up16); XCH2(g_cx, g_c1, 16, up16);
            XCH2(g_cy, g_c2, 16, up16); XCH1(g_cw, 16, up16);
            XCH2(g_mx, g_cy, 8, up8); XCH2(g_my, g_cw, 8, up8); XCH1(g_cx, 8, up8);
            XCH2(g_mx, g_cx, 4, up4); XCH1(g_my, 4, up4);
            XCH2(g_mx, g_my, 2, up2);
            g_mx += __shfl_xor_sync(0xFFFFFFFFu, g_mx, 1);
            if (my_comp >= 0) atomicAdd(&grad_acc[(size_t)lds_u32(a_id + (uint32_t)j * 4u) * ACC_STRIDE + my_comp], g_mx);
          }
        }
    }
}

// ----------------------------------------------------------------------------------
// per-Gaussian backward: one thread per Gaussian, views summed in a fixed order
// ----------------------------------------------------------------------------------
__device__ __forceinline__ float3 dnormvdv3(const float3 v, const float3 dv) {
    const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward for one view (backward.cu:20-139): adds this view's dL/dsh into dL_dsh
// (assign when `first`) and returns the mean gradient through the view direction.
__device__ __forceinline__ float3 sh_backward(int deg, int M, const float* __restrict__ sh, const float3 pos,
                                              const float3 campos, uint8_t clamped, float3 dL_dRGB,
                                              float* __restrict__ dL_dsh, bool first) {
    const float3 dir_orig = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    dL_dRGB.x *= (clamped & 1) ? 0 : 1;
    dL_dRGB.y *= (clamped & 2) ? 0 : 1;
    dL_dRGB.z *= (clamped & 4) ? 0 : 1;
    const float dl[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
    float w[16];
    for (int k = 0; k < 16; ++k) w[k] = 0.f;
    w[0] = kSH_C0;
    float xx = 0, yy = 0, zz = 0, xy = 0, yz = 0, xz = 0;
    if (deg > 0) {
        w[1] = -kSH_C1 * y;
        w[2] = kSH_C1 * z;
        w[3] = -kSH_C1 * x;
        if (deg > 1) {
            xx = x * x; yy = y * y; zz = z * z;
            xy = x * y; yz = y * z; xz = x * z;
            w[4] = kSH_C2[0] * xy;
            w[5] = kSH_C2[1] * yz;
            w[6] = kSH_C2[2] * (2.f * zz - xx - yy);
            w[7] = kSH_C2[3] * xz;
            w[8] = kSH_C2[4] * (xx - yy);
            if (deg > 2) {
                w[9] = kSH_C3[0] * y * (3.f * xx - yy);
                w[10] = kSH_C3[1] * xy * z;
                w[11] = kSH_C3[2] * y * (4.f * zz - xx - yy);
                w[12] = kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = kSH_C3[4] * x * (4.f * zz - xx - yy);
                w[14] = kSH_C3[5] * z * (xx - yy);
                w[15] = kSH_C3[6] * x * (xx - 3.f * yy);
            }
        }
    }
    const int ncoef = (deg + 1) * (deg + 1);
    for (int k = 0; k < M; ++k) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float val = k < ncoef ? w[k] * dl[ch] : 0.f;
            if (first) dL_dsh[k * 3 + ch] = val;
            else if (k < ncoef) dL_dsh[k * 3 + ch] += val;
        }
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
#define SHC(k) sh[(k) * 3 + ch]
        if (deg > 0) {
            dRGBdx[ch] = -kSH_C1 * SHC(3);
            dRGBdy[ch] = -kSH_C1 * SHC(1);
            dRGBdz[ch] = kSH_C1 * SHC(2);
            if (deg > 1) {
                dRGBdx[ch] += kSH_C2[0] * y * SHC(4) + kSH_C2[2] * 2.f * -x * SHC(6) + kSH_C2[3] * z * SHC(7) +
                              kSH_C2[4] * 2.f * x * SHC(8);
                dRGBdy[ch] += kSH_C2[0] * x * SHC(4) + kSH_C2[1] * z * SHC(5) + kSH_C2[2] * 2.f * -y * SHC(6) +
                              kSH_C2[4] * 2.f * -y * SHC(8);
                dRGBdz[ch] += kSH_C2[1] * y * SHC(5) + kSH_C2[2] * 2.f * 2.f * z * SHC(6) + kSH_C2[3] * x * SHC(7);
                if (deg > 2) {
                    dRGBdx[ch] += (kSH_C3[0] * SHC(9) * 3.f * 2.f * xy + kSH_C3[1] * SHC(10) * yz +
                                   kSH_C3[2] * SHC(11) * -2.f * xy + kSH_C3[3] * SHC(12) * -3.f * 2.f * xz +
                                   kSH_C3[4] * SHC(13) * (-3.f * xx + 4.f * zz - yy) +
                                   kSH_C3[5] * SHC(14) * 2.f * xz + kSH_C3[6] * SHC(15) * 3.f * (xx - yy));
                    dRGBdy[ch] += (kSH_C3[0] * SHC(9) * 3.f * (xx - yy) + kSH_C3[1] * SHC(10) * xz +
                                   kSH_C3[2] * SHC(11) * (-3.f * yy + 4.f * zz - xx) +
                                   kSH_C3[3] * SHC(12) * -3.f * 2.f * yz + kSH_C3[4] * SHC(13) * -2.f * xy +
                                   kSH_C3[5] * SHC(14) * -2.f * yz + kSH_C3[6] * SHC(15) * -3.f * 2.f * xy);
                    dRGBdz[ch] += (kSH_C3[1] * SHC(10) * xy + kSH_C3[2] * SHC(11) * 4.f * 2.f * yz +
                                   kSH_C3[3] * SHC(12) * 3.f * (2.f * zz - xx - yy) +
                                   kSH_C3[4] * SHC(13) * 4.f * 2.f * xz + kSH_C3[5] * SHC(14) * (xx - yy));
                }
            }
        }
#undef SHC
    }
    const float3 dL_ddir = make_float3(dRGBdx[0] * dl[0] + dRGBdx[1] * dl[1] + dRGBdx[2] * dl[2],
                                       dRGBdy[0] * dl[0] + dRGBdy[1] * dl[1] + dRGBdy[2] * dl[2],
                                       dRGBdz[0] * dl[0] + dRGBdz[1] * dl[1] + dRGBdz[2] * dl[2]);
    return dnormvdv3(dir_orig, dL_ddir);
}

// What the views of one Gaussian add up to (the sums over views of backward.cu's per-view kernels).
struct GaussPartial {
    float3 dmean;
    float dcov[6];
    float3 dcolor;
    float dop;
    bool any_visible;
};

// Views v_begin, v_begin + v_step, ... of one Gaussian: `sh_row` = its SH record, `my_dsh` = where this walk's dL/dsh
// record goes (both may live in shared memory; zero-filled if no view saw the Gaussian).
__device__ __forceinline__ GaussPartial
gaussian_backward_views(const int idx, const int v_begin, const int v_step, const ViewBatch& vb, int D, int M,
                        const float* __restrict__ means3D, const float* sh_row, const float* __restrict__ cov3D_in,
                        const int* __restrict__ radii, const uint8_t* __restrict__ clamped,
                        const float* __restrict__ grad_acc, float* __restrict__ dL_dmeans2D, float* my_dsh) {
    const float3 mean = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);

    float3 dmean_tot = make_float3(0.f, 0.f, 0.f);
    float dcov_tot[6] = {0, 0, 0, 0, 0, 0};
    float3 dcolor_tot = make_float3(0.f, 0.f, 0.f);
    float dop_tot = 0.f;
    bool first_sh = true, any_visible = false;

    for (int v = v_begin; v < vb.V; v += v_step) {
        const size_t e = (size_t)v * vb.P + idx;
        float* m2d = dL_dmeans2D + e * 3;
        if (!(radii[e] > 0)) {
            m2d[0] = 0.f; m2d[1] = 0.f; m2d[2] = 0.f;
            continue;
        }
        any_visible = true;
        // 12 floats (48 B, 16-byte aligned) per entry: three vector loads instead of nine scalar ones
        const float4* acc4 = reinterpret_cast<const float4*>(grad_acc + e * ACC_STRIDE);
        const float4 acc_a = acc4[0], acc_b = acc4[1], acc_c = acc4[2];
        const float2 dmean2D = make_float2(acc_a.x, acc_a.y);
        const float3 dL_dconic = make_float3(acc_a.z, acc_a.w, acc_b.x);
        m2d[0] = dmean2D.x; m2d[1] = dmean2D.y; m2d[2] = 0.f;
        dop_tot += acc_b.y;
        const float3 dL_dcol = make_float3(acc_b.z, acc_b.w, acc_c.x);
        dcolor_tot.x += dL_dcol.x; dcolor_tot.y += dL_dcol.y; dcolor_tot.z += dL_dcol.z;

        const float* __restrict__ vm = vb.viewmatrix[v];
        const float* __restrict__ proj = vb.projmatrix[v];
        const float h_x = vb.focal_x[v], h_y = vb.focal_y[v];
        const float* cov3D = cov3D_in + 6 * (size_t)idx;

        // ---- conic -> cov2D -> cov3D, T, mean (backward.cu:144-274) ----
        const EwaProj ep = ewa_project(mean, h_x, h_y, vb.tan_fovx[v], vb.tan_fovy[v], vm);
        const float3 t = ep.t;
        const float x_grad_mul = ep.txtz < -ep.limx || ep.txtz > ep.limx ? 0 : 1;
        const float y_grad_mul = ep.tytz < -ep.limy || ep.tytz > ep.limy ? 0 : 1;
        const Mat3& T = ep.T;
        const Mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4],
                                   cov3D[5]);
        const float3 abc = cov2d_from_proj(ep, cov3D);
        const float a = abc.x, b = abc.y, c = abc.z;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float dcov[6] = {0, 0, 0, 0, 0, 0};
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
            dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
            dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
            dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
            dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
            dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
            dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db +
                      2 * T.c[1][0] * T.c[1][1] * dL_dc;
            dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db +
                      2 * T.c[1][0] * T.c[1][2] * dL_dc;
            dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db +
                      2 * T.c[1][1] * T.c[1][2] * dL_dc;
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) dcov_tot[i] += dcov[i];

        const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
        const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
        const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
        const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
        const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
        const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

        // W as in ewa_project: columns (vm0,vm4,vm8), (vm1,vm5,vm9), (vm2,vm6,vm10)
        const float W00 = vm[0], W01 = vm[4], W02 = vm[8];
        const float W10 = vm[1], W11 = vm[5], W12 = vm[9];
        const float W20 = vm[2], W21 = vm[6], W22 = vm[10];
        const float dL_dJ00 = W00 * dL_dT00 + W01 * dL_dT01 + W02 * dL_dT02;
        const float dL_dJ02 = W20 * dL_dT00 + W21 * dL_dT01 + W22 * dL_dT02;
        const float dL_dJ11 = W10 * dL_dT10 + W11 * dL_dT11 + W12 * dL_dT12;
        const float dL_dJ12 = W20 * dL_dT10 + W21 * dL_dT11 + W22 * dL_dT12;

        const float tz = 1.f / t.z;
        const float tz2 = tz * tz;
        const float tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                             (2 * h_y * t.y) * tz3 * dL_dJ12;
        float3 dmean;  // transformVec4x3Transpose
        dmean.x = vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz;
        dmean.y = vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz;
        dmean.z = vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz;

        // ---- mean2D -> mean3D through the projection (backward.cu:370-387) ----
        const float4 m_hom = xform_point_4x4(mean, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        float3 dproj;
        dproj.x = (proj[0] * m_w - proj[3] * mul1) * dmean2D.x + (proj[1] * m_w - proj[3] * mul2) * dmean2D.y;
        dproj.y = (proj[4] * m_w - proj[7] * mul1) * dmean2D.x + (proj[5] * m_w - proj[7] * mul2) * dmean2D.y;
        dproj.z = (proj[8] * m_w - proj[11] * mul1) * dmean2D.x + (proj[9] * m_w - proj[11] * mul2) * dmean2D.y;
        dmean.x += dproj.x; dmean.y += dproj.y; dmean.z += dproj.z;

        // ---- colour -> SH and view direction (backward.cu:20-139) ----
        if (sh_row != nullptr) {
            const float3 campos = make_float3(vb.campos[v][0], vb.campos[v][1], vb.campos[v][2]);
            const float3 dsh_mean = sh_backward(D, M, sh_row, mean, campos, clamped[e], dL_dcol, my_dsh, first_sh);
            first_sh = false;
            dmean.x += dsh_mean.x; dmean.y += dsh_mean.y; dmean.z += dsh_mean.z;
        }
        dmean_tot.x += dmean.x; dmean_tot.y += dmean.y; dmean_tot.z += dmean.z;
    }

    if (my_dsh && first_sh)
        for (int k = 0; k < M * 3; ++k) my_dsh[k] = 0.f;
    GaussPartial r;
    r.dmean = dmean_tot;
#pragma unroll
    for (int i = 0; i < 6; ++i) r.dcov[i] = dcov_tot[i];
    r.dcolor = dcolor_tot;
    r.dop = dop_tot;
    r.any_visible = any_visible;
    return r;
}

// Writes the per-Gaussian outputs from the sum over views, then cov3D -> scale, rotation (backward.cu:278-341;
// linear in dL_dcov3D, so done once).
__device__ __forceinline__ void
gaussian_backward_finish(const int idx, const GaussPartial& tot, const float* __restrict__ scales, float scale_modifier,
                         const float* __restrict__ rotations, float* __restrict__ dL_dcolors,
                         float* __restrict__ dL_dopacity, float* __restrict__ dL_dmeans3D,
                         float* __restrict__ dL_dcov3D, float* __restrict__ dL_dscales, float* __restrict__ dL_drots) {
    const float* dcov_tot = tot.dcov;
    const bool any_visible = tot.any_visible;
    dL_dmeans3D[3 * idx + 0] = tot.dmean.x;
    dL_dmeans3D[3 * idx + 1] = tot.dmean.y;
    dL_dmeans3D[3 * idx + 2] = tot.dmean.z;
    dL_dcolors[3 * idx + 0] = tot.dcolor.x;
    dL_dcolors[3 * idx + 1] = tot.dcolor.y;
    dL_dcolors[3 * idx + 2] = tot.dcolor.z;
    dL_dopacity[idx] = tot.dop;
#pragma unroll
    for (int i = 0; i < 6; ++i) dL_dcov3D[6 * (size_t)idx + i] = dcov_tot[i];

    float3 dscale = make_float3(0.f, 0.f, 0.f);
    float4 drot = make_float4(0.f, 0.f, 0.f, 0.f);
    if (scales != nullptr && any_visible) {
        const float4 q = reinterpret_cast<const float4*>(rotations)[idx];
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const Mat3 R = quat_to_mat3(q);
        Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
        const float3 s = make_float3(scale_modifier * scales[3 * idx], scale_modifier * scales[3 * idx + 1],
                                     scale_modifier * scales[3 * idx + 2]);
        S.c[0][0] = s.x; S.c[1][1] = s.y; S.c[2][2] = s.z;
        const Mat3 Mm = mat3_mul(S, R);
        const float* d = dcov_tot;
        const Mat3 dL_dSigma = mat3_cols(d[0], 0.5f * d[1], 0.5f * d[2], 0.5f * d[1], d[3], 0.5f * d[4],
                                         0.5f * d[2], 0.5f * d[4], d[5]);
        Mat3 M2;
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
#pragma unroll
            for (int rr = 0; rr < 3; ++rr) M2.c[cc][rr] = 2.0f * Mm.c[cc][rr];
        const Mat3 dL_dM = mat3_mul(M2, dL_dSigma);
        const Mat3 Rt = mat3_transpose(R);
        Mat3 dL_dMt = mat3_transpose(dL_dM);
        dscale.x = Rt.c[0][0] * dL_dMt.c[0][0] + Rt.c[0][1] * dL_dMt.c[0][1] + Rt.c[0][2] * dL_dMt.c[0][2];
        dscale.y = Rt.c[1][0] * dL_dMt.c[1][0] + Rt.c[1][1] * dL_dMt.c[1][1] + Rt.c[1][2] * dL_dMt.c[1][2];
        dscale.z = Rt.c[2][0] * dL_dMt.c[2][0] + Rt.c[2][1] * dL_dMt.c[2][1] + Rt.c[2][2] * dL_dMt.c[2][2];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            dL_dMt.c[0][k] *= s.x;
            dL_dMt.c[1][k] *= s.y;
            dL_dMt.c[2][k] *= s.z;
        }
        const Mat3& Q = dL_dMt;
        drot.x = 2 * z * (Q.c[0][1] - Q.c[1][0]) + 2 * y * (Q.c[2][0] - Q.c[0][2]) + 2 * x * (Q.c[1][2] - Q.c[2][1]);
        drot.y = 2 * y * (Q.c[1][0] + Q.c[0][1]) + 2 * z * (Q.c[2][0] + Q.c[0][2]) + 2 * r * (Q.c[1][2] - Q.c[2][1]) -
                 4 * x * (Q.c[2][2] + Q.c[1][1]);
        drot.z = 2 * x * (Q.c[1][0] + Q.c[0][1]) + 2 * r * (Q.c[2][0] - Q.c[0][2]) + 2 * z * (Q.c[1][2] + Q.c[2][1]) -
                 4 * y * (Q.c[2][2] + Q.c[0][0]);
        drot.w = 2 * r * (Q.c[0][1] - Q.c[1][0]) + 2 * x * (Q.c[2][0] + Q.c[0][2]) + 2 * y * (Q.c[1][2] + Q.c[2][1]) -
                 4 * z * (Q.c[1][1] + Q.c[0][0]);
    }
    if (dL_dscales) {
        dL_dscales[3 * idx + 0] = dscale.x;
        dL_dscales[3 * idx + 1] = dscale.y;
        dL_dscales[3 * idx + 2] = dscale.z;
    }
    if (dL_drots) reinterpret_cast<float4*>(dL_drots)[idx] = drot;
}

// One thread per Gaussian; the block's SH records (192 B each at degree 3) and its dL/dsh records are staged through
// shared memory so that global memory only sees coalesced 16-byte accesses (the per-thread walk over an AoS record
// touches 32 different lines per warp instruction).  Row pitch M*3+1 words: conflict-free per-thread access.
__global__ void __launch_bounds__(128)
gaussian_backward_kernel(const ViewBatch vb, int D, int M, const float* __restrict__ means3D,
                         const float* __restrict__ shs, const float* __restrict__ scales, float scale_modifier,
                         const float* __restrict__ rotations, const float* __restrict__ cov3D_in,
                         const int* __restrict__ radii, const uint8_t* __restrict__ clamped,
                         const float* __restrict__ grad_acc, float* __restrict__ dL_dmeans2D,
                         float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
                         float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D,
                         float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                         float* __restrict__ dL_drots) {
    extern __shared__ float s_rec[];  // [2][blockDim.x][M*3+1]: SH in, dL/dsh out
    grid_dep_wait();
    grid_dep_launch();
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int first = blockIdx.x * nthr;
    const int idx = first + tid;
    const int rec = M * 3, pitch = rec + 1;
    const int count = min(nthr, vb.P - first);
    const bool staged = shs != nullptr && dL_dsh != nullptr;
    if (staged) {
        if (rec == 48) stage_records_in<48>(shs + (size_t)first * rec, s_rec, count, rec, tid, nthr);
        else stage_records_in<0>(shs + (size_t)first * rec, s_rec, count, rec, tid, nthr);
        __syncthreads();
    }
    if (idx < vb.P) {
        const GaussPartial tot = gaussian_backward_views(
            idx, 0, 1, vb, D, M, means3D, staged ? s_rec + tid * pitch : (shs ? shs + (size_t)idx * rec : nullptr),
            cov3D_in, radii, clamped, grad_acc, dL_dmeans2D,
            staged ? s_rec + (nthr + tid) * pitch : (dL_dsh ? dL_dsh + (size_t)idx * rec : nullptr));
        gaussian_backward_finish(idx, tot, scales, scale_modifier, rotations, dL_dcolors, dL_dopacity, dL_dmeans3D,
                                 dL_dcov3D, dL_dscales, dL_drots);
    }
    if (staged) {
        __syncthreads();
        if (rec == 48) stage_records_out<48>(dL_dsh + (size_t)first * rec, s_rec + nthr * pitch, count, rec, tid, nthr);
        else stage_records_out<0>(dL_dsh + (size_t)first * rec, s_rec + nthr * pitch, count, rec, tid, nthr);
    }
}

// Few Gaussians, several views (the cloth: 10 k Gaussians x 4 views): one thread per Gaussian leaves the GPU at 3 %
// occupancy and the kernel is one thread's serial latency.  Here a block is 32 Gaussians x S view lanes (warp s walks
// views s, s + S, ...); the lanes' partial sums and dL/dsh records meet in shared memory and are added in lane order
// (deterministic), then warp 0 finishes the Gaussian.
constexpr int GB_SPLIT_G = 32;
constexpr int GB_PART = 14;  // dmean 3, dcov 6, dcolor 3, dop 1, any_visible 1

__global__ void __launch_bounds__(GB_SPLIT_G * 8)
gaussian_backward_split_kernel(const ViewBatch vb, int S, int D, int M, const float* __restrict__ means3D,
                               const float* __restrict__ shs, const float* __restrict__ scales, float scale_modifier,
                               const float* __restrict__ rotations, const float* __restrict__ cov3D_in,
                               const int* __restrict__ radii, const uint8_t* __restrict__ clamped,
                               const float* __restrict__ grad_acc, float* __restrict__ dL_dmeans2D,
                               float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
                               float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D,
                               float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                               float* __restrict__ dL_drots) {
    extern __shared__ float s_rec[];  // SH in [G][pitch] | dL/dsh [S][G][pitch] | partials [S][G][GB_PART]
    grid_dep_wait();
    grid_dep_launch();
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int gl = tid & (GB_SPLIT_G - 1), vl = tid / GB_SPLIT_G;
    const int first = blockIdx.x * GB_SPLIT_G;
    const int idx = first + gl;
    const int rec = M * 3, pitch = rec + 1;
    const int count = min(GB_SPLIT_G, vb.P - first);
    const bool has_sh = shs != nullptr && dL_dsh != nullptr;
    float* s_in = s_rec;
    float* s_dsh = s_rec + (has_sh ? GB_SPLIT_G * pitch : 0);
    float* s_part = s_dsh + (has_sh ? S * GB_SPLIT_G * pitch : 0);
    if (has_sh) {
        if (rec == 48) stage_records_in<48>(shs + (size_t)first * rec, s_in, count, rec, tid, nthr);
        else stage_records_in<0>(shs + (size_t)first * rec, s_in, count, rec, tid, nthr);
        __syncthreads();
    }
    if (idx < vb.P) {
        const GaussPartial p = gaussian_backward_views(
            idx, vl, S, vb, D, M, means3D, has_sh ? s_in + gl * pitch : nullptr, cov3D_in, radii, clamped, grad_acc,
            dL_dmeans2D, has_sh ? s_dsh + (vl * GB_SPLIT_G + gl) * pitch : nullptr);
        float* q = s_part + (vl * GB_SPLIT_G + gl) * GB_PART;
        q[0] = p.dmean.x; q[1] = p.dmean.y; q[2] = p.dmean.z;
#pragma unroll
        for (int i = 0; i < 6; ++i) q[3 + i] = p.dcov[i];
        q[9] = p.dcolor.x; q[10] = p.dcolor.y; q[11] = p.dcolor.z;
        q[12] = p.dop;
        q[13] = p.any_visible ? 1.f : 0.f;
    }
    __syncthreads();
    if (vl == 0 && idx < vb.P) {
        float a[GB_PART];
#pragma unroll
        for (int i = 0; i < GB_PART; ++i) a[i] = s_part[gl * GB_PART + i];
        for (int s = 1; s < S; ++s)
#pragma unroll
            for (int i = 0; i < GB_PART; ++i) a[i] += s_part[(s * GB_SPLIT_G + gl) * GB_PART + i];
        GaussPartial tot;
        tot.dmean = make_float3(a[0], a[1], a[2]);
#pragma unroll
        for (int i = 0; i < 6; ++i) tot.dcov[i] = a[3 + i];
        tot.dcolor = make_float3(a[9], a[10], a[11]);
        tot.dop = a[12];
        tot.any_visible = a[13] > 0.f;
        gaussian_backward_finish(idx, tot, scales, scale_modifier, rotations, dL_dcolors, dL_dopacity, dL_dmeans3D,
                                 dL_dcov3D, dL_dscales, dL_drots);
    }
    if (has_sh) {
        // dL/dsh = sum over the view lanes, in lane order; coalesced 16-byte stores
        float* dst = dL_dsh + (size_t)first * rec;
        const int total = count * rec;
        const bool vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
        for (int i = tid * 4; i < total; i += nthr * 4) {
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = i + j;
                float acc = 0.f;
                if (f < total) {
                    const int o = (f / rec) * pitch + (f % rec);
                    acc = s_dsh[o];
                    for (int s = 1; s < S; ++s) acc += s_dsh[s * GB_SPLIT_G * pitch + o];
                }
                v[j] = acc;
            }
            if (vec && i + 3 < total) *reinterpret_cast<float4*>(dst + i) = make_float4(v[0], v[1], v[2], v[3]);
            else
                for (int j = 0; j < 4 && i + j < total; ++j) dst[i + j] = v[j];
        }
    }
}

}  // namespace csb

using namespace csb;

static int launch_render_backward(cudaStream_t stream, dim3 grid, const ViewBatch& vb, const uint2* ranges,
                                  const uint32_t* point_list, const float2* means2D, const float4* conic_opacity,
                                  const float4* rgb_depth, const float* final_T, const uint32_t* n_contrib,
                                  const float* dL_dpix, float* grad_acc, int view0, const uint32_t* tile_order) {
    // all views of the set in one launch: a 1-D grid over the (view, tile) pairs, composited longest list first when the
    // forward left an order (render_lanes.cuh: tile_of_block); the per-view launches of the staged backward keep (x, y, 1)
    const dim3 lgrid = tile_order ? dim3(grid.x * grid.y * grid.z) : grid;
#define CSB_RB_LANES(NP, MINB, ...)                                                                                            \
    CSB_LAUNCH("render_backward", stream,                                                                                      \
               launch_pdl(render_backward_lanes_kernel<NP, MINB, __VA_ARGS__>, lgrid, dim3(LaneTiling<NP>::THREADS), 0, stream, \
                          vb, ranges, point_list, means2D, conic_opacity, rgb_depth, final_T, n_contrib, dL_dpix, grad_acc, view0, tile_order))
    switch ((g_render_variant >> 4) & 15) {
        case 0:
            CSB_LAUNCH("render_backward", stream, launch_pdl(render_backward_kernel, grid, dim3(CSB_TILE_PIX), 0, stream, vb, ranges, point_list, means2D,
                                                             conic_opacity, rgb_depth, final_T, n_contrib, dL_dpix, grad_acc, view0));
            break;
        case 2: CSB_RB_LANES(2, 1, false); break;
        case 3: CSB_RB_LANES(1, 6, false); break;
        case 4: CSB_RB_LANES(1, 1, true); break;
        case 5: CSB_RB_LANES(1, 6, true); break;
        case 6: CSB_RB_LANES(2, 1, true); break;
        case 7: CSB_RB_LANES(1, 5, true, true); break;
        default: CSB_RB_LANES(1, 1, false);
    }
#undef CSB_RB_LANES
    return 0;
}

extern "C" {

static int backward_impl(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp,
                         const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const int* radii, const void* geom_ws,
                         const void* image_ws, const void* binning_ws, size_t binning_capacity,
                         const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                         float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                         float* dL_drotations, void* const* view_ready, void* stream_) {
    (void)colors_precomp;
    cudaStream_t stream = (cudaStream_t)stream_;
    ViewBatch vb;
    int rc = fill_view_batch(vb, num_views, views, P, width, height);
    if (rc) return rc;
    if (!geom_ws || !image_ws || !radii || !dL_dpix) return CSB_ERR_BAD_ARG;
    if (!dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dcov3D) return CSB_ERR_BAD_ARG;
    if (shs != nullptr && dL_dsh == nullptr) return CSB_ERR_BAD_ARG;
    if (P == 0) return 0;
    GeomState g = GeomState::carve((void*)geom_ws, num_views, P);
    ImageState im = ImageState::carve((void*)image_ws, num_views, width, height);
    BinState b = BinState::carve((void*)binning_ws, num_views, P, width, height, binning_capacity);
    const size_t n = (size_t)num_views * P;
    const size_t total_tiles = (size_t)num_views * vb.tiles;
    CSB_CUDA_TRY(cudaMemsetAsync(g.grad_acc, 0, n * ACC_STRIDE * sizeof(float), stream));
    if (binning_capacity > 0 && binning_ws) {
        const uint32_t* point_list = sorted_side(total_tiles) ? b.vals_b : b.vals_a;
        if (view_ready) {
            // dL/dpixel arrives view by view (e.g. over PCIe on another stream): view v is composited backwards as soon as
            // its event has fired, while the later views are still in flight
            for (int v = 0; v < num_views; ++v) {
                if (view_ready[v]) CSB_CUDA_TRY(cudaStreamWaitEvent(stream, (cudaEvent_t)view_ready[v], 0));
                rc = launch_render_backward(stream, dim3(vb.gx, vb.gy, 1), vb, im.ranges, point_list, g.means2D, g.conic_opacity,
                                            g.rgb_depth, im.final_T, im.n_contrib, dL_dpix, g.grad_acc, v, nullptr);
                if (rc) return rc;
            }
        } else {
            dim3 tgrid(vb.gx, vb.gy, num_views);
            rc = launch_render_backward(stream, tgrid, vb, im.ranges, point_list, g.means2D, g.conic_opacity, g.rgb_depth,
                                        im.final_T, im.n_contrib, dL_dpix, g.grad_acc, 0, im.tile_order);
            if (rc) return rc;
        }
    }
    const float* cov3D = cov3D_precomp ? cov3D_precomp : g.cov3D;
    const int gb = P < 148 * 256 ? 64 : 128;  // few Gaussians (C2: 10 k): smaller blocks reach every SM
    const size_t gsmem = (shs && dL_dsh) ? (size_t)2 * gb * (M * 3 + 1) * sizeof(float) : 0;
    {
        static unsigned long long configured = 0;
        CSB_CUDA_TRY(opt_in_dynamic_smem(gaussian_backward_kernel, 64 * 1024, &configured));
    }
    if (gsmem > 64 * 1024) return CSB_ERR_UNSUPPORTED;
    const int S = num_views >= 8 ? 8 : (num_views >= 4 ? 4 : num_views);
    const size_t pitch = (size_t)M * 3 + 1;
    const size_t ssmem = ((shs && dL_dsh) ? (size_t)(1 + S) * GB_SPLIT_G * pitch : 0) * sizeof(float) +
                         (size_t)S * GB_SPLIT_G * GB_PART * sizeof(float);
    if (S >= 2 && P < 148 * 256 && ssmem <= 64 * 1024) {
        static unsigned long long configured_split = 0;
        CSB_CUDA_TRY(opt_in_dynamic_smem(gaussian_backward_split_kernel, 64 * 1024, &configured_split));
        CSB_LAUNCH("gaussian_backward", stream, launch_pdl(gaussian_backward_split_kernel, dim3((P + GB_SPLIT_G - 1) / GB_SPLIT_G), dim3(GB_SPLIT_G * S), ssmem, stream,
            vb, S, D, M, means3D, shs, cov3D_precomp ? nullptr : scales, scale_modifier, rotations, cov3D, radii, g.clamped,
            g.grad_acc, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations));
        return 0;
    }
    CSB_LAUNCH("gaussian_backward", stream, launch_pdl(gaussian_backward_kernel, dim3((P + gb - 1) / gb), dim3(gb), gsmem, stream, 
        vb, D, M, means3D, shs, cov3D_precomp ? nullptr : scales, scale_modifier, rotations, cov3D, radii, g.clamped,
        g.grad_acc, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations));
    return 0;
}

int csb_rasterize_backward(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* scales, float scale_modifier, const float* rotations,
                           const float* cov3D_precomp, const int* radii, const void* geom_ws,
                           const void* image_ws, const void* binning_ws, size_t binning_capacity,
                           const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                           float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                           float* dL_drotations, void* stream_) {
    return backward_impl(num_views, views, P, D, M, width, height, means3D, shs, colors_precomp, scales, scale_modifier,
                         rotations, cov3D_precomp, radii, geom_ws, image_ws, binning_ws, binning_capacity, dL_dpix,
                         dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
                         nullptr, stream_);
}

int csb_rasterize_backward_staged(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                                  const float* means3D, const float* shs, const float* colors_precomp,
                                  const float* scales, float scale_modifier, const float* rotations,
                                  const float* cov3D_precomp, const int* radii, const void* geom_ws,
                                  const void* image_ws, const void* binning_ws, size_t binning_capacity,
                                  const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                                  float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                                  float* dL_drotations, void* const* view_ready, void* stream_) {
    if (!view_ready) return CSB_ERR_BAD_ARG;
    return backward_impl(num_views, views, P, D, M, width, height, means3D, shs, colors_precomp, scales, scale_modifier,
                         rotations, cov3D_precomp, radii, geom_ws, image_ws, binning_ws, binning_capacity, dL_dpix,
                         dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
                         view_ready, stream_);
}

}  // extern "C"
