/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference rasterizer's backward pass.
 * See rast_oracle.c for the rules; same parity status.
 *
 * Follows RAST/cuda_rasterizer/backward.cu:399-557 (per-pixel replay), :144-274 (conic -> cov2D ->
 * cov3D / mean), :346-396 (projection, SH, cov3D -> scale/rotation), :20-139, :278-341.
 * Per-Gaussian sums over pixels are accumulated in double (the GPU adds floats atomically in an
 * unspecified order; the comparison tolerance is 1e-4 relative).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* Back-to-front replay.  acc is double[P*9]: dmean2D.xy, dconic.x/.y/.w, dopacity, dcolor.rgb. */
void oracle_render_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                            const float* means2D, const float* rgb, const float* conic_opacity, const float* bg,
                            const float* final_T, const uint32_t* n_contrib, const float* dL_dpix, double* acc) {
    const int gx = (W + TILE - 1) / TILE;
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    double* priv = (double*)calloc((size_t)nthreads * P * 9, sizeof(double));
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
#pragma omp parallel
    {
        int tidx = 0;
#ifdef _OPENMP
        tidx = omp_get_thread_num();
#endif
        double* my = priv + (size_t)tidx * P * 9;
#pragma omp for schedule(dynamic, 4)
        for (int py = 0; py < H; ++py) {
            for (int px = 0; px < W; ++px) {
                const int tile = (py / TILE) * gx + (px / TILE);
                const uint32_t beg = ranges[2 * tile], end = ranges[2 * tile + 1];
                const size_t pix = (size_t)py * W + px;
                const float pxf = (float)px, pyf = (float)py;
                const float T_final = final_T[pix];
                float T = T_final;
                const uint32_t last_contributor = n_contrib[pix];
                float dpix[3];
                for (int ch = 0; ch < 3; ++ch) dpix[ch] = dL_dpix[(size_t)ch * H * W + pix];
                float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0;
                uint32_t contributor = end - beg;
                for (uint32_t k = end; k-- > beg;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    const uint32_t id = point_list[k];
                    const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                    const float* co = conic_opacity + 4 * (size_t)id;
                    const float s = fmaf(co[0] * dx, dx, co[2] * dy * dy);
                    const float power = fmaf(s, -0.5f, -(co[1] * dx * dy));
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    double* g = my + (size_t)id * 9;
                    for (int ch = 0; ch < 3; ++ch) {
                        const float c = rgb[3 * id + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dalpha += (c - accum_rec[ch]) * dpix[ch];
                        g[6 + ch] += (double)(dchannel_dcolor * dpix[ch]);
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    float bg_dot = 0;
                    for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dpix[ch];
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                    const float dL_dG = co[3] * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
                    g[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
                    g[1] += (double)(dL_dG * dG_ddely * ddely_dy);
                    g[2] += (double)(-0.5f * gdx * dx * dL_dG);
                    g[3] += (double)(-0.5f * gdx * dy * dL_dG);
                    g[4] += (double)(-0.5f * gdy * dy * dL_dG);
                    g[5] += (double)(G * dL_dalpha);
                }
            }
        }
    }
    for (int t = 0; t < nthreads; ++t)
        for (size_t i = 0; i < (size_t)P * 9; ++i) acc[i] += priv[(size_t)t * P * 9 + i];
    free(priv);
}

typedef struct { float c[3][3]; } mat3;
static mat3 mmul(const mat3* A, const mat3* B) {
    mat3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            R.c[col][row] = A->c[0][row] * B->c[col][0] + A->c[1][row] * B->c[col][1] + A->c[2][row] * B->c[col][2];
    return R;
}
static mat3 mtr(const mat3* A) {
    mat3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row) R.c[col][row] = A->c[row][col];
    return R;
}

/* Per-Gaussian backward for ONE view.  Inputs acc (double[P*9], from oracle_render_backward) are
 * rounded to float first, like the GPU's float accumulators.  All outputs are assigned
 * (zero for Gaussians with radii <= 0). */
void oracle_gaussian_backward(int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                              const float* scales, float scale_modifier, const float* rotations,
                              const float* cov3D, const float* vm, const float* proj, const float* campos,
                              float tan_fovx, float tan_fovy, const int* radii, const uint8_t* clamped,
                              const double* acc, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                              float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                              float* dL_drots) {
    const float h_y = H / (2.0f * tan_fovy);
    const float h_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        for (int k = 0; k < 3; ++k) { dL_dmeans2D[3 * i + k] = 0; dL_dcolors[3 * i + k] = 0; dL_dmeans3D[3 * i + k] = 0; }
        dL_dopacity[i] = 0;
        for (int k = 0; k < 6; ++k) dL_dcov3D[6 * i + k] = 0;
        if (dL_dsh) for (int k = 0; k < M * 3; ++k) dL_dsh[(size_t)i * M * 3 + k] = 0;
        if (dL_dscales) for (int k = 0; k < 3; ++k) dL_dscales[3 * i + k] = 0;
        if (dL_drots) for (int k = 0; k < 4; ++k) dL_drots[4 * i + k] = 0;
        if (!(radii[i] > 0)) continue;
        const double* a9 = acc + (size_t)i * 9;
        const float dmean2D[2] = {(float)a9[0], (float)a9[1]};
        const float dconic[3] = {(float)a9[2], (float)a9[3], (float)a9[4]};
        const float dcol[3] = {(float)a9[6], (float)a9[7], (float)a9[8]};
        dL_dmeans2D[3 * i] = dmean2D[0];
        dL_dmeans2D[3 * i + 1] = dmean2D[1];
        dL_dopacity[i] = (float)a9[5];
        for (int k = 0; k < 3; ++k) dL_dcolors[3 * i + k] = dcol[k];
        const float* mean = means3D + 3 * i;
        const float* c6 = cov3D + 6 * (size_t)i;

        /* ---- conic -> cov2D -> cov3D / T / mean ---- */
        float t[3];
        t[0] = vm[0] * mean[0] + vm[4] * mean[1] + vm[8] * mean[2] + vm[12];
        t[1] = vm[1] * mean[0] + vm[5] * mean[1] + vm[9] * mean[2] + vm[13];
        t[2] = vm[2] * mean[0] + vm[6] * mean[1] + vm[10] * mean[2] + vm[14];
        const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
        const float txtz = t[0] / t[2], tytz = t[1] / t[2];
        t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
        t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
        const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
        const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
        mat3 J, Wm, Vrk;
        memset(&J, 0, sizeof J);
        J.c[0][0] = h_x / t[2]; J.c[0][2] = -(h_x * t[0]) / (t[2] * t[2]);
        J.c[1][1] = h_y / t[2]; J.c[1][2] = -(h_y * t[1]) / (t[2] * t[2]);
        Wm.c[0][0] = vm[0]; Wm.c[0][1] = vm[4]; Wm.c[0][2] = vm[8];
        Wm.c[1][0] = vm[1]; Wm.c[1][1] = vm[5]; Wm.c[1][2] = vm[9];
        Wm.c[2][0] = vm[2]; Wm.c[2][1] = vm[6]; Wm.c[2][2] = vm[10];
        Vrk.c[0][0] = c6[0]; Vrk.c[0][1] = c6[1]; Vrk.c[0][2] = c6[2];
        Vrk.c[1][0] = c6[1]; Vrk.c[1][1] = c6[3]; Vrk.c[1][2] = c6[4];
        Vrk.c[2][0] = c6[2]; Vrk.c[2][1] = c6[4]; Vrk.c[2][2] = c6[5];
        const mat3 T = mmul(&Wm, &J);
        const mat3 Tt = mtr(&T), Vt = mtr(&Vrk);
        const mat3 A = mmul(&Tt, &Vt);
        const mat3 cov2D = mmul(&A, &T);
        const float a = cov2D.c[0][0] + 0.3f, b = cov2D.c[0][1], c = cov2D.c[1][1] + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dconic[0] + 2 * b * c * dconic[1] + (denom - a * c) * dconic[2]);
            dL_dc = denom2inv * (-a * a * dconic[2] + 2 * a * b * dconic[1] + (denom - a * c) * dconic[0]);
            dL_db = denom2inv * 2 * (b * c * dconic[0] - (denom + 2 * b * b) * dconic[1] + a * b * dconic[2]);
            float* dc = dL_dcov3D + 6 * (size_t)i;
#define TT(x, y) T.c[x][y]
            dc[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
            dc[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
            dc[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
            dc[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
            dc[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
            dc[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
        }
#define VV(x, y) Vrk.c[x][y]
        const float dT00 = 2 * (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_da + (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_db;
        const float dT01 = 2 * (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_da + (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_db;
        const float dT02 = 2 * (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_da + (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_db;
        const float dT10 = 2 * (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_dc + (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_db;
        const float dT11 = 2 * (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_dc + (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_db;
        const float dT12 = 2 * (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_dc + (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_db;
#undef VV
#undef TT
        const float dJ00 = Wm.c[0][0] * dT00 + Wm.c[0][1] * dT01 + Wm.c[0][2] * dT02;
        const float dJ02 = Wm.c[2][0] * dT00 + Wm.c[2][1] * dT01 + Wm.c[2][2] * dT02;
        const float dJ11 = Wm.c[1][0] * dT10 + Wm.c[1][1] * dT11 + Wm.c[1][2] * dT12;
        const float dJ12 = Wm.c[2][0] * dT10 + Wm.c[2][1] * dT11 + Wm.c[2][2] * dT12;
        const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = x_grad_mul * -h_x * tz2 * dJ02;
        const float dty = y_grad_mul * -h_y * tz2 * dJ12;
        const float dtz = -h_x * tz2 * dJ00 - h_y * tz2 * dJ11 + (2 * h_x * t[0]) * tz3 * dJ02 + (2 * h_y * t[1]) * tz3 * dJ12;
        float dmean[3];
        dmean[0] = vm[0] * dtx + vm[1] * dty + vm[2] * dtz;
        dmean[1] = vm[4] * dtx + vm[5] * dty + vm[6] * dtz;
        dmean[2] = vm[8] * dtx + vm[9] * dty + vm[10] * dtz;

        /* ---- mean2D -> mean3D ---- */
        const float mw_ = proj[3] * mean[0] + proj[7] * mean[1] + proj[11] * mean[2] + proj[15];
        const float m_w = 1.0f / (mw_ + 0.0000001f);
        const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * m_w * m_w;
        dmean[0] += (proj[0] * m_w - proj[3] * mul1) * dmean2D[0] + (proj[1] * m_w - proj[3] * mul2) * dmean2D[1];
        dmean[1] += (proj[4] * m_w - proj[7] * mul1) * dmean2D[0] + (proj[5] * m_w - proj[7] * mul2) * dmean2D[1];
        dmean[2] += (proj[8] * m_w - proj[11] * mul1) * dmean2D[0] + (proj[9] * m_w - proj[11] * mul2) * dmean2D[1];

        /* ---- colour -> SH, view direction ---- */
        if (shs) {
            const float* sh = shs + (size_t)i * M * 3;
            float* dsh = dL_dsh + (size_t)i * M * 3;
            const float dor[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
            const float len = sqrtf(dor[0] * dor[0] + dor[1] * dor[1] + dor[2] * dor[2]);
            const float x = dor[0] / len, y = dor[1] / len, z = dor[2] / len;
            float dl[3];
            for (int ch = 0; ch < 3; ++ch) dl[ch] = dcol[ch] * (clamped[3 * i + ch] ? 0 : 1);
            float w[16];
            memset(w, 0, sizeof w);
            float xx = 0, yy = 0, zz = 0, xy = 0, yz = 0, xz = 0;
            w[0] = SH_C0;
            if (D > 0) {
                w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x;
                if (D > 1) {
                    xx = x * x; yy = y * y; zz = z * z; xy = x * y; yz = y * z; xz = x * z;
                    w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2.f * zz - xx - yy);
                    w[7] = SH_C2[3] * xz; w[8] = SH_C2[4] * (xx - yy);
                    if (D > 2) {
                        w[9] = SH_C3[0] * y * (3.f * xx - yy);
                        w[10] = SH_C3[1] * xy * z;
                        w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
                        w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                        w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
                        w[14] = SH_C3[5] * z * (xx - yy);
                        w[15] = SH_C3[6] * x * (xx - 3.f * yy);
                    }
                }
            }
            const int ncoef = (D + 1) * (D + 1);
            for (int k = 0; k < ncoef; ++k)
                for (int ch = 0; ch < 3; ++ch) dsh[k * 3 + ch] = w[k] * dl[ch];
            float ddir[3] = {0, 0, 0};
            for (int ch = 0; ch < 3; ++ch) {
#define S(k) sh[(k) * 3 + ch]
                float gx_ = 0, gy_ = 0, gz_ = 0;
                if (D > 0) {
                    gx_ = -SH_C1 * S(3); gy_ = -SH_C1 * S(1); gz_ = SH_C1 * S(2);
                    if (D > 1) {
                        gx_ += SH_C2[0] * y * S(4) + SH_C2[2] * 2.f * -x * S(6) + SH_C2[3] * z * S(7) + SH_C2[4] * 2.f * x * S(8);
                        gy_ += SH_C2[0] * x * S(4) + SH_C2[1] * z * S(5) + SH_C2[2] * 2.f * -y * S(6) + SH_C2[4] * 2.f * -y * S(8);
                        gz_ += SH_C2[1] * y * S(5) + SH_C2[2] * 2.f * 2.f * z * S(6) + SH_C2[3] * x * S(7);
                        if (D > 2) {
                            gx_ += (SH_C3[0] * S(9) * 3.f * 2.f * xy + SH_C3[1] * S(10) * yz + SH_C3[2] * S(11) * -2.f * xy +
                                    SH_C3[3] * S(12) * -3.f * 2.f * xz + SH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                                    SH_C3[5] * S(14) * 2.f * xz + SH_C3[6] * S(15) * 3.f * (xx - yy));
                            gy_ += (SH_C3[0] * S(9) * 3.f * (xx - yy) + SH_C3[1] * S(10) * xz +
                                    SH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * S(12) * -3.f * 2.f * yz +
                                    SH_C3[4] * S(13) * -2.f * xy + SH_C3[5] * S(14) * -2.f * yz + SH_C3[6] * S(15) * -3.f * 2.f * xy);
                            gz_ += (SH_C3[1] * S(10) * xy + SH_C3[2] * S(11) * 4.f * 2.f * yz +
                                    SH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * S(13) * 4.f * 2.f * xz +
                                    SH_C3[5] * S(14) * (xx - yy));
                        }
                    }
                }
#undef S
                ddir[0] += gx_ * dl[ch]; ddir[1] += gy_ * dl[ch]; ddir[2] += gz_ * dl[ch];
            }
            const float sum2 = dor[0] * dor[0] + dor[1] * dor[1] + dor[2] * dor[2];
            const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
            dmean[0] += ((+sum2 - dor[0] * dor[0]) * ddir[0] - dor[1] * dor[0] * ddir[1] - dor[2] * dor[0] * ddir[2]) * inv32;
            dmean[1] += (-dor[0] * dor[1] * ddir[0] + (sum2 - dor[1] * dor[1]) * ddir[1] - dor[2] * dor[1] * ddir[2]) * inv32;
            dmean[2] += (-dor[0] * dor[2] * ddir[0] - dor[1] * dor[2] * ddir[1] + (sum2 - dor[2] * dor[2]) * ddir[2]) * inv32;
        }
        for (int k = 0; k < 3; ++k) dL_dmeans3D[3 * i + k] = dmean[k];

        /* ---- cov3D -> scale, rotation ---- */
        if (scales) {
            const float* q = rotations + 4 * i;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            mat3 R, S, dSig;
            R.c[0][0] = 1.f - 2.f * (y * y + z * z); R.c[0][1] = 2.f * (x * y - r * z); R.c[0][2] = 2.f * (x * z + r * y);
            R.c[1][0] = 2.f * (x * y + r * z); R.c[1][1] = 1.f - 2.f * (x * x + z * z); R.c[1][2] = 2.f * (y * z - r * x);
            R.c[2][0] = 2.f * (x * z - r * y); R.c[2][1] = 2.f * (y * z + r * x); R.c[2][2] = 1.f - 2.f * (x * x + y * y);
            memset(&S, 0, sizeof S);
            const float s[3] = {scale_modifier * scales[3 * i], scale_modifier * scales[3 * i + 1], scale_modifier * scales[3 * i + 2]};
            S.c[0][0] = s[0]; S.c[1][1] = s[1]; S.c[2][2] = s[2];
            mat3 Mm = mmul(&S, &R);
            const float* d = dL_dcov3D + 6 * (size_t)i;
            dSig.c[0][0] = d[0]; dSig.c[0][1] = 0.5f * d[1]; dSig.c[0][2] = 0.5f * d[2];
            dSig.c[1][0] = 0.5f * d[1]; dSig.c[1][1] = d[3]; dSig.c[1][2] = 0.5f * d[4];
            dSig.c[2][0] = 0.5f * d[2]; dSig.c[2][1] = 0.5f * d[4]; dSig.c[2][2] = d[5];
            for (int cc = 0; cc < 3; ++cc) for (int rr = 0; rr < 3; ++rr) Mm.c[cc][rr] *= 2.0f;
            const mat3 dM = mmul(&Mm, &dSig);
            const mat3 Rt = mtr(&R);
            mat3 Q = mtr(&dM);
            dL_dscales[3 * i + 0] = Rt.c[0][0] * Q.c[0][0] + Rt.c[0][1] * Q.c[0][1] + Rt.c[0][2] * Q.c[0][2];
            dL_dscales[3 * i + 1] = Rt.c[1][0] * Q.c[1][0] + Rt.c[1][1] * Q.c[1][1] + Rt.c[1][2] * Q.c[1][2];
            dL_dscales[3 * i + 2] = Rt.c[2][0] * Q.c[2][0] + Rt.c[2][1] * Q.c[2][1] + Rt.c[2][2] * Q.c[2][2];
            for (int k = 0; k < 3; ++k) { Q.c[0][k] *= s[0]; Q.c[1][k] *= s[1]; Q.c[2][k] *= s[2]; }
            float* dq = dL_drots + 4 * i;
            dq[0] = 2 * z * (Q.c[0][1] - Q.c[1][0]) + 2 * y * (Q.c[2][0] - Q.c[0][2]) + 2 * x * (Q.c[1][2] - Q.c[2][1]);
            dq[1] = 2 * y * (Q.c[1][0] + Q.c[0][1]) + 2 * z * (Q.c[2][0] + Q.c[0][2]) + 2 * r * (Q.c[1][2] - Q.c[2][1]) - 4 * x * (Q.c[2][2] + Q.c[1][1]);
            dq[2] = 2 * x * (Q.c[1][0] + Q.c[0][1]) + 2 * r * (Q.c[2][0] - Q.c[0][2]) + 2 * z * (Q.c[1][2] + Q.c[2][1]) - 4 * y * (Q.c[2][2] + Q.c[0][0]);
            dq[3] = 2 * r * (Q.c[0][1] - Q.c[1][0]) + 2 * x * (Q.c[2][0] + Q.c[0][2]) + 2 * y * (Q.c[1][2] + Q.c[2][1]) - 4 * z * (Q.c[1][1] + Q.c[0][0]);
        }
    }
}
