/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's tile rasterizer.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it; the product
 * path (cloth-splatting_b200/) never links or calls anything in oracle/.
 *
 * Parity status: the reference ships no tests or golden vectors for this path,
 * so the oracle is pinned against outputs of the reference's own CUDA code, compiled unmodified
 * into oracle/_ref/libref_cuda.so and run on the B200 box (tests/golden/, tests/make_golden.py).
 *
 * Follows (RAST = submodules/depth-diff-gaussian-rasterization of KTH-RPL/cloth-splatting):
 *   oracle_preprocess      RAST/cuda_rasterizer/forward.cu:20-71,74-113,118-152,155-256, auxiliary.h:41-77,139-164
 *   oracle_bin             RAST/cuda_rasterizer/rasterizer_impl.cu:35-50,70-138,276-319
 *   oracle_render          RAST/cuda_rasterizer/forward.cu:261-379
 *   oracle_render_backward RAST/cuda_rasterizer/backward.cu:399-557
 *   oracle_gaussian_backward RAST/cuda_rasterizer/backward.cu:20-139,144-274,278-341,346-396
 *
 * Floating point: fp32 throughout (ndc2Pix in fp64, auxiliary.h:41-44).  The GPU code is compiled
 * with FMA contraction; to stay as close to it as possible the sums of products below are written
 * with explicit fmaf() the way nvcc contracts them: in A*B + C*D the LEFT product is fused and the
 * right one rounded (read off the reference's SASS: FMUL dy*dy; FFMA dx*dx+.; FFMA dz*dz+. for
 * dx*dx + dy*dy + dz*dz), a product added to a non-product is fused into it,
 * and the file is compiled with -ffp-contract=off.  Integer results can still differ from the GPU
 * for inputs that sit exactly on a rounding boundary; the tests bound how often.
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

/* a*b + c*d + e*f the way nvcc contracts it */
static inline float dot3(float a, float b, float c, float d, float e, float f) {
    return fmaf(e, f, fmaf(a, b, c * d));
}
/* m0*x + m1*y + m2*z + m3 */
static inline float affine(float m0, float x, float m1, float y, float m2, float z, float m3) {
    return dot3(m0, x, m1, y, m2, z) + m3;
}

typedef struct { float c[3][3]; } mat3; /* c[col][row] */

static mat3 mat3_mul(const mat3* A, const mat3* B) {
    mat3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            R.c[col][row] = dot3(A->c[0][row], B->c[col][0], A->c[1][row], B->c[col][1], A->c[2][row], B->c[col][2]);
    return R;
}
static mat3 mat3_t(const mat3* A) {
    mat3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row) R.c[col][row] = A->c[row][col];
    return R;
}

static mat3 quat_to_mat3(const float* q) {
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    mat3 R;
    R.c[0][0] = fmaf(-2.f, fmaf(y, y, z * z), 1.f);
    R.c[0][1] = 2.f * fmaf(x, y, -(r * z));
    R.c[0][2] = 2.f * fmaf(x, z, r * y);
    R.c[1][0] = 2.f * fmaf(x, y, r * z);
    R.c[1][1] = fmaf(-2.f, fmaf(x, x, z * z), 1.f);
    R.c[1][2] = 2.f * fmaf(y, z, -(r * x));
    R.c[2][0] = 2.f * fmaf(x, z, -(r * y));
    R.c[2][1] = 2.f * fmaf(y, z, r * x);
    R.c[2][2] = fmaf(-2.f, fmaf(x, x, y * y), 1.f);
    return R;
}

static void cov3d_from_scale_rot(const float* s, float mod, const float* q, float* cov6) {
    mat3 S;
    memset(&S, 0, sizeof S);
    S.c[0][0] = mod * s[0];
    S.c[1][1] = mod * s[1];
    S.c[2][2] = mod * s[2];
    mat3 R = quat_to_mat3(q);
    mat3 M = mat3_mul(&S, &R);
    mat3 Mt = mat3_t(&M);
    mat3 Sg = mat3_mul(&Mt, &M);
    cov6[0] = Sg.c[0][0]; cov6[1] = Sg.c[0][1]; cov6[2] = Sg.c[0][2];
    cov6[3] = Sg.c[1][1]; cov6[4] = Sg.c[1][2]; cov6[5] = Sg.c[2][2];
}

typedef struct {
    mat3 T;
    float t[3];
    float txtz, tytz, limx, limy;
} ewa_t;

static ewa_t ewa_project(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy, const float* vm) {
    ewa_t e;
    float t[3];
    t[0] = affine(vm[0], mean[0], vm[4], mean[1], vm[8], mean[2], vm[12]);
    t[1] = affine(vm[1], mean[0], vm[5], mean[1], vm[9], mean[2], vm[13]);
    t[2] = affine(vm[2], mean[0], vm[6], mean[1], vm[10], mean[2], vm[14]);
    e.limx = 1.3f * tan_fovx;
    e.limy = 1.3f * tan_fovy;
    e.txtz = t[0] / t[2];
    e.tytz = t[1] / t[2];
    t[0] = fminf(e.limx, fmaxf(-e.limx, e.txtz)) * t[2];
    t[1] = fminf(e.limy, fmaxf(-e.limy, e.tytz)) * t[2];
    e.t[0] = t[0]; e.t[1] = t[1]; e.t[2] = t[2];
    mat3 J, W;
    memset(&J, 0, sizeof J);
    J.c[0][0] = fx / t[2];
    J.c[0][2] = -(fx * t[0]) / (t[2] * t[2]);
    J.c[1][1] = fy / t[2];
    J.c[1][2] = -(fy * t[1]) / (t[2] * t[2]);
    W.c[0][0] = vm[0]; W.c[0][1] = vm[4]; W.c[0][2] = vm[8];
    W.c[1][0] = vm[1]; W.c[1][1] = vm[5]; W.c[1][2] = vm[9];
    W.c[2][0] = vm[2]; W.c[2][1] = vm[6]; W.c[2][2] = vm[10];
    e.T = mat3_mul(&W, &J);
    return e;
}

static void cov2d_from_proj(const ewa_t* e, const float* c6, float* abc) {
    mat3 Vrk;
    Vrk.c[0][0] = c6[0]; Vrk.c[0][1] = c6[1]; Vrk.c[0][2] = c6[2];
    Vrk.c[1][0] = c6[1]; Vrk.c[1][1] = c6[3]; Vrk.c[1][2] = c6[4];
    Vrk.c[2][0] = c6[2]; Vrk.c[2][1] = c6[4]; Vrk.c[2][2] = c6[5];
    mat3 Tt = mat3_t(&e->T), Vt = mat3_t(&Vrk);
    mat3 A = mat3_mul(&Tt, &Vt);
    mat3 cov = mat3_mul(&A, &e->T);
    abc[0] = cov.c[0][0] + 0.3f;
    abc[1] = cov.c[0][1];
    abc[2] = cov.c[1][1] + 0.3f;
}

static inline float ndc_to_pix(float v, int S) { return (float)((((double)v + 1.0) * S - 1.0) * 0.5); }

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

static void tile_rect(float px, float py, int r, int gx, int gy, int* mn, int* mx) {
    mn[0] = imin(gx, imax(0, (int)((px - (float)r) / (float)TILE)));
    mn[1] = imin(gy, imax(0, (int)((py - (float)r) / (float)TILE)));
    mx[0] = imin(gx, imax(0, (int)((((px + (float)r) + (float)TILE) - 1.0f) / (float)TILE)));
    mx[1] = imin(gy, imax(0, (int)((((py + (float)r) + (float)TILE) - 1.0f) / (float)TILE)));
}

static void sh_to_rgb(int deg, const float* sh, const float* pos, const float* campos, float* rgb, uint8_t* clamped) {
    float d[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    const float len = sqrtf(fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1])));
    const float x = d[0] / len, y = d[1] / len, z = d[2] / len;
    for (int ch = 0; ch < 3; ++ch) {
#define S(k) sh[(k) * 3 + ch]
        float r = SH_C0 * S(0);
        if (deg > 0) {
            r = fmaf(-(SH_C1 * y), S(1), r);
            r = fmaf(SH_C1 * z, S(2), r);
            r = fmaf(-(SH_C1 * x), S(3), r);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = fmaf(SH_C2[0] * xy, S(4), r);
                r = fmaf(SH_C2[1] * yz, S(5), r);
                r = fmaf(SH_C2[2] * ((fmaf(2.0f, zz, -xx)) - yy), S(6), r);
                r = fmaf(SH_C2[3] * xz, S(7), r);
                r = fmaf(SH_C2[4] * (xx - yy), S(8), r);
                if (deg > 2) {
                    r = fmaf(SH_C3[0] * y * fmaf(3.0f, xx, -yy), S(9), r);
                    r = fmaf(SH_C3[1] * xy * z, S(10), r);
                    r = fmaf(SH_C3[2] * y * (fmaf(4.0f, zz, -xx) - yy), S(11), r);
                    r = fmaf(SH_C3[3] * z * fmaf(-3.0f, yy, fmaf(2.0f, zz, -(3.0f * xx))), S(12), r);
                    r = fmaf(SH_C3[4] * x * (fmaf(4.0f, zz, -xx) - yy), S(13), r);
                    r = fmaf(SH_C3[5] * z * (xx - yy), S(14), r);
                    r = fmaf(SH_C3[6] * x * fmaf(-3.0f, yy, xx), S(15), r);
                }
            }
        }
#undef S
        r += 0.5f;
        clamped[ch] = r < 0;
        rgb[ch] = fmaxf(r, 0.0f);
    }
}

/* Per-Gaussian forward.  Outputs are indexed by Gaussian; entries of culled Gaussians are zeroed
 * (the reference leaves them uninitialised and gates on radii > 0).  Returns sum(tiles_touched). */
int64_t oracle_preprocess(int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                          const float* colors_precomp, const float* opacities, const float* scales,
                          float scale_modifier, const float* rotations, const float* cov3D_precomp,
                          const float* vm, const float* pm, const float* campos, float tan_fovx, float tan_fovy,
                          int* radii, float* means2D, float* depths, float* cov3D, float* conic_opacity, float* rgb,
                          uint8_t* clamped, uint32_t* tiles_touched) {
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    int64_t total = 0;
#pragma omp parallel for schedule(static) reduction(+ : total)
    for (int i = 0; i < P; ++i) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        means2D[2 * i] = means2D[2 * i + 1] = 0.f;
        depths[i] = 0.f;
        for (int k = 0; k < 4; ++k) conic_opacity[4 * i + k] = 0.f;
        for (int k = 0; k < 3; ++k) { rgb[3 * i + k] = 0.f; clamped[3 * i + k] = 0; }
        if (!cov3D_precomp) for (int k = 0; k < 6; ++k) cov3D[6 * i + k] = 0.f;
        const float* p = means3D + 3 * i;
        const float pvz = affine(vm[2], p[0], vm[6], p[1], vm[10], p[2], vm[14]);
        if (pvz <= 0.2f) continue;
        const float hx = affine(pm[0], p[0], pm[4], p[1], pm[8], p[2], pm[12]);
        const float hy = affine(pm[1], p[0], pm[5], p[1], pm[9], p[2], pm[13]);
        const float hw = affine(pm[3], p[0], pm[7], p[1], pm[11], p[2], pm[15]);
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float projx = hx * p_w, projy = hy * p_w;
        float c6[6];
        if (cov3D_precomp) {
            memcpy(c6, cov3D_precomp + 6 * (size_t)i, sizeof c6);
        } else {
            cov3d_from_scale_rot(scales + 3 * i, scale_modifier, rotations + 4 * i, c6);
            memcpy(cov3D + 6 * (size_t)i, c6, sizeof c6);
        }
        ewa_t e = ewa_project(p, focal_x, focal_y, tan_fovx, tan_fovy, vm);
        float abc[3];
        cov2d_from_proj(&e, c6, abc);
        const float det = fmaf(abc[0], abc[2], -(abc[1] * abc[1]));
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float conic[3] = {abc[2] * det_inv, -abc[1] * det_inv, abc[0] * det_inv};
        const float mid = 0.5f * (abc[0] + abc[2]);
        const float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        const float l1 = mid + disc, l2 = mid - disc;
        const float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        const float px = ndc_to_pix(projx, W), py = ndc_to_pix(projy, H);
        int mn[2], mx[2];
        tile_rect(px, py, (int)my_radius, gx, gy, mn, mx);
        const int area = (mx[0] - mn[0]) * (mx[1] - mn[1]);
        if (area == 0) continue;
        if (!colors_precomp) {
            sh_to_rgb(D, shs + (size_t)i * M * 3, p, campos, rgb + 3 * i, clamped + 3 * i);
        } else {
            for (int k = 0; k < 3; ++k) rgb[3 * i + k] = colors_precomp[3 * i + k];
        }
        depths[i] = pvz;
        radii[i] = (int)my_radius;
        means2D[2 * i] = px;
        means2D[2 * i + 1] = py;
        conic_opacity[4 * i + 0] = conic[0];
        conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2];
        conic_opacity[4 * i + 3] = opacities[i];
        tiles_touched[i] = (uint32_t)area;
        total += area;
    }
    return total;
}

uint32_t oracle_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Binning: emit tile|depth keys in Gaussian order, stable LSD radix sort on the low
 * 32 + higher_msb(tiles) bits, tile ranges.  keys/point_list have R entries, ranges 2*tiles. */
int oracle_bin(int P, int W, int H, const int* radii, const float* means2D, const float* depths, int64_t R,
               uint64_t* keys, uint32_t* point_list, uint32_t* ranges) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    uint64_t* k0 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R ? R : 1));
    uint32_t* v0 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R ? R : 1));
    uint64_t* k1 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R ? R : 1));
    uint32_t* v1 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R ? R : 1));
    if (!k0 || !v0 || !k1 || !v1) return -1;
    int64_t off = 0;
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        int mn[2], mx[2];
        tile_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, mn, mx);
        uint32_t dbits;
        memcpy(&dbits, depths + i, 4);
        for (int y = mn[1]; y < mx[1]; ++y)
            for (int x = mn[0]; x < mx[0]; ++x) {
                if (off >= R) return -2;
                k0[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
                v0[off] = (uint32_t)i;
                ++off;
            }
    }
    if (off != R) return -3;
    const int end_bit = 32 + (int)oracle_higher_msb((uint32_t)(gx * gy));
    for (int shift = 0; shift < end_bit; shift += 8) {
        size_t cnt[257];
        memset(cnt, 0, sizeof cnt);
        for (int64_t i = 0; i < R; ++i) cnt[((k0[i] >> shift) & 255) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int64_t i = 0; i < R; ++i) {
            const size_t dst = cnt[(k0[i] >> shift) & 255]++;
            k1[dst] = k0[i];
            v1[dst] = v0[i];
        }
        uint64_t* tk = k0; k0 = k1; k1 = tk;
        uint32_t* tv = v0; v0 = v1; v1 = tv;
    }
    memcpy(keys, k0, sizeof(uint64_t) * (size_t)R);
    memcpy(point_list, v0, sizeof(uint32_t) * (size_t)R);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (int64_t i = 0; i < R; ++i) {
        const uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
    free(k0); free(v0); free(k1); free(v1);
    return 0;
}

static inline float pair_power(const float* co, float dx, float dy) {
    /* -0.5f*(a*dx*dx + c*dy*dy) - b*dx*dy with nvcc's contraction */
    const float s = fmaf(co[0] * dx, dx, co[2] * dy * dy);
    return fmaf(s, -0.5f, -(co[1] * dx * dy));
}

/* Front-to-back compositing of every pixel.  counts[0] += evaluated pairs, counts[1] += pairs that
 * passed power<=0 (one exp each), counts[2] += blended pairs (may be NULL). */
void oracle_render(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                   const float* rgb, const float* depths, const float* conic_opacity, const float* bg,
                   float* final_T, uint32_t* n_contrib, float* out_color, float* out_depth, int64_t* counts) {
    const int gx = (W + TILE - 1) / TILE;
    int64_t c_eval = 0, c_exp = 0, c_blend = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c_eval, c_exp, c_blend)
    for (int py = 0; py < H; ++py) {
        for (int px = 0; px < W; ++px) {
            const int tile = (py / TILE) * gx + (px / TILE);
            const uint32_t beg = ranges[2 * tile], end = ranges[2 * tile + 1];
            const float pxf = (float)px, pyf = (float)py;
            float T = 1.0f, C[3] = {0, 0, 0}, Dd = 0;
            uint32_t contributor = 0, last = 0;
            for (uint32_t k = beg; k < end; ++k) {
                contributor++;
                c_eval++;
                const uint32_t id = point_list[k];
                const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                const float* co = conic_opacity + 4 * (size_t)id;
                const float power = pair_power(co, dx, dy);
                if (power > 0.0f) continue;
                c_exp++;
                const float alpha = fminf(0.99f, co[3] * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) break;
                c_blend++;
                for (int ch = 0; ch < 3; ++ch) C[ch] = fmaf(rgb[3 * id + ch] * alpha, T, C[ch]);
                Dd = fmaf(depths[id] * alpha, T, Dd);
                T = test_T;
                last = contributor;
            }
            const size_t pix = (size_t)py * W + px;
            final_T[pix] = T;
            n_contrib[pix] = last;
            for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pix] = fmaf(T, bg[ch], C[ch]);
            out_depth[pix] = Dd;
        }
    }
    if (counts) { counts[0] += c_eval; counts[1] += c_exp; counts[2] += c_blend; }
}
