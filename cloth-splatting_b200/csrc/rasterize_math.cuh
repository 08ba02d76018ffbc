// Per-Gaussian EWA arithmetic shared by the forward and backward kernels.
//
// The integer results (radii, tile rectangles, depth keys) must be bit-identical to the
// reference, so every expression below keeps the operand order and association of the formula
// it implements (RAST/cuda_rasterizer/forward.cu:74-152,
// auxiliary.h:41-77) and is compiled with nvcc's default FMA contraction, like the reference.
#pragma once
#include "common.cuh"

namespace csb {

// 3x3 matrix stored as columns (c[col][row]), the convention the reference's formulas use.
struct Mat3 {
    float c[3][3];
};

__device__ __forceinline__ Mat3 mat3_cols(float a0, float a1, float a2, float b0, float b1, float b2, float c0,
                                          float c1, float c2) {
    Mat3 m;
    m.c[0][0] = a0; m.c[0][1] = a1; m.c[0][2] = a2;
    m.c[1][0] = b0; m.c[1][1] = b1; m.c[1][2] = b2;
    m.c[2][0] = c0; m.c[2][1] = c1; m.c[2][2] = c2;
    return m;
}

// (A*B)[col][row] = sum_k A[k][row] * B[col][k]
__device__ __forceinline__ Mat3 mat3_mul(const Mat3& A, const Mat3& B) {
    Mat3 R;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            R.c[col][row] = A.c[0][row] * B.c[col][0] + A.c[1][row] * B.c[col][1] + A.c[2][row] * B.c[col][2];
    return R;
}

__device__ __forceinline__ Mat3 mat3_transpose(const Mat3& A) {
    Mat3 R;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row) R.c[col][row] = A.c[row][col];
    return R;
}

__device__ __forceinline__ float3 xform_point_4x3(const float3 p, const float* __restrict__ m) {
    float3 r;
    r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
    r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
    r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
    return r;
}

__device__ __forceinline__ float4 xform_point_4x4(const float3 p, const float* __restrict__ m) {
    float4 r;
    r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
    r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
    r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
    r.w = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
    return r;
}

// Rotation (columns as the reference writes them) from an un-normalised quaternion (r,x,y,z).
__device__ __forceinline__ Mat3 quat_to_mat3(const float4 q) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                     2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                     2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// Sigma = (S R)^T (S R); upper triangle (forward.cu:118-152).
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 rot, float* cov6) {
    Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    S.c[0][0] = mod * scale.x;
    S.c[1][1] = mod * scale.y;
    S.c[2][2] = mod * scale.z;
    const Mat3 R = quat_to_mat3(rot);
    const Mat3 M = mat3_mul(S, R);
    const Mat3 Sigma = mat3_mul(mat3_transpose(M), M);
    cov6[0] = Sigma.c[0][0];
    cov6[1] = Sigma.c[0][1];
    cov6[2] = Sigma.c[0][2];
    cov6[3] = Sigma.c[1][1];
    cov6[4] = Sigma.c[1][2];
    cov6[5] = Sigma.c[2][2];
}

// Intermediate matrices of the EWA projection, needed again by the backward pass.
struct EwaProj {
    Mat3 T;        // W * J
    float3 t;      // clamped view-space mean
    float txtz, tytz, limx, limy;
};

__device__ __forceinline__ EwaProj ewa_project(const float3 mean, float focal_x, float focal_y, float tan_fovx,
                                               float tan_fovy, const float* __restrict__ vm) {
    EwaProj e;
    float3 t = xform_point_4x3(mean, vm);
    e.limx = 1.3f * tan_fovx;
    e.limy = 1.3f * tan_fovy;
    e.txtz = t.x / t.z;
    e.tytz = t.y / t.z;
    t.x = fminf(e.limx, fmaxf(-e.limx, e.txtz)) * t.z;
    t.y = fminf(e.limy, fmaxf(-e.limy, e.tytz)) * t.z;
    e.t = t;
    const Mat3 J = mat3_cols(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
                             0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
                             0.0f, 0.0f, 0.0f);
    const Mat3 W = mat3_cols(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
    e.T = mat3_mul(W, J);
    return e;
}

// cov2D = T^T Vrk^T T with the 0.3 px low-pass on the diagonal; returns (a, b, c) (forward.cu:74-113).
__device__ __forceinline__ float3 cov2d_from_proj(const EwaProj& e, const float* __restrict__ cov6) {
    const Mat3 Vrk = mat3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
    const Mat3 cov = mat3_mul(mat3_mul(mat3_transpose(e.T), mat3_transpose(Vrk)), e.T);
    return make_float3(cov.c[0][0] + 0.3f, cov.c[0][1], cov.c[1][1] + 0.3f);
}

// Pixel coordinate of an NDC coordinate; evaluated in double like the reference (auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle touched by a disc of integer radius around p (auxiliary.h:46-56).
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, int gx, int gy, uint2& rect_min,
                                          uint2& rect_max) {
    rect_min.x = min(gx, max((int)0, (int)((p.x - max_radius) / CSB_TILE_X)));
    rect_min.y = min(gy, max((int)0, (int)((p.y - max_radius) / CSB_TILE_Y)));
    rect_max.x = min(gx, max((int)0, (int)((p.x + max_radius + CSB_TILE_X - 1) / CSB_TILE_X)));
    rect_max.y = min(gy, max((int)0, (int)((p.y + max_radius + CSB_TILE_Y - 1) / CSB_TILE_Y)));
}

// Real spherical-harmonics constants (auxiliary.h:21-39).
__device__ const float kSH_C0 = 0.28209479177387814f;
__device__ const float kSH_C1 = 0.4886025119029199f;
__device__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                    -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                    0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                    -0.5900435899266435f};


// Which of the eight 8x4 pixel patches of a 16x16 tile (bit w = patch of warp w: column w & 1, row w >> 1)
// a Gaussian can reach.  A pixel is only blended when alpha = o * exp(-q/2) >= 1/255 with
// q = a dx^2 + 2 b dx dy + c dy^2, i.e. inside the ellipse q <= t = 2 ln(255 o); the axis-aligned box of that
// ellipse has half extents sqrt(t c / det), sqrt(t a / det).  The test is conservative (0.1 % + 0.02 px
// of slack, far above the fp32 rounding of the per-pixel expression), so skipping the other patches does
// not change any result of the per-pixel loops of forward.cu:309-367 / backward.cu:464-556.
__device__ __forceinline__ uint32_t patch_mask(const float2 xy, const float4 con_o, float tile_x0, float tile_y0) {
    const float o = con_o.w;
    if (!(o * 255.0f >= 0.999f)) return 0u;  // alpha <= o < 1/255 everywhere
    const float det = con_o.x * con_o.z - con_o.y * con_o.y;
    if (!(det > 0.0f) || !(con_o.x > 0.0f) || !(con_o.z > 0.0f)) return 0xFFu;
    const float t = 2.0f * logf(fminf(o, 1e30f) * 255.0f) * 1.001f + 1e-3f;
    const float inv = t / det;
    const float ex = sqrtf(inv * con_o.z) * 1.001f + 0.02f;
    const float ey = sqrtf(inv * con_o.x) * 1.001f + 0.02f;
    if (!(ex < 1e30f) || !(ey < 1e30f)) return 0xFFu;
    const float x_lo = xy.x - ex - tile_x0, x_hi = xy.x + ex - tile_x0;  // reach in tile-local pixels
    const float y_lo = xy.y - ey - tile_y0, y_hi = xy.y + ey - tile_y0;
    uint32_t cols = 0, rows = 0;
    if (x_lo <= 7.0f && x_hi >= 0.0f) cols |= 1u;
    if (x_lo <= 15.0f && x_hi >= 8.0f) cols |= 2u;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (y_lo <= (float)(4 * j + 3) && y_hi >= (float)(4 * j)) rows |= 1u << j;
    uint32_t m = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w)
        if (((cols >> (w & 1)) & 1u) && ((rows >> (w >> 1)) & 1u)) m |= 1u << w;
    return m;
}

// 2^x on the SFU (MUFU.EX2), no range reduction
__device__ __forceinline__ float exp2f_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// Shared-memory loads through a 32-bit shared-window address computed once per kernel: indexing a __shared__
// array makes the compiler rebuild the window base (S2R SR_CgaCtaId + MOV + LEA ...) at every use in the hot loops.
// Packed fp32x2 arithmetic of sm_100 (FFMA2 / FMUL2): IEEE round-to-nearest per lane, one issue slot for two lanes.
__device__ __forceinline__ float2 ffma2_bcast(float w, float2 b, float2 c) {
    unsigned long long rw, rb, rc, rd;
    asm("mov.b64 %0, {%1, %1};" : "=l"(rw) : "f"(w));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(rw), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ float2 fmul2_bcast(float w, float2 b) {
    unsigned long long rw, rb, rd;
    asm("mov.b64 %0, {%1, %1};" : "=l"(rw) : "f"(w));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(rw), "l"(rb));
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
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long ra, rb, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float2 lds_f2(uint32_t a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}

// Cooperative copy between `count` consecutive AoS records of `rec` floats in global memory and shared-memory rows of
// pitch rec+1 words (conflict-free per-thread walks): coalesced 16-byte global accesses.  REC > 0 fixes the record
// length at compile time (48 = SH degree 3): the row/column split of the flat index becomes a multiply-shift.
template <int REC>
__device__ __forceinline__ void stage_records_in(const float* __restrict__ src, float* s_rows, int count, int rec_rt,
                                                 int tid, int nthr) {
    const int rec = REC > 0 ? REC : rec_rt, pitch = rec + 1, total = count * rec;
    const bool vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    // Six 16-byte loads of a thread are in flight before the first of their words is stored: one iteration at a time
    // (load, four dependent stores, next load) this copy was twelve serial memory round trips -- a fifth of the stall
    // samples of preprocess_kernel at cloth size (ncu source page: 206 of 923 on the STS).
    constexpr int U = 6;
    for (int base = tid * 4; base < total; base += nthr * 4 * U) {
        float v[U][4];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * nthr * 4;
            v[u][0] = v[u][1] = v[u][2] = v[u][3] = 0.f;
            if (i < total) {
                if (vec && i + 3 < total) {
                    const float4 q = *reinterpret_cast<const float4*>(src + i);
                    v[u][0] = q.x; v[u][1] = q.y; v[u][2] = q.z; v[u][3] = q.w;
                } else {
                    for (int j = 0; j < 4 && i + j < total; ++j) v[u][j] = src[i + j];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * nthr * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = i + j;
                if (f < total) s_rows[(f / rec) * pitch + (f % rec)] = v[u][j];
            }
        }
    }
}
template <int REC>
__device__ __forceinline__ void stage_records_out(float* __restrict__ dst, const float* s_rows, int count, int rec_rt,
                                                  int tid, int nthr) {
    const int rec = REC > 0 ? REC : rec_rt, pitch = rec + 1, total = count * rec;
    const bool vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    for (int i = tid * 4; i < total; i += nthr * 4) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int f = i + j;
            v[j] = f < total ? s_rows[(f / rec) * pitch + (f % rec)] : 0.f;
        }
        if (vec && i + 3 < total) *reinterpret_cast<float4*>(dst + i) = make_float4(v[0], v[1], v[2], v[3]);
        else
            for (int j = 0; j < 4 && i + j < total; ++j) dst[i + j] = v[j];
    }
}

}  // namespace csb
