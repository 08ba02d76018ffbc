// Shared device code of the mesh -> Gaussian transform kernels (mesh_rotation.cu, mesh_transform.cu): the closed-form
// Kabsch rotation of a face as a quaternion, quaternion normalisation / composition, all templated on a scalar type so that
// the same code runs on floats (forward) and on (value, tangent) pairs (forward-mode derivative for the backward).
// Reference: scene_reconstruction/gaussian_mesh.py:171-188 (roma.rigid_points_registration, rotmat_to_unitquat,
// quat_composition).
#pragma once
#include "common.cuh"

namespace csb {
namespace rot {

struct Dual {
    float v, d;
};
__device__ __forceinline__ Dual operator+(Dual a, Dual b) { return {a.v + b.v, a.d + b.d}; }
__device__ __forceinline__ Dual operator-(Dual a, Dual b) { return {a.v - b.v, a.d - b.d}; }
__device__ __forceinline__ Dual operator-(Dual a) { return {-a.v, -a.d}; }
__device__ __forceinline__ Dual operator*(Dual a, Dual b) { return {a.v * b.v, a.d * b.v + a.v * b.d}; }
__device__ __forceinline__ Dual operator/(Dual a, Dual b) {
    const float q = a.v / b.v;
    return {q, (a.d - q * b.d) / b.v};
}
__device__ __forceinline__ Dual operator*(float a, Dual b) { return {a * b.v, a * b.d}; }
__device__ __forceinline__ Dual operator*(Dual a, float b) { return {a.v * b, a.d * b}; }
__device__ __forceinline__ Dual operator+(Dual a, float b) { return {a.v + b, a.d}; }
__device__ __forceinline__ Dual operator+(float a, Dual b) { return {a + b.v, b.d}; }
__device__ __forceinline__ Dual operator-(float a, Dual b) { return {a - b.v, -b.d}; }
__device__ __forceinline__ Dual operator-(Dual a, float b) { return {a.v - b, a.d}; }
__device__ __forceinline__ float val(float a) { return a; }
__device__ __forceinline__ float val(Dual a) { return a.v; }
__device__ __forceinline__ float tsqrt(float a) { return sqrtf(a); }
__device__ __forceinline__ Dual tsqrt(Dual a) {
    const float s = sqrtf(a.v);
    return {s, s > 0.f ? 0.5f * a.d / s : 0.f};
}
// max(a, floor): below the floor the value is the constant
__device__ __forceinline__ float tfloor(float a, float f) { return fmaxf(a, f); }
__device__ __forceinline__ Dual tfloor(Dual a, float f) { return a.v > f ? a : Dual{f, 0.f}; }
template <typename T> __device__ __forceinline__ T lift(float a);
template <> __device__ __forceinline__ float lift<float>(float a) { return a; }
template <> __device__ __forceinline__ Dual lift<Dual>(float a) { return {a, 0.f}; }

template <typename T> __device__ __forceinline__ void cross3(const T* a, const T* b, T* c) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
template <typename T> __device__ __forceinline__ T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> __device__ __forceinline__ void normalize3(T* a) {
    const T n = tfloor(tsqrt(dot3(a, a)), 1e-30f);
    a[0] = a[0] / n; a[1] = a[1] / n; a[2] = a[2] / n;
}

// normalize(q): q / max(|q|, 1e-12)  (torch.nn.functional.normalize); returns the floored norm
template <typename T> __device__ __forceinline__ T quat_normalize(const T (&q)[4], T (&p)[4]) {
    const T qn = tfloor(tsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
    p[0] = q[0] / qn; p[1] = q[1] / qn; p[2] = q[2] / qn; p[3] = q[3] / qn;
    return qn;
}

// The rotation of a FACE: special-Procrustes (Kabsch) rotation from the rest triangle x to the deformed triangle y in the
// closed form of the header comment of mesh_rotation.cu, as an XYZW unit quaternion (roma.rotmat_to_unitquat's branches).
// It depends on the face only: all Gaussians anchored on a face share it.
template <typename T>
__device__ void face_quat(const float (&xf)[3][3], const T (&y)[3][3], T (&u)[4]) {
    T x[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c) x[i][c] = lift<T>(xf[i][c]);
    // centred point sets
    T xh[3][3], yh[3][3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const T xm = (x[0][c] + x[1][c] + x[2][c]) * (1.0f / 3.0f);
        const T ym = (y[0][c] + y[1][c] + y[2][c]) * (1.0f / 3.0f);
#pragma unroll
        for (int i = 0; i < 3; ++i) { xh[i][c] = x[i][c] - xm; yh[i][c] = y[i][c] - ym; }
    }
    // plane frames
    T e1[3], e2[3], nx[3], f1[3], f2[3], ny[3], t[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { e1[c] = x[1][c] - x[0][c]; t[c] = x[2][c] - x[0][c]; }
    cross3(e1, t, nx); normalize3(nx); normalize3(e1); cross3(nx, e1, e2);
#pragma unroll
    for (int c = 0; c < 3; ++c) { f1[c] = y[1][c] - y[0][c]; t[c] = y[2][c] - y[0][c]; }
    cross3(f1, t, ny); normalize3(ny); normalize3(f1); cross3(ny, f1, f2);
    // m = [f1 f2]^T M [e1 e2],  M = sum_i yh_i xh_i^T
    T m00 = lift<T>(0.f), m01 = m00, m10 = m00, m11 = m00;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const T a0 = dot3(f1, yh[i]), a1 = dot3(f2, yh[i]), b0 = dot3(e1, xh[i]), b1 = dot3(e2, xh[i]);
        m00 = m00 + a0 * b0; m01 = m01 + a0 * b1; m10 = m10 + a1 * b0; m11 = m11 + a1 * b1;
    }
    const float s = (val(m00) * val(m11) - val(m01) * val(m10)) >= 0.f ? 1.f : -1.f;
    T r00 = m00 + s * m11, r01 = m01 - s * m10, r10 = m10 - s * m01, r11 = m11 + s * m00;
    const T h = tfloor(tsqrt(r00 * r00 + r10 * r10), 1e-30f);
    r00 = r00 / h; r01 = r01 / h; r10 = r10 / h; r11 = r11 / h;
    T R[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b)
            R[a][b] = f1[a] * (r00 * e1[b] + r01 * e2[b]) + f2[a] * (r10 * e1[b] + r11 * e2[b]) + s * (ny[a] * nx[b]);
    // rotmat_to_unitquat (XYZW), the branch picks the largest of (R00, R11, R22, trace)
    const T tr = R[0][0] + R[1][1] + R[2][2];
    int choice = 3;
    float best = val(tr);
    if (val(R[2][2]) >= best) { best = val(R[2][2]); choice = 2; }
    if (val(R[1][1]) >= best) { best = val(R[1][1]); choice = 1; }
    if (val(R[0][0]) >= best) { best = val(R[0][0]); choice = 0; }
    if (choice == 3) {
        u[0] = R[2][1] - R[1][2]; u[1] = R[0][2] - R[2][0]; u[2] = R[1][0] - R[0][1]; u[3] = 1.0f + tr;
    } else {
        const int i = choice, j = (i + 1) % 3, k = (j + 1) % 3;
        u[i] = 1.0f - tr + 2.0f * R[i][i];
        u[j] = R[j][i] + R[i][j];
        u[k] = R[k][i] + R[i][k];
        u[3] = R[k][j] - R[j][k];
    }
    const T un = tfloor(tsqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2] + u[3] * u[3]), 1e-30f);
    u[0] = u[0] / un; u[1] = u[1] / un; u[2] = u[2] / un; u[3] = u[3] / un;
}

// quat_product(p, u), XYZW (roma.quat_composition([p, u]))
template <typename T> __device__ __forceinline__ void quat_compose(const T (&p)[4], const T (&u)[4], T (&out)[4]) {
    out[0] = p[3] * u[0] + u[3] * p[0] + (p[1] * u[2] - p[2] * u[1]);
    out[1] = p[3] * u[1] + u[3] * p[1] + (p[2] * u[0] - p[0] * u[2]);
    out[2] = p[3] * u[2] + u[3] * p[2] + (p[0] * u[1] - p[1] * u[0]);
    out[3] = p[3] * u[3] - (p[0] * u[0] + p[1] * u[1] + p[2] * u[2]);
}

// x: rest triangle [3][3], y: deformed triangle [3][3], q: raw quaternion parameter [4] -> out [4]
template <typename T>
__device__ void rotation_chain(const float (&xf)[3][3], const T (&y)[3][3], const T (&q)[4], bool has_face, T (&out)[4]) {
    T p[4];
    quat_normalize(q, p);
    if (!has_face) {
        out[0] = p[0]; out[1] = p[1]; out[2] = p[2]; out[3] = p[3];
        return;
    }
    T u[4];
    face_quat(xf, y, u);
    quat_compose(p, u, out);
}

struct FaceRef {
    long long v[3];
    bool ok;
};
__device__ __forceinline__ FaceRef face_of(int p, int Nv, int F, const long long* __restrict__ face,
                                           const long long* __restrict__ face_ids) {
    FaceRef r;
    const long long f = face_ids[p];
    r.ok = (unsigned long long)f < (unsigned long long)F;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        r.v[k] = r.ok ? face[(size_t)k * F + f] : 0;
        if ((unsigned long long)r.v[k] >= (unsigned long long)Nv) r.ok = false;
    }
    return r;
}

}  // namespace rot
}  // namespace csb
