// Mesh-anchored Gaussian orientation: MultiGaussianMesh.get_rotation forward + backward
// (scene_reconstruction/gaussian_mesh.py:171-188 of the reference):
//     q     = normalize(_rotation)                                            (torch.nn.functional.normalize)
//     R, _  = roma.rigid_points_registration(V0[vid], V[vid])                 (Kabsch on the face's 3 vertices)
//     q_rel = roma.rotmat_to_unitquat(R)                                      (XYZW)
//     out   = roma.quat_composition([q, q_rel])                               (= quat_product(q, q_rel), XYZW)
// The reference runs a batched 3x3 SVD (and its autograd, the slowest torch op left in render()); roma is a third
// party package that is not vendored, its published algorithms are restated in oracle/mesh_oracle.py.
//
// Here: one thread per Gaussian, no SVD.  The cross-covariance M = sum_i (y_i - ybar)(x_i - xbar)^T of two triangles
// has rank 2 (both point sets are planar), so the special-Procrustes solution U diag(1,1,det(U V^T)) V^T has the
// closed form
//     R = [f1 f2] r2 [e1 e2]^T + det(r2) n_y n_x^T
// with (e1,e2,n_x), (f1,f2,n_y) right-handed orthonormal frames of the two triangle planes, m = [f1 f2]^T M [e1 e2]
// (2x2) and r2 = (m + sign(det m) cof(m)) / |.| its orthogonal polar factor.
// Backward: forward-mode differentiation of the same templated function with a (value, tangent) scalar, one pass
// per input (9 vertex coordinates + 4 quaternion parameters); ~13 x 200 flops per Gaussian, no stored state.
#include "mesh_math.cuh"

namespace csb {
namespace rot {

__global__ void __launch_bounds__(128)
mesh_rotation_forward_kernel(int P, int Nv, int F, const float* __restrict__ V0, const float* __restrict__ V,
                             const long long* __restrict__ face, const long long* __restrict__ face_ids,
                             const float* __restrict__ rot, float* __restrict__ out) {
    const int p = blockIdx.x * 128 + threadIdx.x;
    if (p >= P) return;
    const FaceRef fr = face_of(p, Nv, F, face, face_ids);
    float x[3][3] = {}, y[3][3] = {};
    if (fr.ok)
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int c = 0; c < 3; ++c) { x[k][c] = V0[3 * fr.v[k] + c]; y[k][c] = V[3 * fr.v[k] + c]; }
    const float4 qv = reinterpret_cast<const float4*>(rot)[p];
    const float q[4] = {qv.x, qv.y, qv.z, qv.w};
    float o[4];
    rotation_chain<float>(x, y, q, fr.ok, o);
    reinterpret_cast<float4*>(out)[p] = make_float4(o[0], o[1], o[2], o[3]);
}

// Backward = forward-mode differentiation of the same templated chain, one pass per input (9 vertex coordinates of the
// face, 4 quaternion components).  The 13 passes of a Gaussian run on 13 lanes of a 16-lane group (two Gaussians per
// warp): with one thread per Gaussian doing them one after another the kernel was a single 13 x ~300-instruction
// dependency chain on a third of the SMs (65 us for 10 108 Gaussians; the passes share nothing but their inputs).
__global__ void __launch_bounds__(128)
mesh_rotation_backward_kernel(int P, int Nv, int F, const float* __restrict__ V0, const float* __restrict__ V,
                              const long long* __restrict__ face, const long long* __restrict__ face_ids,
                              const float* __restrict__ rot, const float* __restrict__ dout, float* __restrict__ dV,
                              float* __restrict__ drot) {
    const long long t = (long long)blockIdx.x * 128 + threadIdx.x;
    const int p = (int)(t >> 4), in = (int)(t & 15);
    if (p >= P || in >= 13) return;
    const FaceRef fr = face_of(p, Nv, F, face, face_ids);
    if (in < 9 && (!fr.ok || !dV)) return;
    if (in >= 9 && !drot) return;
    float x[3][3] = {}, yf[3][3] = {};
    if (fr.ok)
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int c = 0; c < 3; ++c) { x[k][c] = V0[3 * fr.v[k] + c]; yf[k][c] = V[3 * fr.v[k] + c]; }
    const float4 qv = reinterpret_cast<const float4*>(rot)[p];
    const float qf[4] = {qv.x, qv.y, qv.z, qv.w};
    const float4 gv = reinterpret_cast<const float4*>(dout)[p];
    const float g[4] = {gv.x, gv.y, gv.z, gv.w};
    Dual y[3][3], q[4], o[4];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) y[k][c] = {yf[k][c], (in == 3 * k + c) ? 1.f : 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c) q[c] = {qf[c], (in == 9 + c) ? 1.f : 0.f};
    rotation_chain<Dual>(x, y, q, fr.ok, o);
    const float grad = g[0] * o[0].d + g[1] * o[1].d + g[2] * o[2].d + g[3] * o[3].d;
    if (in < 9) atomicAdd(dV + 3 * fr.v[in / 3] + (in % 3), grad);
    else drot[4 * (size_t)p + (in - 9)] = grad;
}

}  // namespace rot
}  // namespace csb

using namespace csb;

extern "C" {

int csb_mesh_rotation_forward(int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                              const int64_t* face, const int64_t* face_ids, const float* rotation, float* out,
                              void* stream_) {
    if (P < 0 || Nv < 0 || F < 0) return CSB_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!rest_vertices || !vertices || !face || !face_ids || !rotation || !out) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    CSB_LAUNCH("mesh_rotation_forward", stream, rot::mesh_rotation_forward_kernel<<<(P + 127) / 128, 128, 0, stream>>>(
        P, Nv, F, rest_vertices, vertices, (const long long*)face, (const long long*)face_ids, rotation, out));
    return 0;
}

int csb_mesh_rotation_backward(int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                               const int64_t* face, const int64_t* face_ids, const float* rotation,
                               const float* dL_dout, float* dL_dvertices, float* dL_drotation, void* stream_) {
    if (P < 0 || Nv < 0 || F < 0) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (dL_dvertices && Nv > 0) CSB_CUDA_TRY(cudaMemsetAsync(dL_dvertices, 0, (size_t)Nv * 3 * sizeof(float), stream));
    if (P == 0) return 0;
    if (!rest_vertices || !vertices || !face || !face_ids || !rotation || !dL_dout) return CSB_ERR_BAD_ARG;
    CSB_LAUNCH("mesh_rotation_backward", stream, rot::mesh_rotation_backward_kernel<<<(unsigned)(((long long)P * 16 + 127) / 128), 128, 0, stream>>>(
        P, Nv, F, rest_vertices, vertices, (const long long*)face, (const long long*)face_ids, rotation, dL_dout,
        dL_dvertices, dL_drotation));
    return 0;
}

}  // extern "C"
