// Mesh-anchored Gaussian placement: barycentric transform forward + backward.
//
// Replaces MultiGaussianMesh.get_xyz (scene_reconstruction/gaussian_mesh.py:151-169): four torch kernels
// (index, sum, div, bmm) and, in backward, index_put atomics + bmm transposes.  One thread per Gaussian:
//   vid_k = face[k, face_ids[p]];  nb = bary / (b0 + b1 + b2);  pos = nb0 V[vid0] + nb1 V[vid1] + nb2 V[vid2]
// backward:  dV[vid_k] += nb_k dpos;  dnb_k = <dpos, V[vid_k]>;  dbary_i = (dnb_i - sum_k nb_k dnb_k) / s
#include "common.cuh"

namespace csb {

__global__ void __launch_bounds__(256)
mesh_xyz_forward_kernel(int P, int Nv, int F, const float* __restrict__ V, const long long* __restrict__ face,
                        const long long* __restrict__ face_ids, const float* __restrict__ bary, float* __restrict__ out) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p >= P) return;
    const long long f = face_ids[p];
    float r[3] = {0.f, 0.f, 0.f};
    if ((unsigned long long)f < (unsigned long long)F) {
        const float b0 = bary[3 * p], b1 = bary[3 * p + 1], b2 = bary[3 * p + 2];
        const float s = b0 + b1 + b2;
        const float nb[3] = {b0 / s, b1 / s, b2 / s};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const long long v = face[(size_t)k * F + f];
            if ((unsigned long long)v >= (unsigned long long)Nv) continue;
            r[0] = fmaf(nb[k], V[3 * v], r[0]);
            r[1] = fmaf(nb[k], V[3 * v + 1], r[1]);
            r[2] = fmaf(nb[k], V[3 * v + 2], r[2]);
        }
    }
    out[3 * p] = r[0];
    out[3 * p + 1] = r[1];
    out[3 * p + 2] = r[2];
}

__global__ void __launch_bounds__(256)
mesh_xyz_backward_kernel(int P, int Nv, int F, const float* __restrict__ V, const long long* __restrict__ face,
                         const long long* __restrict__ face_ids, const float* __restrict__ bary,
                         const float* __restrict__ dpos, float* __restrict__ dV, float* __restrict__ dbary) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p >= P) return;
    const long long f = face_ids[p];
    float db[3] = {0.f, 0.f, 0.f};
    if ((unsigned long long)f < (unsigned long long)F) {
        const float b0 = bary[3 * p], b1 = bary[3 * p + 1], b2 = bary[3 * p + 2];
        const float s = b0 + b1 + b2;
        const float nb[3] = {b0 / s, b1 / s, b2 / s};
        const float g[3] = {dpos[3 * p], dpos[3 * p + 1], dpos[3 * p + 2]};
        float dnb[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const long long v = face[(size_t)k * F + f];
            if ((unsigned long long)v >= (unsigned long long)Nv) continue;
            dnb[k] = g[0] * V[3 * v] + g[1] * V[3 * v + 1] + g[2] * V[3 * v + 2];
            if (dV) {
                atomicAdd(dV + 3 * v, nb[k] * g[0]);
                atomicAdd(dV + 3 * v + 1, nb[k] * g[1]);
                atomicAdd(dV + 3 * v + 2, nb[k] * g[2]);
            }
        }
        const float dot = nb[0] * dnb[0] + nb[1] * dnb[1] + nb[2] * dnb[2];
#pragma unroll
        for (int k = 0; k < 3; ++k) db[k] = (dnb[k] - dot) / s;
    }
    if (dbary) {
        dbary[3 * p] = db[0];
        dbary[3 * p + 1] = db[1];
        dbary[3 * p + 2] = db[2];
    }
}

}  // namespace csb

using namespace csb;

extern "C" {

int csb_mesh_xyz_forward(int P, int Nv, int F, const float* vertices, const int64_t* face, const int64_t* face_ids,
                         const float* face_bary, float* xyz, void* stream_) {
    if (P < 0 || Nv < 0 || F < 0) return CSB_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!vertices || !face || !face_ids || !face_bary || !xyz) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    CSB_LAUNCH("mesh_xyz_forward", stream, mesh_xyz_forward_kernel<<<(P + 255) / 256, 256, 0, stream>>>(
        P, Nv, F, vertices, (const long long*)face, (const long long*)face_ids, face_bary, xyz));
    return 0;
}

int csb_mesh_xyz_backward(int P, int Nv, int F, const float* vertices, const int64_t* face, const int64_t* face_ids,
                          const float* face_bary, const float* dL_dxyz, float* dL_dvertices, float* dL_dface_bary,
                          void* stream_) {
    if (P < 0 || Nv < 0 || F < 0) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (dL_dvertices && Nv > 0) CSB_CUDA_TRY(cudaMemsetAsync(dL_dvertices, 0, (size_t)Nv * 3 * sizeof(float), stream));
    if (P == 0) return 0;
    if (!vertices || !face || !face_ids || !face_bary || !dL_dxyz) return CSB_ERR_BAD_ARG;
    CSB_LAUNCH("mesh_xyz_backward", stream, mesh_xyz_backward_kernel<<<(P + 255) / 256, 256, 0, stream>>>(
        P, Nv, F, vertices, (const long long*)face, (const long long*)face_ids, face_bary, dL_dxyz, dL_dvertices,
        dL_dface_bary));
    return 0;
}

}  // extern "C"
