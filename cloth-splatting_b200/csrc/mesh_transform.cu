// Fused mesh -> Gaussian transform: MultiGaussianMesh.get_xyz AND get_rotation of
// scene_reconstruction/gaussian_mesh.py:151-188 for all views of a refinement step in ONE forward launch, with the
// residual simulator's output (meshnet/meshnet_network.py:361-373: vertices = mesh_prediction + residual) consumed in
// place: the kernels read `vertices[v] + residual[v]`, the gradient they return is that of both.
//
//   forward   one thread per (view, Gaussian): the three vertices of its face are gathered once and feed both the
//             barycentric position and the face's rest -> deformed rotation (closed-form Kabsch, mesh_math.cuh),
//             composed with the normalised quaternion parameter.
//   backward  A  one thread per Gaussian, views in a fixed order (deterministic sums over views for dL/dbary and
//                dL/drotation): position gradient to the vertices and the barycentric coordinates, quaternion-product
//                gradient to the parameter (hand-written reverse mode incl. the normalisation) and to the FACE's
//                relative quaternion u, accumulated per (view, face) -- a face carries ~14 Gaussians that share u;
//             B  one thread per (view, face, vertex coordinate): forward-mode derivative of u (9 passes per face instead
//                of 13 per Gaussian: 6.5 k chains instead of 131 k at C2) contracted with the face's dL/du.
// Replaces, per view: 4 torch kernels + roma's batched 3x3 SVD forward, and their autograd (index_put atomics, SVD
// backward); and the two separate kernel pairs of mesh_xyz.cu / mesh_rotation.cu, which stay as the single-purpose
// entry points of the accessors.
#include "mesh_math.cuh"

namespace csb {
namespace xform {
using namespace csb::rot;

struct Args {
    int V, P, Nv, F;
    const float* V0;         // [Nv,3] rest vertices
    const float* base;       // [V,Nv,3]
    const float* residual;   // [V,Nv,3] or null
    const long long* face;   // [3,F]
    const long long* face_ids;  // [P]
    const float* bary;       // [P,3]
    const float* rotation;   // [P,4]
};

__device__ __forceinline__ void load_triangles(const Args& a, int v, const FaceRef& fr, float (&x)[3][3], float (&y)[3][3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            x[k][c] = 0.f;
            y[k][c] = 0.f;
            if (fr.ok) {
                const size_t i = 3 * (size_t)fr.v[k] + c, iv = (size_t)v * a.Nv * 3 + i;
                x[k][c] = a.V0[i];
                y[k][c] = a.residual ? a.base[iv] + a.residual[iv] : a.base[iv];
            }
        }
}

__global__ void __launch_bounds__(128)
mesh_transform_forward_kernel(const Args a, float* __restrict__ xyz, float* __restrict__ rot_out) {
    const int p = blockIdx.x * 128 + threadIdx.x, v = blockIdx.y;
    if (p >= a.P) return;
    const FaceRef fr = face_of(p, a.Nv, a.F, a.face, a.face_ids);
    float x[3][3], y[3][3];
    load_triangles(a, v, fr, x, y);
    const size_t o = (size_t)v * a.P + p;
    if (xyz) {
        // the arithmetic of mesh_xyz_forward_kernel
        float r[3] = {0.f, 0.f, 0.f};
        if ((unsigned long long)a.face_ids[p] < (unsigned long long)a.F) {
            const float b0 = a.bary[3 * p], b1 = a.bary[3 * p + 1], b2 = a.bary[3 * p + 2];
            const float s = b0 + b1 + b2;
            const float nb[3] = {b0 / s, b1 / s, b2 / s};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const long long vid = a.face[(size_t)k * a.F + a.face_ids[p]];
                if ((unsigned long long)vid >= (unsigned long long)a.Nv) continue;
                const size_t iv = (size_t)v * a.Nv * 3 + 3 * (size_t)vid;
                const float y0 = a.residual ? a.base[iv] + a.residual[iv] : a.base[iv];
                const float y1 = a.residual ? a.base[iv + 1] + a.residual[iv + 1] : a.base[iv + 1];
                const float y2 = a.residual ? a.base[iv + 2] + a.residual[iv + 2] : a.base[iv + 2];
                r[0] = fmaf(nb[k], y0, r[0]);
                r[1] = fmaf(nb[k], y1, r[1]);
                r[2] = fmaf(nb[k], y2, r[2]);
            }
        }
        xyz[3 * o] = r[0];
        xyz[3 * o + 1] = r[1];
        xyz[3 * o + 2] = r[2];
    }
    if (rot_out) {
        const float4 qv = reinterpret_cast<const float4*>(a.rotation)[p];
        const float q[4] = {qv.x, qv.y, qv.z, qv.w};
        float out[4];
        rotation_chain<float>(x, y, q, fr.ok, out);
        reinterpret_cast<float4*>(rot_out)[o] = make_float4(out[0], out[1], out[2], out[3]);
    }
}

__global__ void __launch_bounds__(128)
mesh_transform_backward_gaussians_kernel(const Args a, const float* __restrict__ dxyz, const float* __restrict__ drot,
                                         float* __restrict__ dV, float* __restrict__ dbary, float* __restrict__ drotation,
                                         float* __restrict__ du_face) {
    const int p = blockIdx.x * 128 + threadIdx.x;
    if (p >= a.P) return;
    const FaceRef fr = face_of(p, a.Nv, a.F, a.face, a.face_ids);
    const long long f = a.face_ids[p];
    const bool has_f = (unsigned long long)f < (unsigned long long)a.F;
    float nb[3] = {0.f, 0.f, 0.f}, s = 1.f;
    if (has_f) {
        const float b0 = a.bary[3 * p], b1 = a.bary[3 * p + 1], b2 = a.bary[3 * p + 2];
        s = b0 + b1 + b2;
        nb[0] = b0 / s; nb[1] = b1 / s; nb[2] = b2 / s;
    }
    const float4 qv = reinterpret_cast<const float4*>(a.rotation)[p];
    const float q[4] = {qv.x, qv.y, qv.z, qv.w};
    float pn[4];
    const float qn = quat_normalize<float>(q, pn);
    float db[3] = {0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
    for (int v = 0; v < a.V; ++v) {
        const size_t o = (size_t)v * a.P + p;
        float x[3][3], y[3][3];
        if (dxyz && has_f) {
            // mesh_xyz_backward_kernel's arithmetic (it skips a vertex that is out of range, not the whole face)
            const float g[3] = {dxyz[3 * o], dxyz[3 * o + 1], dxyz[3 * o + 2]};
            float dnb[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const long long vid = a.face[(size_t)k * a.F + f];
                if ((unsigned long long)vid >= (unsigned long long)a.Nv) continue;
                const size_t iv = (size_t)v * a.Nv * 3 + 3 * (size_t)vid;
                const float y0 = a.residual ? a.base[iv] + a.residual[iv] : a.base[iv];
                const float y1 = a.residual ? a.base[iv + 1] + a.residual[iv + 1] : a.base[iv + 1];
                const float y2 = a.residual ? a.base[iv + 2] + a.residual[iv + 2] : a.base[iv + 2];
                dnb[k] = g[0] * y0 + g[1] * y1 + g[2] * y2;
                if (dV) {
                    atomicAdd(dV + iv, nb[k] * g[0]);
                    atomicAdd(dV + iv + 1, nb[k] * g[1]);
                    atomicAdd(dV + iv + 2, nb[k] * g[2]);
                }
            }
            const float dot = nb[0] * dnb[0] + nb[1] * dnb[1] + nb[2] * dnb[2];
#pragma unroll
            for (int k = 0; k < 3; ++k) db[k] += (dnb[k] - dot) / s;
        }
        if (drot) {
            const float4 gv = reinterpret_cast<const float4*>(drot)[o];
            const float g[4] = {gv.x, gv.y, gv.z, gv.w};
            if (!fr.ok) {  // out = normalize(q)
#pragma unroll
                for (int c = 0; c < 4; ++c) dp[c] += g[c];
                continue;
            }
            load_triangles(a, v, fr, x, y);
            float u[4];
            face_quat<float>(x, y, u);
            // out = quat_product(pn, u): the transposes of its two linear maps
            dp[0] += g[0] * u[3] - g[1] * u[2] + g[2] * u[1] - g[3] * u[0];
            dp[1] += g[0] * u[2] + g[1] * u[3] - g[2] * u[0] - g[3] * u[1];
            dp[2] += -g[0] * u[1] + g[1] * u[0] + g[2] * u[3] - g[3] * u[2];
            dp[3] += g[0] * u[0] + g[1] * u[1] + g[2] * u[2] + g[3] * u[3];
            if (du_face) {
                float* d = du_face + 4 * ((size_t)v * a.F + f);
                atomicAdd(d + 0, g[0] * pn[3] + g[1] * pn[2] - g[2] * pn[1] - g[3] * pn[0]);
                atomicAdd(d + 1, -g[0] * pn[2] + g[1] * pn[3] + g[2] * pn[0] - g[3] * pn[1]);
                atomicAdd(d + 2, g[0] * pn[1] - g[1] * pn[0] + g[2] * pn[3] - g[3] * pn[2]);
                atomicAdd(d + 3, g[0] * pn[0] + g[1] * pn[1] + g[2] * pn[2] + g[3] * pn[3]);
            }
        }
    }
    if (dbary && dxyz) {
        dbary[3 * p] = db[0];
        dbary[3 * p + 1] = db[1];
        dbary[3 * p + 2] = db[2];
    }
    if (drotation && drot) {
        // p = q / max(|q|, 1e-12): below the floor the denominator is a constant
        const float nrm = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
        float dq[4];
        if (nrm > 1e-12f) {
            const float dot = pn[0] * dp[0] + pn[1] * dp[1] + pn[2] * dp[2] + pn[3] * dp[3];
#pragma unroll
            for (int c = 0; c < 4; ++c) dq[c] = (dp[c] - pn[c] * dot) / qn;
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) dq[c] = dp[c] / qn;
        }
        reinterpret_cast<float4*>(drotation)[p] = make_float4(dq[0], dq[1], dq[2], dq[3]);
    }
}

// 16-lane groups: group = (view, face), lanes 0..8 = the vertex coordinate whose tangent the pass carries
__global__ void __launch_bounds__(128)
mesh_transform_backward_faces_kernel(const Args a, const float* __restrict__ du_face, float* __restrict__ dV) {
    const long long t = (long long)blockIdx.x * 128 + threadIdx.x;
    const long long grp = t >> 4;
    const int in = (int)(t & 15);
    if (grp >= (long long)a.V * a.F || in >= 9) return;
    const int v = (int)(grp / a.F), f = (int)(grp % a.F);
    const float4 d = reinterpret_cast<const float4*>(du_face)[grp];
    if (d.x == 0.f && d.y == 0.f && d.z == 0.f && d.w == 0.f) return;  // no Gaussian on this face (or no gradient)
    FaceRef fr;
    fr.ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        fr.v[k] = a.face[(size_t)k * a.F + f];
        if ((unsigned long long)fr.v[k] >= (unsigned long long)a.Nv) fr.ok = false;
    }
    if (!fr.ok) return;
    float x[3][3], yf[3][3];
    load_triangles(a, v, fr, x, yf);
    Dual y[3][3], u[4];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) y[k][c] = {yf[k][c], (in == 3 * k + c) ? 1.f : 0.f};
    face_quat<Dual>(x, y, u);
    const float grad = d.x * u[0].d + d.y * u[1].d + d.z * u[2].d + d.w * u[3].d;
    atomicAdd(dV + (size_t)v * a.Nv * 3 + 3 * (size_t)fr.v[in / 3] + (in % 3), grad);
}

}  // namespace xform
}  // namespace csb

using namespace csb;

extern "C" {

size_t csb_mesh_transform_workspace_bytes(int num_views, int F) {
    return (size_t)(num_views > 0 ? num_views : 0) * (size_t)(F > 0 ? F : 0) * 4 * sizeof(float);
}

int csb_mesh_transform_forward(int num_views, int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                               const float* residual, const int64_t* face, const int64_t* face_ids,
                               const float* face_bary, const float* rotation, float* xyz, float* rot, void* stream_) {
    if (num_views < 0 || P < 0 || Nv < 0 || F < 0 || num_views > 65535) return CSB_ERR_BAD_ARG;
    if (P == 0 || num_views == 0) return 0;
    if (!rest_vertices || !vertices || !face || !face_ids || (!xyz && !rot)) return CSB_ERR_BAD_ARG;
    if ((xyz && !face_bary) || (rot && !rotation)) return CSB_ERR_BAD_ARG;
    xform::Args a{num_views, P, Nv, F, rest_vertices, vertices, residual, (const long long*)face,
                  (const long long*)face_ids, face_bary, rotation};
    cudaStream_t stream = (cudaStream_t)stream_;
    CSB_LAUNCH("mesh_transform_forward", stream,
               xform::mesh_transform_forward_kernel<<<dim3((P + 127) / 128, num_views), 128, 0, stream>>>(a, xyz, rot));
    return 0;
}

int csb_mesh_transform_backward(int num_views, int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                                const float* residual, const int64_t* face, const int64_t* face_ids,
                                const float* face_bary, const float* rotation, const float* dL_dxyz,
                                const float* dL_drot, float* dL_dvertices, float* dL_dface_bary, float* dL_drotation,
                                void* workspace, size_t workspace_bytes, void* stream_) {
    if (num_views < 0 || P < 0 || Nv < 0 || F < 0) return CSB_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (dL_dvertices && Nv > 0 && num_views > 0)
        CSB_CUDA_TRY(cudaMemsetAsync(dL_dvertices, 0, (size_t)num_views * Nv * 3 * sizeof(float), stream));
    if (P == 0 || num_views == 0) return 0;
    if (!rest_vertices || !vertices || !face || !face_ids || !face_bary || !rotation) return CSB_ERR_BAD_ARG;
    if (!dL_dxyz && !dL_drot) return CSB_ERR_BAD_ARG;
    const bool faces = dL_drot && dL_dvertices && F > 0;
    const size_t need = csb_mesh_transform_workspace_bytes(num_views, F);
    if (faces) {
        if (!workspace || workspace_bytes < need) return CSB_ERR_WORKSPACE;
        CSB_CUDA_TRY(cudaMemsetAsync(workspace, 0, need, stream));
    }
    xform::Args a{num_views, P, Nv, F, rest_vertices, vertices, residual, (const long long*)face,
                  (const long long*)face_ids, face_bary, rotation};
    CSB_LAUNCH("mesh_transform_backward", stream,
               xform::mesh_transform_backward_gaussians_kernel<<<(P + 127) / 128, 128, 0, stream>>>(
                   a, dL_dxyz, dL_drot, dL_dvertices, dL_dface_bary, dL_drotation, faces ? (float*)workspace : nullptr));
    if (faces) {
        const long long threads = (long long)num_views * F * 16;
        CSB_LAUNCH("mesh_transform_backward_faces", stream,
                   xform::mesh_transform_backward_faces_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, stream>>>(
                       a, (const float*)workspace, dL_dvertices));
    }
    return 0;
}

}  // extern "C"
