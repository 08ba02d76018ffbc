// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Thin extern "C" shim around the *unmodified* reference CUDA sources, compiled
// where they lie under /root/reference (see oracle/Makefile; outputs go to
// oracle/_ref/ only).  It lets tests/ and bench.py's baseline legs drive the
// reference rasterizer and simple-knn with raw device pointers through ctypes:
//
//   CudaRasterizer::Rasterizer::{forward,backward,markVisible}
//        RAST/cuda_rasterizer/rasterizer.h:20-86, rasterizer_impl.cu:141-437
//   SimpleKNN::knn            KNN/simple_knn.h:15-19, simple_knn.cu:185-221
//
// The shim owns the three growable byte buffers that rasterize_points.cu:27-33
// obtains from torch (geometry / binning / image), so no torch is involved.
// It also exposes the carved sub-arrays (GeometryState / BinningState /
// ImageState, rasterizer_impl.h:21-73) so that tests can compare intermediate
// results (means2D, conics, keys, point list, ranges, ...) bit for bit.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"

namespace {
struct Buf {
    char* ptr = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
            cap = want;
        }
        return ptr;
    }
};
struct RefCtx {
    Buf geom, binning, img;
    int P = 0, W = 0, H = 0, R = 0;
};
}  // namespace

extern "C" {

void* ref_ctx_create() { return new RefCtx(); }

void ref_ctx_destroy(void* h) {
    RefCtx* c = (RefCtx*)h;
    if (!c) return;
    if (c->geom.ptr) cudaFree(c->geom.ptr);
    if (c->binning.ptr) cudaFree(c->binning.ptr);
    if (c->img.ptr) cudaFree(c->img.ptr);
    delete c;
}

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-117) minus torch.
// Null pointers stand for the "empty tensor" optional inputs.  Returns
// num_rendered, or a negative cudaError_t.
int ref_rasterize_forward(void* h, int P, int D, int M, const float* background, int W, int H,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                          const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                          int prefiltered, float* out_color, float* out_depth, int* radii) {
    RefCtx* c = (RefCtx*)h;
    c->P = P; c->W = W; c->H = H;
    std::function<char*(size_t)> g = [c](size_t n) { return c->geom.get(n); };
    std::function<char*(size_t)> b = [c](size_t n) { return c->binning.get(n); };
    std::function<char*(size_t)> i = [c](size_t n) { return c->img.get(n); };
    int R = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, background, W, H, means3D, shs,
                                                colors_precomp, opacities, scales, scale_modifier,
                                                rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
                                                tan_fovx, tan_fovy, prefiltered != 0, out_color,
                                                out_depth, radii, false);
    c->R = R;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return -(int)e;
    return R;
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:119-198) minus
// torch; the caller provides zero-initialised gradient arrays.
int ref_rasterize_backward(void* h, int P, int D, int M, int R, const float* background, int W, int H,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* scales, float scale_modifier, const float* rotations,
                           const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                           const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                           float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                           float* dL_dscale, float* dL_drot) {
    RefCtx* c = (RefCtx*)h;
    CudaRasterizer::Rasterizer::backward(P, D, M, R, background, W, H, means3D, shs, colors_precomp, scales,
                                         scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
                                         campos, tan_fovx, tan_fovy, radii, c->geom.ptr, c->binning.ptr,
                                         c->img.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                                         dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, false);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

int ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, unsigned char* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, (bool*)present);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

// Device pointers of the carved sub-arrays of the last forward call.
// names: depths clamped means2D cov3D conic_opacity rgb tiles_touched point_offsets
//        accum_alpha n_contrib ranges point_list point_list_unsorted keys keys_unsorted
void* ref_state_ptr(void* h, const char* name) {
    RefCtx* c = (RefCtx*)h;
    char* gp = c->geom.ptr;
    CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(gp, c->P);
    char* ip = c->img.ptr;
    CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(ip, (size_t)c->W * c->H);
    char* bp = c->binning.ptr;
    CudaRasterizer::BinningState bs = CudaRasterizer::BinningState::fromChunk(bp, c->R);
    if (!strcmp(name, "depths")) return gs.depths;
    if (!strcmp(name, "clamped")) return gs.clamped;
    if (!strcmp(name, "means2D")) return gs.means2D;
    if (!strcmp(name, "cov3D")) return gs.cov3D;
    if (!strcmp(name, "conic_opacity")) return gs.conic_opacity;
    if (!strcmp(name, "rgb")) return gs.rgb;
    if (!strcmp(name, "tiles_touched")) return gs.tiles_touched;
    if (!strcmp(name, "point_offsets")) return gs.point_offsets;
    if (!strcmp(name, "accum_alpha")) return is.accum_alpha;
    if (!strcmp(name, "n_contrib")) return is.n_contrib;
    if (!strcmp(name, "ranges")) return is.ranges;
    if (!strcmp(name, "point_list")) return bs.point_list;
    if (!strcmp(name, "point_list_unsorted")) return bs.point_list_unsorted;
    if (!strcmp(name, "keys")) return bs.point_list_keys;
    if (!strcmp(name, "keys_unsorted")) return bs.point_list_keys_unsorted;
    return nullptr;
}

// Device-to-device copy of one carved sub-array into a caller buffer (synchronous).
int ref_copy_state(void* h, const char* name, void* dst, size_t nbytes) {
    void* src = ref_state_ptr(h, name);
    if (!src) return -1;
    cudaError_t e = cudaMemcpy(dst, src, nbytes, cudaMemcpyDeviceToDevice);
    return e == cudaSuccess ? 0 : -(int)e;
}

// simple-knn: KNN/spatial.cu:15-26 minus torch.
int ref_dist_cuda2(int P, float* points, float* mean_dists) {
    SimpleKNN::knn(P, (float3*)points, mean_dists);
    cudaError_t e = cudaDeviceSynchronize();
    return e == cudaSuccess ? 0 : -(int)e;
}

}  // extern "C"
