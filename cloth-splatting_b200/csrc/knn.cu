// distCUDA2: mean squared distance to the 3 nearest other points, on a Morton-ordered uniform grid.
//
// Replaces SimpleKNN::knn (KNN/simple_knn.cu:185-221) and its kernels coord2Morton / boxMinMax /
// boxMeanDist (simple_knn.cu:45-183).  The reference sorts points by a 30-bit Morton code, cuts the
// sorted list into boxes of 1024 points and lets every point scan all boxes whose AABB is close
// enough, brute-forcing inside a box.  Its pruning is conservative, so the result is the exact
// 3-NN of every point over all OTHER indices; that is what this file computes too, with the same
// distance expression, but the candidate set comes from a dense cell list:
//   bounds   -> AABB of the cloud (ordered-int atomics, no host round trip)
//   cells    -> Morton code of each point's cell on a 2^L-per-axis cubic grid (L from P)
//   sort     -> the library's onesweep radix sort on the 3L code bits (radix_sort.cu)
//   ranges   -> [start, end) of every cell in the sorted list + points gathered into sorted order
//   query    -> one thread per point (sorted order: a warp shares its cells), rings of cells
//               around the own cell until the third-best distance is inside the searched block.
#include <float.h>
#include "common.cuh"

namespace csb {

enum {
    KNN_CTRL_N = 0,        // P, element count of the sort
    KNN_CTRL_TICKETS = 4,  // 4 radix pass tickets
    KNN_CTRL_BOUNDS = 8,   // min.xyz, max.xyz as ordered ints
    KNN_CTRL_HIST = 16,    // 4 x 256
    KNN_CTRL_WORDS = 16 + 4 * 256,
};

static inline int knn_level(int P) {
    // about 4 points per cell for a cloud that fills its box; 2^L cells per axis, L <= 8
    int L = 0;
    while (L < 8 && ((size_t)1 << (3 * (L + 1))) * 4 <= (size_t)P) ++L;
    return L;
}

struct KnnState {
    uint32_t* control;
    uint32_t* lookback;
    size_t lookback_words;
    uint32_t *keys_a, *keys_b, *vals_a, *vals_b;
    float4* sorted;  // x, y, z, original index bits
    uint2* cells;    // [8^L]
    size_t n_cells;
    int L;
    size_t bytes;
    static KnnState carve(void* ws, int P) {
        KnnState s;
        Carver c(ws);
        s.L = knn_level(P);
        s.n_cells = (size_t)1 << (3 * s.L);
        s.control = c.take<uint32_t>(KNN_CTRL_WORDS);
        s.lookback_words = sort_lookback_words((size_t)P, 4);
        s.lookback = c.take<uint32_t>(s.lookback_words);
        s.keys_a = c.take<uint32_t>(P);
        s.keys_b = c.take<uint32_t>(P);
        s.vals_a = c.take<uint32_t>(P);
        s.vals_b = c.take<uint32_t>(P);
        s.sorted = c.take<float4>(P);
        s.cells = c.take<uint2>(s.n_cells);
        s.bytes = c.total();
        return s;
    }
};

__device__ __forceinline__ int float_to_ordered(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ordered_to_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__global__ void knn_init_kernel(uint32_t* control, int P) {
    const int t = threadIdx.x;
    if (t == 0) control[KNN_CTRL_N] = (uint32_t)P;
    if (t < 3) ((int*)control)[KNN_CTRL_BOUNDS + t] = 0x7FFFFFFF;          // +max ordered
    if (t >= 3 && t < 6) ((int*)control)[KNN_CTRL_BOUNDS + t] = (int)0x80000000;  // -max ordered
}

__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ pts, uint32_t* control) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < (size_t)P; i += (size_t)gridDim.x * 256) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float v = pts[3 * i + k];
            mn[k] = fminf(mn[k], v);
            mx[k] = fmaxf(mx[k], v);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xFFFFFFFFu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xFFFFFFFFu, mx[k], o));
        }
    if ((threadIdx.x & 31) == 0) {
        int* b = (int*)control + KNN_CTRL_BOUNDS;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(&b[k], float_to_ordered(mn[k]));
            atomicMax(&b[3 + k], float_to_ordered(mx[k]));
        }
    }
}

struct KnnGrid {
    float3 mn;
    float h, inv_h;
    int G;
};

__device__ __forceinline__ KnnGrid knn_grid(const uint32_t* control, int L) {
    const int* b = (const int*)control + KNN_CTRL_BOUNDS;
    KnnGrid g;
    g.mn = make_float3(ordered_to_float(b[0]), ordered_to_float(b[1]), ordered_to_float(b[2]));
    const float ex = ordered_to_float(b[3]) - g.mn.x, ey = ordered_to_float(b[4]) - g.mn.y,
                ez = ordered_to_float(b[5]) - g.mn.z;
    g.G = 1 << L;
    const float ext = fmaxf(ex, fmaxf(ey, ez));
    g.h = ext / (float)g.G;
    g.inv_h = g.h > 0.f ? 1.0f / g.h : 0.f;
    if (!(g.inv_h < 3.0e38f)) g.inv_h = 0.f;  // degenerate cloud: everything lands in cell 0
    return g;
}

// Monotone map coordinate -> cell index (the ring bound below relies on monotonicity only).
__device__ __forceinline__ float cell_coord(float v, float mn, float inv_h) { return (v - mn) * inv_h; }
__device__ __forceinline__ int cell_index(float f, int G) { return min(G - 1, max(0, (int)f)); }

__device__ __forceinline__ uint32_t spread3(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__device__ __forceinline__ uint32_t cell_code(int ix, int iy, int iz) {
    return spread3((uint32_t)ix) | (spread3((uint32_t)iy) << 1) | (spread3((uint32_t)iz) << 2);
}

__global__ void __launch_bounds__(256)
knn_cells_kernel(int P, int L, const float* __restrict__ pts, const uint32_t* __restrict__ control,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (size_t)P) return;
    const KnnGrid g = knn_grid(control, L);
    const int ix = cell_index(cell_coord(pts[3 * i], g.mn.x, g.inv_h), g.G);
    const int iy = cell_index(cell_coord(pts[3 * i + 1], g.mn.y, g.inv_h), g.G);
    const int iz = cell_index(cell_coord(pts[3 * i + 2], g.mn.z, g.inv_h), g.G);
    keys[i] = cell_code(ix, iy, iz);
    vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256)
knn_ranges_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ keys,
                  const uint32_t* __restrict__ vals, float4* __restrict__ sorted, uint2* __restrict__ cells) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (size_t)P) return;
    const uint32_t id = vals[i];
    sorted[i] = make_float4(pts[3 * (size_t)id], pts[3 * (size_t)id + 1], pts[3 * (size_t)id + 2], __uint_as_float(id));
    const uint32_t cur = keys[i];
    if (i == 0) cells[cur].x = 0;
    else {
        const uint32_t prev = keys[i - 1];
        if (prev != cur) {
            cells[prev].y = (uint32_t)i;
            cells[cur].x = (uint32_t)i;
        }
    }
    if (i == (size_t)P - 1) cells[cur].y = (uint32_t)P;
}

__device__ __forceinline__ void update3(float dist, float* best) {
    // insertion into the ascending triple (simple_knn.cu:131-145)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        if (best[j] > dist) {
            const float t = best[j];
            best[j] = dist;
            dist = t;
        }
    }
}

__global__ void __launch_bounds__(128)
knn_query_kernel(int P, int L, const uint32_t* __restrict__ control, const float4* __restrict__ sorted,
                 const uint2* __restrict__ cells, float* __restrict__ dists) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= P) return;
    const KnnGrid g = knn_grid(control, L);
    const float4 me = sorted[i];
    const float fx = cell_coord(me.x, g.mn.x, g.inv_h), fy = cell_coord(me.y, g.mn.y, g.inv_h),
                fz = cell_coord(me.z, g.mn.z, g.inv_h);
    const int cx = cell_index(fx, g.G), cy = cell_index(fy, g.G), cz = cell_index(fz, g.G);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    const float margin = 1.0e-3f * g.h;

    for (int r = 1;; ++r) {
        const int x0 = max(cx - r, 0), x1 = min(cx + r, g.G - 1);
        const int y0 = max(cy - r, 0), y1 = min(cy + r, g.G - 1);
        const int z0 = max(cz - r, 0), z1 = min(cz + r, g.G - 1);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                // only the shell of ring r: inner cells were visited by earlier rings
                const bool edge_zy = (r == 1) || abs(z - cz) == r || abs(y - cy) == r;
                for (int x = x0; x <= x1; ++x) {
                    if (!edge_zy && abs(x - cx) != r) {
                        x = max(x, cx + r - 1);  // jump over the interior of this row
                        continue;
                    }
                    const uint2 range = cells[cell_code(x, y, z)];
                    for (uint32_t j = range.x; j < range.y; ++j) {
                        if ((int)j == i) continue;
                        const float4 o = sorted[j];
                        const float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
                        // d.x*d.x + d.y*d.y + d.z*d.z as nvcc contracts it in the reference
                        // (SASS of boxMeanDist: FMUL dy*dy; FFMA dx*dx + .; FFMA dz*dz + .)
                        const float dist = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
                        update3(dist, best);
                    }
                }
            }
        // lower bound of the distance to any point outside the searched block
        float bound = FLT_MAX;
        if (cx + r + 1 <= g.G - 1) bound = fminf(bound, ((float)(cx + r + 1) - fx) * g.h);
        if (cy + r + 1 <= g.G - 1) bound = fminf(bound, ((float)(cy + r + 1) - fy) * g.h);
        if (cz + r + 1 <= g.G - 1) bound = fminf(bound, ((float)(cz + r + 1) - fz) * g.h);
        if (cx - r >= 1) bound = fminf(bound, (fx - (float)(cx - r)) * g.h);
        if (cy - r >= 1) bound = fminf(bound, (fy - (float)(cy - r)) * g.h);
        if (cz - r >= 1) bound = fminf(bound, (fz - (float)(cz - r)) * g.h);
        if (bound == FLT_MAX) break;  // the block covers the whole grid
        bound = fmaxf(bound - margin, 0.f);
        if (best[2] <= bound * bound) break;
    }
    dists[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

}  // namespace csb

using namespace csb;

extern "C" {

size_t csb_dist_cuda2_bytes(int P) { return KnnState::carve(nullptr, P < 0 ? 0 : P).bytes; }

int csb_dist_cuda2(int P, const float* points, float* mean_dists, void* ws, size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || (P > 0 && (!points || !mean_dists))) return CSB_ERR_BAD_ARG;
    if (P == 0) return 0;
    if ((size_t)P >= (1ull << 30)) return CSB_ERR_UNSUPPORTED;
    KnnState s = KnnState::carve(ws, P);
    if (!ws || ws_bytes < s.bytes) return CSB_ERR_WORKSPACE;
    CSB_CUDA_TRY(cudaMemsetAsync(s.control, 0, KNN_CTRL_WORDS * sizeof(uint32_t), stream));
    CSB_CUDA_TRY(cudaMemsetAsync(s.lookback, 0, s.lookback_words * sizeof(uint32_t), stream));
    CSB_CUDA_TRY(cudaMemsetAsync(s.cells, 0, s.n_cells * sizeof(uint2), stream));
    CSB_LAUNCH("knn_init", stream, knn_init_kernel<<<1, 32, 0, stream>>>(s.control, P));
    int rblocks = (P + 255) / 256;
    if (rblocks > 148 * 8) rblocks = 148 * 8;
    CSB_LAUNCH("knn_bounds", stream, knn_bounds_kernel<<<rblocks, 256, 0, stream>>>(P, points, s.control));
    const unsigned blocks = (unsigned)((P + 255) / 256);
    CSB_LAUNCH("knn_cells", stream, knn_cells_kernel<<<blocks, 256, 0, stream>>>(P, s.L, points, s.control, s.keys_a, s.vals_a));
    const uint32_t *keys = s.keys_a, *vals = s.vals_a;
    if (s.L > 0) {
        int side = 0;
        const int rc = radix_sort_pairs(s.keys_a, s.keys_b, s.vals_a, s.vals_b, s.control + KNN_CTRL_N, (size_t)P,
                                        3 * s.L, s.control + KNN_CTRL_HIST, s.control + KNN_CTRL_TICKETS, s.lookback,
                                        &side, stream);
        if (rc) return rc;
        keys = side ? s.keys_b : s.keys_a;
        vals = side ? s.vals_b : s.vals_a;
    }
    CSB_LAUNCH("knn_ranges", stream, knn_ranges_kernel<<<blocks, 256, 0, stream>>>(P, points, keys, vals, s.sorted, s.cells));
    CSB_LAUNCH("knn_query", stream, knn_query_kernel<<<(P + 127) / 128, 128, 0, stream>>>(P, s.L, s.control, s.sorted, s.cells, mean_dists));
    return 0;
}

}  // extern "C"
