// Forward pass of the tile rasterizer: preprocess -> depth order -> emit -> tile sort -> ranges -> render.
//
// Replaces CudaRasterizer::Rasterizer::forward (RAST/cuda_rasterizer/rasterizer_impl.cu:198-339)
// and its kernels preprocessCUDA / duplicateWithKeys / identifyTileRanges / renderCUDA
// (forward.cu:155-256, rasterizer_impl.cu:70-138, forward.cu:261-379).
#include "render_lanes.cuh"

namespace csb {

// ----------------------------------------------------------------------------------
// workspaces
// ----------------------------------------------------------------------------------
static inline size_t scan_blocks(int V, int P) { return ((size_t)V * P + 255) / 256; }

GeomState GeomState::carve(void* ws, int V, int P) {
    GeomState g;
    Carver c(ws);
    const size_t n = (size_t)V * P;
    g.means2D = c.take<float2>(n);
    g.conic_opacity = c.take<float4>(n);
    g.rgb_depth = c.take<float4>(n);
    g.cov3D = c.take<float>((size_t)P * 6);
    g.clamped = c.take<uint8_t>(n);
    g.tiles_touched = c.take<uint32_t>(n);
    g.rects = c.take<uint4>(n);
    g.depth_keys = c.take<uint32_t>(n);
    g.order_a = c.take<uint32_t>(n);
    g.order_b = c.take<uint32_t>(n);
    g.keys_b = c.take<uint32_t>(n);
    g.grad_acc = c.take<float>(n * 12);
    g.control_words = CTRL_FIXED_WORDS + 2 * (scan_blocks(V, P) + 1) + sort_lookback_words(n, 4);
    g.control = c.take<uint32_t>(g.control_words);
    g.bytes = c.total();
    return g;
}

ImageState ImageState::carve(void* ws, int V, int W, int H) {
    ImageState s;
    Carver c(ws);
    const size_t npix = (size_t)V * W * H;
    const size_t tiles = (size_t)V * ((W + CSB_TILE_X - 1) / CSB_TILE_X) * ((H + CSB_TILE_Y - 1) / CSB_TILE_Y);
    s.final_T = c.take<float>(npix);
    s.n_contrib = c.take<uint32_t>(npix);
    s.ranges = c.take<uint2>(tiles);
    s.tile_order = c.take<uint32_t>(tiles + TILE_ORDER_HEAD);
    s.bytes = c.total();
    return s;
}

BinState BinState::carve(void* ws, int V, int P, int W, int H, size_t cap) {
    (void)V; (void)P; (void)W; (void)H;
    BinState b;
    Carver c(ws);
    b.tiles_a = c.take<uint32_t>(cap);
    b.tiles_b = c.take<uint32_t>(cap);
    b.vals_a = c.take<uint32_t>(cap);
    b.vals_b = c.take<uint32_t>(cap);
    b.keys = c.take<uint64_t>(cap);
    b.lookback_words = sort_lookback_words(cap, 4);
    b.lookback = c.take<uint32_t>(b.lookback_words);
    b.bytes = c.total();
    return b;
}

// ----------------------------------------------------------------------------------
// preprocess: one thread per (view, Gaussian)
// ----------------------------------------------------------------------------------
__device__ __forceinline__ float3 sh_to_rgb(int deg, int M, const float* __restrict__ sh, const float3 pos,
                                            const float3 campos, uint8_t& clamped) {
    // direction from the camera, normalised by division (forward.cu:27-29)
    float3 dir = make_float3(pos.x - campos.x, pos.y - campos.y, pos.z - campos.z);
    const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x = dir.x / len;
    dir.y = dir.y / len;
    dir.z = dir.z / len;
    const float x = dir.x, y = dir.y, z = dir.z;
    float res[3];
    (void)M;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
#define SHC(k) sh[(k) * 3 + ch]
        float r = kSH_C0 * SHC(0);
        if (deg > 0) {
            r = r - kSH_C1 * y * SHC(1) + kSH_C1 * z * SHC(2) - kSH_C1 * x * SHC(3);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z;
                const float xy = x * y, yz = y * z, xz = x * z;
                r = r + kSH_C2[0] * xy * SHC(4) + kSH_C2[1] * yz * SHC(5) +
                    kSH_C2[2] * (2.0f * zz - xx - yy) * SHC(6) + kSH_C2[3] * xz * SHC(7) +
                    kSH_C2[4] * (xx - yy) * SHC(8);
                if (deg > 2) {
                    r = r + kSH_C3[0] * y * (3.0f * xx - yy) * SHC(9) + kSH_C3[1] * xy * z * SHC(10) +
                        kSH_C3[2] * y * (4.0f * zz - xx - yy) * SHC(11) +
                        kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHC(12) +
                        kSH_C3[4] * x * (4.0f * zz - xx - yy) * SHC(13) + kSH_C3[5] * z * (xx - yy) * SHC(14) +
                        kSH_C3[6] * x * (xx - 3.0f * yy) * SHC(15);
                }
            }
        }
#undef SHC
        r += 0.5f;
        res[ch] = r;
    }
    clamped = (uint8_t)((res[0] < 0 ? 1 : 0) | (res[1] < 0 ? 2 : 0) | (res[2] < 0 ? 4 : 0));
    return make_float3(fmaxf(res[0], 0.0f), fmaxf(res[1], 0.0f), fmaxf(res[2], 0.0f));
}

// One (view, Gaussian) entry; returns its depth-sort key (0xFFFFFFFF when culled).
__device__ __forceinline__ uint32_t
preprocess_entry(const size_t e, const ViewBatch& vb, int D, int M, const float* __restrict__ means3D,
                 const float* __restrict__ scales, float scale_modifier, const float* __restrict__ rotations,
                 const float* __restrict__ opacities, const float* __restrict__ shs,
                 const float* __restrict__ cov3D_precomp, const float* __restrict__ colors_precomp,
                 int* __restrict__ radii, const GeomState& g, const float* sh_row) {
    const int v = (int)(e / vb.P);
    const int idx = (int)(e - (size_t)v * vb.P);

    radii[e] = 0;
    g.tiles_touched[e] = 0;
    g.order_a[e] = (uint32_t)e;
    g.depth_keys[e] = 0xFFFFFFFFu;

    const float* __restrict__ vm = vb.viewmatrix[v];
    const float* __restrict__ pm = vb.projmatrix[v];
    const float3 p_orig = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);

    // near culling (auxiliary.h:139-164): only view-space z is tested
    const float3 p_view = xform_point_4x3(p_orig, vm);
    if (p_view.z <= 0.2f) return 0xFFFFFFFFu;

    const float4 p_hom = xform_point_4x4(p_orig, pm);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = make_float3(p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w);

    float cov6[6];
    if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; ++i) cov6[i] = cov3D_precomp[6 * (size_t)idx + i];
    } else {
        const float3 s = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
        const float4 q = reinterpret_cast<const float4*>(rotations)[idx];
        cov3d_from_scale_rot(s, scale_modifier, q, cov6);
#pragma unroll
        for (int i = 0; i < 6; ++i) g.cov3D[6 * (size_t)idx + i] = cov6[i];  // same bits from every view
    }

    const EwaProj ep = ewa_project(p_orig, vb.focal_x[v], vb.focal_y[v], vb.tan_fovx[v], vb.tan_fovy[v], vm);
    const float3 cov = cov2d_from_proj(ep, cov6);

    const float det = (cov.x * cov.z - cov.y * cov.y);
    if (det == 0.0f) return 0xFFFFFFFFu;
    const float det_inv = 1.f / det;
    const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);

    const float mid = 0.5f * (cov.x + cov.z);
    const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    const float2 point_image = make_float2(ndc_to_pix(p_proj.x, vb.W), ndc_to_pix(p_proj.y, vb.H));
    uint2 rect_min, rect_max;
    tile_rect(point_image, (int)my_radius, vb.gx, vb.gy, rect_min, rect_max);
    const uint32_t area = (rect_max.x - rect_min.x) * (rect_max.y - rect_min.y);
    if (area == 0) return 0xFFFFFFFFu;

    float3 rgb;
    uint8_t clamped = 0;
    if (colors_precomp == nullptr) {
        const float3 campos = make_float3(vb.campos[v][0], vb.campos[v][1], vb.campos[v][2]);
        rgb = sh_to_rgb(D, M, sh_row ? sh_row : shs + (size_t)idx * M * 3, p_orig, campos, clamped);
    } else {
        rgb = make_float3(colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]);
    }

    radii[e] = (int)my_radius;
    g.means2D[e] = point_image;
    g.conic_opacity[e] = make_float4(conic.x, conic.y, conic.z, opacities[idx]);
    g.rgb_depth[e] = make_float4(rgb.x, rgb.y, rgb.z, p_view.z);
    g.clamped[e] = clamped;
    g.tiles_touched[e] = area;
    g.rects[e] = make_uint4(rect_min.x, rect_min.y, rect_max.x - rect_min.x, __float_as_uint(p_view.z));
    g.depth_keys[e] = __float_as_uint(p_view.z);
    return __float_as_uint(p_view.z);
}

// One thread per (view, Gaussian) entry.  The digit histograms of the depth sort's four passes are accumulated
// here (shared-memory bins, flushed once per block), so the sort needs no histogram pass of its own.
__global__ void __launch_bounds__(256)
preprocess_kernel(const ViewBatch vb, int D, int M, const float* __restrict__ means3D,
                  const float* __restrict__ scales, float scale_modifier, const float* __restrict__ rotations,
                  const float* __restrict__ opacities, const float* __restrict__ shs,
                  const float* __restrict__ cov3D_precomp, const float* __restrict__ colors_precomp,
                  int* __restrict__ radii, GeomState g, int stage_sh, uint2* __restrict__ count_ranges) {
    __shared__ uint32_t sh[4 * 256];
    extern __shared__ float s_rec[];  // [256][M*3+1] SH records of the block (only when the launch provides it)
    grid_dep_wait();
    grid_dep_launch();
    for (int i = threadIdx.x; i < 4 * 256; i += 256) sh[i] = 0;
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;  // entry = v*P + idx
    const size_t n = (size_t)vb.V * vb.P;
    if (e == 0) g.control[CTRL_N_ENTRIES] = (uint32_t)n;  // element count of the depth sort
    // Stage the block's SH records with coalesced 16-byte loads (a thread walking its own 192-byte record touches
    // 32 lines per warp instruction).  Blocks that straddle a view boundary read their records directly.
    const float* sh_row = nullptr;
    {
        const size_t e0 = (size_t)blockIdx.x * 256;
        const size_t e1 = (e0 + 255 < n ? e0 + 255 : n - 1);
        const bool one_view = stage_sh && e0 < n && (e0 / vb.P) == (e1 / vb.P);
        if (one_view) {
            const int rec = M * 3, pitch = rec + 1;
            const size_t first = e0 % vb.P;
            const int count = (int)(e1 - e0 + 1);
            if (rec == 48) stage_records_in<48>(shs + first * rec, s_rec, count, rec, threadIdx.x, 256);
            else stage_records_in<0>(shs + first * rec, s_rec, count, rec, threadIdx.x, 256);
            sh_row = s_rec + threadIdx.x * pitch;
        }
    }
    __syncthreads();
    uint32_t key = 0;
    if (e < n)
        key = preprocess_entry(e, vb, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, cov3D_precomp,
                               colors_precomp, radii, g, sh_row);
    if (count_ranges && e < n) {
        // per-tile binning path: the entry counts itself into the tiles of its rectangle right here (this was a kernel
        // of its own, bin_entries_kernel<false>: ~8 us of launch + load latency for 40 k entries at C2 size)
        const uint32_t cnt = g.tiles_touched[e];
        if (cnt) {
            const uint4 rect = g.rects[e];
            const uint32_t view = (uint32_t)(e / (size_t)vb.P);
            const uint32_t first = view * (uint32_t)vb.tiles + rect.y * (uint32_t)vb.gx + rect.x;
            const float inv_w = 1.0f / (float)rect.z;
            for (uint32_t k = 0; k < cnt; ++k) {
                const uint32_t ry = (uint32_t)(((float)k + 0.5f) * inv_w);
                atomicAdd(&count_ranges[first + ry * (uint32_t)vb.gx + (k - ry * rect.z)].y, 1u);
            }
        }
    }
    warp_hist_add(sh, 4, key, e < n, threadIdx.x & 31);
    __syncthreads();
    uint32_t* hist = g.control + CTRL_HIST_DEPTH;
    for (int i = threadIdx.x; i < 4 * 256; i += 256) {
        const uint32_t c = sh[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// ----------------------------------------------------------------------------------
// emit: chained scan of tiles_touched in depth order + warp-cooperative instance emission
// ----------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

constexpr uint64_t SCAN_AGG = 1ull << 62;
constexpr uint64_t SCAN_INC = 2ull << 62;
constexpr uint64_t SCAN_VAL = (1ull << 62) - 1;

__global__ void __launch_bounds__(256)
emit_kernel(const ViewBatch vb, const uint32_t* __restrict__ order, GeomState g, uint64_t* __restrict__ scan_state,
            uint32_t* __restrict__ out_tiles, uint32_t* __restrict__ out_vals, size_t capacity,
            uint32_t* __restrict__ num_rendered_out, uint32_t* num_rendered_mapped, int tile_passes) {
    __shared__ uint32_t sh[3 * 256];  // digit histograms of the tile sort's passes
    __shared__ uint32_t s_block;
    __shared__ uint32_t s_warp_tot[8];
    __shared__ uint64_t s_excl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    grid_dep_wait();
    grid_dep_launch();
    if (tid == 0) s_block = atomicAdd(&g.control[CTRL_TICKET_SCAN], 1u);
    for (int i = tid; i < 3 * 256; i += 256) sh[i] = 0;
    __syncthreads();
    const uint32_t blk = s_block;
    const size_t n = (size_t)vb.V * vb.P;
    const size_t j = (size_t)blk * 256 + tid;

    uint32_t e = 0, cnt = 0;
    uint4 rect = make_uint4(0, 0, 1, 0);
    if (j < n) {
        e = order[j];
        cnt = g.tiles_touched[e];
        if (cnt) rect = g.rects[e];
    }
    // exclusive scan inside the warp, then across the block's warps
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    const uint32_t wexcl = incl - cnt;
    const uint32_t wtotal = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (lane == 31) s_warp_tot[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_tot = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        if (w < warp) warp_off += s_warp_tot[w];
        block_tot += s_warp_tot[w];
    }
    // decoupled look-back over earlier blocks (warp 0)
    if (warp == 0) {
        uint64_t excl = 0;
        if (blk == 0) {
            if (lane == 0) st_volatile_u64(&scan_state[0], SCAN_INC | block_tot);
        } else {
            if (lane == 0) st_volatile_u64(&scan_state[blk], SCAN_AGG | block_tot);
            int pos = (int)blk - 1;
            while (true) {
                const int q = pos - lane;
                const uint64_t s = q >= 0 ? ld_volatile_u64(&scan_state[q]) : SCAN_INC;
                const uint32_t ready = __ballot_sync(0xFFFFFFFFu, (s & ~SCAN_VAL) != 0);
                const uint32_t inc = __ballot_sync(0xFFFFFFFFu, (s & ~SCAN_VAL) == SCAN_INC);
                if (inc) {
                    const int fi = __ffs(inc) - 1;  // nearest predecessor holding an inclusive prefix
                    const uint32_t need = fi == 31 ? 0xFFFFFFFFu : ((1u << (fi + 1)) - 1);
                    if ((ready & need) != need) continue;
                    uint64_t part = lane <= fi ? (s & SCAN_VAL) : 0;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
                    excl += part;
                    break;
                }
                if (ready != 0xFFFFFFFFu) continue;
                uint64_t part = s & SCAN_VAL;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
                excl += part;
                pos -= 32;
            }
            if (lane == 0) st_volatile_u64(&scan_state[blk], SCAN_INC | (excl + block_tot));
        }
        if (lane == 0) {
            s_excl = excl;
            if (blk == gridDim.x - 1) {  // the last block knows the total
                const uint64_t total = excl + block_tot;
                const bool overflow = total > capacity;
                g.control[CTRL_NUM_RENDERED] = overflow ? 0u : (uint32_t)total;  // n of the later stages
                g.control[CTRL_OVERFLOW] = overflow ? 1u : 0u;
                const uint32_t clamped = (uint32_t)min(total, (uint64_t)0xFFFFFFFEull);
                num_rendered_out[0] = clamped;
                if (num_rendered_mapped) {
                    // pinned host word the caller polls: the count reaches the host while the tile sort and the
                    // render kernel are still running, so the forward needs no stream synchronisation
                    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(num_rendered_mapped), "r"(clamped) : "memory");
                    __threadfence_system();
                }
            }
        }
    }
    __syncthreads();
    const uint64_t warp_base = s_excl + warp_off;
    // overflow: nothing is emitted, the caller grows the workspace and retries
    const uint32_t wemit = (warp_base + wtotal > capacity) ? 0u : wtotal;

    // flat emission: item k of the warp belongs to the lane g with wexcl[g] <= k < wexcl[g] + cnt[g]
    const float inv_w = 1.0f / (float)rect.z;
    for (uint32_t k = lane; k < ((wemit + 31) & ~31u); k += 32) {
        int gl = 0;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            const int c = gl + s;
            const uint32_t v = __shfl_sync(0xFFFFFFFFu, wexcl, c & 31);
            if (c < 32 && v <= k) gl = c;
        }
        // skip over lanes without instances that share the same prefix
        const uint32_t g_excl = __shfl_sync(0xFFFFFFFFu, wexcl, gl);
        const uint32_t g_min_x = __shfl_sync(0xFFFFFFFFu, rect.x, gl);
        const uint32_t g_min_y = __shfl_sync(0xFFFFFFFFu, rect.y, gl);
        const uint32_t g_w = __shfl_sync(0xFFFFFFFFu, rect.z, gl);
        const float g_inv_w = __shfl_sync(0xFFFFFFFFu, inv_w, gl);
        const uint32_t g_e = __shfl_sync(0xFFFFFFFFu, e, gl);
        uint32_t tile = 0;
        if (k < wemit) {
            const uint32_t local = k - g_excl;
            const uint32_t ry = (uint32_t)(((float)local + 0.5f) * g_inv_w);
            const uint32_t rx = local - ry * g_w;
            const uint32_t view = g_e / (uint32_t)vb.P;
            tile = view * (uint32_t)vb.tiles + (g_min_y + ry) * (uint32_t)vb.gx + (g_min_x + rx);
            out_tiles[warp_base + k] = tile;
            out_vals[warp_base + k] = g_e;
        }
        warp_hist_add(sh, tile_passes, tile, k < wemit, lane);
    }
    __syncthreads();
    uint32_t* hist = g.control + CTRL_HIST_TILE;
    for (int i = tid; i < tile_passes * 256; i += 256) {
        const uint32_t c = sh[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// ----------------------------------------------------------------------------------
// Per-tile binning for scenes with short tile lists (flags bit2): count -> scan -> fill -> per-tile sort.
//
// The global path above sorts V*P depth keys (4 radix passes), emits, and sorts the instances by tile (2-3 passes):
// 7-8 launches whose cost for a cloth-sized scene (40 k entries, 270 k instances) is launch latency, not traffic.
// Here every (view, Gaussian) entry counts itself into the tiles of its rectangle, one block scans the per-tile
// counts into the final `ranges`, the entries are filled in arbitrary (atomic) order, and one block per tile sorts
// its list by (depth bits, entry): keys are unique, so the result is the reference's order -- stable by depth, then by
// Gaussian index (rasterizer_impl.cu:70-111, 304-309) -- bit for bit and independent of the atomics' order.
// A tile list longer than TILE_SORT_MAX makes the scan kernel refuse (count word 0xFFFFFFFD): the caller falls back
// to the global path.
// ----------------------------------------------------------------------------------
constexpr uint32_t TILE_SORT_MAX = 2048;
constexpr uint32_t TILE_RANK_MAX = 256;  // lists up to here are rank-sorted (no barriers), longer ones bitonic
constexpr uint32_t COUNT_REFUSED = 0xFFFFFFFDu;

// Control words of this path (the depth sort's tickets are free here).
constexpr int CTRL_BIN_MAX = CTRL_TICKET_DEPTH + 1;     // longest tile list
constexpr int CTRL_BIN_TICKET = CTRL_TICKET_DEPTH + 2;  // block ticket of bin_ranges_kernel

// FILL = false: every (view, Gaussian) entry counts itself into the tiles of its rectangle.  FILL = true: the same walk claims a slot per tile and stores
// (depth bits, entry) there.
// Longest tile list first.  The compositing kernels give every tile a CTA and the hardware hands CTAs to the SMs in index
// order, so the order of the indices IS the schedule: in image order the cloth's heavy tiles (lists of 100-220 entries in
// a quarter of the tiles, the rest empty) arrive in clumps and the kernels end on a tail of half-empty SMs (ncu: achieved
// occupancy 76-82 % of the theoretical one).  Sorted by descending list length the big tiles start first and the small
// and empty ones fill the gaps -- the classic longest-processing-time rule, for the price of one indirection per CTA.
// Counting sort of the tiles by length class (8 entries per class, class 0 = the longest lists): bin_ranges_kernel adds its
// tiles to the 256-bin class histogram, extra blocks of the fill launch (one per 256 tiles, so they run beside the fill)
// scan it and place their tiles with one cursor atomic each.
__device__ __forceinline__ uint32_t tile_len_class(uint32_t len) { return 255u - min((len + 7u) >> 3, 255u); }

__device__ void build_tile_order(const uint2* __restrict__ ranges, uint32_t n_tiles, uint32_t total, uint32_t* __restrict__ order,
                                 uint32_t chunk, bool closed_ranges = false) {
    __shared__ uint32_t s_base[256], s_warp[8];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // exclusive scan of the class histogram (256 bins, one per thread)
    const uint32_t h = __ldcg(&order[1 + tid]);
    uint32_t inc = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= (uint32_t)o) inc += u;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w = 0; w < warp; ++w) off += s_warp[w];
    s_base[tid] = off + inc - h;
    __syncthreads();
    const uint32_t t = chunk * 256u + tid;
    uint32_t cls = 0xFFFFFFFFu;
    if (t < n_tiles) {
        // per-tile binning: only the starts are final while the fill runs; global path: finished (start, end) pairs
        const uint32_t a = __ldcg(&ranges[t].x);
        const uint32_t b = closed_ranges ? __ldcg(&ranges[t].y) : (t + 1 < n_tiles ? __ldcg(&ranges[t + 1].x) : total);
        cls = tile_len_class(b - a);
    }
    // one cursor atomic per (warp, class): three quarters of a cloth frame's tiles share the class "empty"
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, cls);
    const uint32_t leader = __ffs(peers) - 1;
    uint32_t slot0 = 0;
    if (lane == leader && t < n_tiles) slot0 = atomicAdd(&order[257 + cls], (uint32_t)__popc(peers));
    slot0 = __shfl_sync(0xFFFFFFFFu, slot0, leader);
    if (t < n_tiles) order[TILE_ORDER_HEAD + s_base[cls] + slot0 + __popc(peers & ((1u << lane) - 1u))] = t;
    if (chunk == 0 && tid == 0) order[0] = 1u;  // the consumers start after this whole grid has finished
}

// The same order for the global sort path, where the ranges come from tile_ranges_kernel: two small launches behind it
// (class histogram; scan + placement).
__global__ void __launch_bounds__(256)
tile_order_hist_kernel(const uint2* __restrict__ ranges, uint32_t n_tiles, uint32_t* __restrict__ order) {
    grid_dep_wait();
    grid_dep_launch();
    const uint32_t t = blockIdx.x * 256u + threadIdx.x, lane = threadIdx.x & 31;
    uint32_t cls = 0xFFFFFFFFu;
    if (t < n_tiles) {
        const uint2 r = ranges[t];
        cls = tile_len_class(r.y - r.x);
    }
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, cls);
    if (lane == (uint32_t)__ffs(peers) - 1u && t < n_tiles) atomicAdd(&order[1 + cls], (uint32_t)__popc(peers));
}
__global__ void __launch_bounds__(256)
tile_order_place_kernel(const uint2* __restrict__ ranges, uint32_t n_tiles, uint32_t* __restrict__ order) {
    grid_dep_wait();
    grid_dep_launch();
    build_tile_order(ranges, n_tiles, 0u, order, blockIdx.x, true);
}

template <bool FILL>
__global__ void __launch_bounds__(256)
bin_entries_kernel(const ViewBatch vb, GeomState g, uint2* __restrict__ ranges, uint32_t* __restrict__ groups,
                   uint32_t* __restrict__ fill_depth, uint32_t* __restrict__ fill_vals, uint32_t* __restrict__ tile_order) {
    grid_dep_wait();
    grid_dep_launch();
    if (FILL && g.control[CTRL_OVERFLOW]) return;
    const uint32_t entry_blocks = (uint32_t)(((size_t)vb.V * vb.P + 255) / 256);
    if (FILL && tile_order && blockIdx.x >= entry_blocks) {  // the extra blocks: no entries of their own
        build_tile_order(ranges, (uint32_t)vb.V * (uint32_t)vb.tiles, g.control[CTRL_NUM_RENDERED], tile_order,
                         blockIdx.x - entry_blocks);
        return;
    }
    const size_t n = (size_t)vb.V * vb.P;
    const size_t j = (size_t)blockIdx.x * 256 + threadIdx.x;
    const uint32_t cnt = j < n ? g.tiles_touched[j] : 0u;
    if (!cnt) return;
    const uint4 rect = g.rects[j];
    const uint32_t view = (uint32_t)(j / (size_t)vb.P);
    const uint32_t first = view * (uint32_t)vb.tiles + rect.y * (uint32_t)vb.gx + rect.x;
    const float inv_w = 1.0f / (float)rect.z;
    // tile k of the rectangle (row-major); four slot claims in flight per round trip
    for (uint32_t k = 0; k < cnt; k += 4) {
        uint32_t t[4], slot[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t kk = min(k + u, cnt - 1);
            const uint32_t ry = (uint32_t)(((float)kk + 0.5f) * inv_w);
            t[u] = first + ry * (uint32_t)vb.gx + (kk - ry * rect.z);
        }
        if (FILL) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (k + u < cnt) slot[u] = atomicAdd(&ranges[t[u]].y, 1u);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (k + u < cnt) {
                    fill_depth[slot[u]] = rect.w;
                    fill_vals[slot[u]] = (uint32_t)j;
                }
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (k + u < cnt) atomicAdd(&ranges[t[u]].y, 1u);
        }
    }
}

// Single-pass exclusive scan of the per-tile counts (chained scan with decoupled look-back, 256 tiles per block):
// ranges[t] = (start, start); the fill kernel then advances .y to the end of the run.  The block holding the last
// ticket ends up with the instance count and publishes it (device word, control word, pinned host word) -- or the
// refusal, when a tile list is too long for the per-tile sort.
constexpr uint32_t BS_AGG = 1u << 30, BS_INC = 2u << 30, BS_VAL = (1u << 30) - 1;
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256)
bin_ranges_kernel(uint2* __restrict__ ranges, uint32_t* __restrict__ status, uint32_t n_tiles, size_t capacity,
                  uint32_t* __restrict__ control, uint32_t* __restrict__ num_rendered_out,
                  uint32_t* num_rendered_mapped, uint32_t* __restrict__ tile_order) {
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_blk, s_excl;
    grid_dep_wait();
    grid_dep_launch();
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_blk = atomicAdd(&control[CTRL_BIN_TICKET], 1u);
    __syncthreads();
    const uint32_t blk = s_blk;
    const uint32_t t = blk * 256u + tid;
    const uint32_t c = t < n_tiles ? __ldcg(&ranges[t].y) : 0u;
    if (tile_order) {  // class histogram of build_tile_order, one atomic per (warp, class)
        const uint32_t cls = t < n_tiles ? tile_len_class(c) : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, cls);
        if (lane == (uint32_t)__ffs(peers) - 1u && t < n_tiles) atomicAdd(&tile_order[1 + cls], (uint32_t)__popc(peers));
    }
    uint32_t incl = c, mx = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
        mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    }
    if (lane == 31) s_warp[warp] = incl;
    if (lane == 0 && mx > TILE_SORT_MAX) atomicMax(&control[CTRL_BIN_MAX], mx);
    __syncthreads();
    uint32_t warp_off = 0, block_tot = 0;
#pragma unroll
    for (uint32_t w = 0; w < 8; ++w) {
        if (w < warp) warp_off += s_warp[w];
        block_tot += s_warp[w];
    }
    if (warp == 0) {
        uint32_t excl = 0;
        __threadfence();  // this block's atomicMax is visible before its status is
        if (blk == 0) {
            if (lane == 0) st_volatile_u32(&status[0], BS_INC | block_tot);
        } else {
            if (lane == 0) st_volatile_u32(&status[blk], BS_AGG | block_tot);
            int pos = (int)blk - 1;
            while (true) {
                const int q = pos - (int)lane;
                const uint32_t sv = q >= 0 ? ld_volatile_u32(&status[q]) : BS_INC;
                const uint32_t ready = __ballot_sync(0xFFFFFFFFu, (sv & ~BS_VAL) != 0);
                const uint32_t inc = __ballot_sync(0xFFFFFFFFu, (sv & ~BS_VAL) == BS_INC);
                if (inc) {
                    const int fi = __ffs(inc) - 1;  // nearest predecessor holding an inclusive prefix
                    const uint32_t need = fi == 31 ? 0xFFFFFFFFu : ((1u << (fi + 1)) - 1);
                    if ((ready & need) != need) continue;
                    uint32_t part = (int)lane <= fi ? (sv & BS_VAL) : 0u;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
                    excl += part;
                    break;
                }
                if (ready != 0xFFFFFFFFu) continue;
                uint32_t part = sv & BS_VAL;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
                excl += part;
                pos -= 32;
            }
            if (lane == 0) st_volatile_u32(&status[blk], BS_INC | (excl + block_tot));
        }
        if (lane == 0) {
            s_excl = excl;
            if (blk == gridDim.x - 1) {
                __threadfence();
                const uint32_t total = excl + block_tot;
                const bool refused = __ldcg(&control[CTRL_BIN_MAX]) > TILE_SORT_MAX;
                const bool overflow = (size_t)total > capacity;
                control[CTRL_NUM_RENDERED] = (refused || overflow) ? 0u : total;
                control[CTRL_OVERFLOW] = refused ? 2u : (overflow ? 1u : 0u);
                const uint32_t word = refused ? COUNT_REFUSED : total;
                num_rendered_out[0] = word;
                if (num_rendered_mapped) {
                    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(num_rendered_mapped), "r"(word) : "memory");
                    __threadfence_system();
                }
            }
        }
    }
    __syncthreads();
    if (t < n_tiles) {
        const uint32_t start = s_excl + warp_off + incl - c;
        ranges[t] = make_uint2(start, start);
    }
}

// Persistent blocks walk the tiles; each sorts its tile's (depth bits, entry) keys in shared memory: rank sort (no
// barriers) for short lists, bitonic network for long ones.
__global__ void __launch_bounds__(256)
bin_sort_kernel(uint2* __restrict__ ranges, const uint32_t* __restrict__ control, uint32_t n_tiles,
                const uint32_t* __restrict__ fill_depth, const uint32_t* __restrict__ fill_vals,
                uint32_t* __restrict__ out_tiles, uint32_t* __restrict__ out_vals, uint64_t* __restrict__ keys_out,
                const uint32_t* __restrict__ tile_order) {
    __shared__ uint64_t s_key[TILE_SORT_MAX];
    grid_dep_wait();
    grid_dep_launch();
    if (control[CTRL_OVERFLOW]) return;
    const uint32_t tid = threadIdx.x;
    // With the longest-first order of build_tile_order the blocks take the long lists first and never walk an empty tile
    // (three quarters of a cloth frame: each cost a block a global-load round trip); the empty tiles -- the tail of the
    // order, their count is the last class of its histogram -- are reset by a thread-parallel sweep.
    const bool ordered = tile_order != nullptr && tile_order[0] != 0;
    const uint32_t n_work = ordered ? n_tiles - tile_order[1 + 255] : n_tiles;
    if (ordered) {
        for (uint32_t i = n_work + blockIdx.x * 256u + tid; i < n_tiles; i += gridDim.x * 256u) {
            const uint32_t t = tile_order[TILE_ORDER_HEAD + i];
            if (ranges[t].x != 0) ranges[t] = make_uint2(0u, 0u);  // the reference leaves empty tiles at (0, 0)
        }
    }
    for (uint32_t idx = blockIdx.x; idx < n_work; idx += gridDim.x) {
        const uint32_t tile = ordered ? tile_order[TILE_ORDER_HEAD + idx] : idx;
        const uint2 rg = ranges[tile];
        const uint32_t n = rg.y - rg.x;
        if (n == 0) {
            if (tid == 0 && rg.x != 0) ranges[tile] = make_uint2(0u, 0u);  // the reference leaves empty tiles at (0, 0)
            continue;
        }
        __syncthreads();  // the previous tile's readers are done with s_key
        if (n <= TILE_RANK_MAX) {
            uint64_t mine = 0;
            if (tid < n) {
                mine = ((uint64_t)fill_depth[rg.x + tid] << 32) | fill_vals[rg.x + tid];
                s_key[tid] = mine;
            }
            __syncthreads();
            if (tid < n) {
                uint32_t rank = 0;
                for (uint32_t q = 0; q < n; ++q) rank += s_key[q] < mine ? 1u : 0u;  // keys are unique
                out_vals[rg.x + rank] = (uint32_t)mine;
                out_tiles[rg.x + rank] = tile;
                if (keys_out) keys_out[rg.x + rank] = ((uint64_t)tile << 32) | (mine >> 32);
            }
            continue;
        }
        uint32_t m = 1;
        while (m < n) m <<= 1;
        for (uint32_t i = tid; i < m; i += 256)
            s_key[i] = i < n ? ((uint64_t)fill_depth[rg.x + i] << 32) | fill_vals[rg.x + i] : ~0ull;
        __syncthreads();
        for (uint32_t k = 2; k <= m; k <<= 1) {
            for (uint32_t jj = k >> 1; jj > 0; jj >>= 1) {
                for (uint32_t i = tid; i < m; i += 256) {
                    const uint32_t p = i ^ jj;
                    if (p > i) {
                        const uint64_t a = s_key[i], b = s_key[p];
                        if ((a > b) == ((i & k) == 0)) {
                            s_key[i] = b;
                            s_key[p] = a;
                        }
                    }
                }
                __syncthreads();
            }
        }
        for (uint32_t i = tid; i < n; i += 256) {
            const uint64_t key = s_key[i];
            out_vals[rg.x + i] = (uint32_t)key;
            out_tiles[rg.x + i] = tile;
            if (keys_out) keys_out[rg.x + i] = ((uint64_t)tile << 32) | (key >> 32);
        }
    }
}

// ----------------------------------------------------------------------------------
// tile ranges (+ optional u64 keys) from the sorted tile ids
// ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ n_dev, const uint32_t* __restrict__ tiles,
                   const uint32_t* __restrict__ vals, const float4* __restrict__ rgb_depth,
                   uint2* __restrict__ ranges, uint64_t* __restrict__ keys_out, uint32_t tiles_per_view) {
    grid_dep_wait();
    grid_dep_launch();
    const uint32_t n = *n_dev;
    const size_t stride = (size_t)gridDim.x * 256;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const uint32_t cur = tiles[i];
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = tiles[i - 1];
            if (cur != prev) {
                ranges[prev].y = (uint32_t)i;
                ranges[cur].x = (uint32_t)i;
            }
        }
        if (i == n - 1) ranges[cur].y = n;
        if (keys_out) {
            const uint32_t d = __float_as_uint(rgb_depth[vals[i]].w);
            keys_out[i] = ((uint64_t)(cur % tiles_per_view) << 32) | d;
        }
    }
}

// ----------------------------------------------------------------------------------
// render: one 16x16 tile per block, front-to-back alpha compositing
// ----------------------------------------------------------------------------------
constexpr int RENDER_BATCH = 256;

__global__ void __launch_bounds__(CSB_TILE_PIX)
render_forward_kernel(const ViewBatch vb, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                      const float4* __restrict__ rgb_depth, float* __restrict__ final_T,
                      uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                      float* __restrict__ out_depth) {
    __shared__ float2 s_xy[RENDER_BATCH];
    __shared__ float4 s_co[RENDER_BATCH];
    __shared__ float4 s_cd[RENDER_BATCH];
    __shared__ uint32_t s_mask[RENDER_BATCH];  // bit w: the Gaussian can reach the 8x4 pixel patch of warp w

    grid_dep_wait();
    grid_dep_launch();
    const int v = blockIdx.z;
    const int tid = threadIdx.x;
    // a warp covers an 8x4 pixel patch of the tile (compact footprint -> coherent early-outs)
    const int warp = tid >> 5, lane = tid & 31;
    const int lx = (warp & 1) * 8 + (lane & 7);
    const int ly = (warp >> 1) * 4 + (lane >> 3);
    const uint32_t px = blockIdx.x * CSB_TILE_X + lx;
    const uint32_t py = blockIdx.y * CSB_TILE_Y + ly;
    const bool inside = px < (uint32_t)vb.W && py < (uint32_t)vb.H;
    const float pxf = (float)px, pyf = (float)py;
    const float tile_x0 = (float)(blockIdx.x * CSB_TILE_X), tile_y0 = (float)(blockIdx.y * CSB_TILE_Y);

    const uint2 range = ranges[(size_t)v * vb.tiles + blockIdx.y * vb.gx + blockIdx.x];
    int todo = (int)(range.y - range.x);
    const int rounds = (todo + RENDER_BATCH - 1) / RENDER_BATCH;

    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float2 C01 = make_float2(0.f, 0.f), C2D = make_float2(0.f, 0.f);  // (C0, C1), (C2, depth)

    // shared-window addresses, made opaque so that they stay in registers: otherwise the compiler rematerialises
    // them (S2R SR_CgaCtaId + LEA, ~8 instructions) in every iteration of the divergent inner loop
    uint32_t a_xy = smem_addr(s_xy), a_co = smem_addr(s_co), a_cd = smem_addr(s_cd);
    asm volatile("" : "+r"(a_xy), "+r"(a_co), "+r"(a_cd));
    for (int r = 0; r < rounds; ++r, todo -= RENDER_BATCH) {
        if (__syncthreads_count(done) == CSB_TILE_PIX) break;
        const int progress = r * RENDER_BATCH + tid;
        if (range.x + progress < range.y) {
            const uint32_t id = point_list[range.x + progress];
            const float2 xy = means2D[id];
            const float4 co = conic_opacity[id];
            s_xy[tid] = xy;
            s_co[tid] = co;
            s_cd[tid] = rgb_depth[id];
            s_mask[tid] = patch_mask(xy, co, tile_x0, tile_y0);
        }
        __syncthreads();
        const int cnt = min(RENDER_BATCH, todo);
        const uint32_t base = (uint32_t)r * RENDER_BATCH;
        for (int g = 0; g < cnt; g += 32) {
            const uint32_t m = (g + lane < cnt) ? s_mask[g + lane] : 0u;
            uint32_t bits = __ballot_sync(0xFFFFFFFFu, (m >> warp) & 1u);
            while (bits) {
                const int j = g + __ffs(bits) - 1;
                bits &= bits - 1;
                if (done) continue;
                const float2 xy = lds_f2(a_xy + (uint32_t)j * 8u);
                const float dx = xy.x - pxf, dy = xy.y - pyf;
                const float4 con_o = lds_f4(a_co + (uint32_t)j * 16u);
                const float power = -0.5f * (con_o.x * dx * dx + con_o.z * dy * dy) - con_o.y * dx * dy;
                if (power > 0.0f) continue;
                // expf as in the reference (forward.cu:336): an SFU ex2 would be 7 instructions shorter, but alpha decides
                // the 1/255 and T < 1e-4 cut-offs, and a flipped decision moves a pixel by up to 4e-3 (measured: a
                // handful of pixels of a 1920x1080 frame) -- parity with the reference's decisions comes first
                const float alpha = fminf(0.99f, con_o.w * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) {
                    done = true;
                    continue;
                }
                const float4 cd = lds_f4(a_cd + (uint32_t)j * 16u);
                // (c * alpha) * T + C, two channels per packed instruction (same roundings as the scalar form)
                C01 = ffma2_bcast(T, fmul2_bcast(alpha, make_float2(cd.x, cd.y)), C01);
                C2D = ffma2_bcast(T, fmul2_bcast(alpha, make_float2(cd.z, cd.w)), C2D);
                T = test_T;
                last_contributor = base + (uint32_t)j + 1u;  // 1-based list position, skipped entries count
            }
            if (__all_sync(0xFFFFFFFFu, done)) break;
        }
    }
    if (inside) {
        const size_t hw = (size_t)vb.W * vb.H;
        const size_t pix = (size_t)py * vb.W + px;
        const float* __restrict__ bg = vb.background[v];
        final_T[v * hw + pix] = T;
        n_contrib[v * hw + pix] = last_contributor;
        out_color[(v * 3 + 0) * hw + pix] = C01.x + T * bg[0];
        out_color[(v * 3 + 1) * hw + pix] = C01.y + T * bg[1];
        out_color[(v * 3 + 2) * hw + pix] = C2D.x + T * bg[2];
        out_depth[v * hw + pix] = C2D.y;
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ vm,
                                    uint8_t* __restrict__ present) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float3 p = make_float3(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]);
    present[i] = xform_point_4x3(p, vm).z > 0.2f ? 1 : 0;
}

}  // namespace csb

// ----------------------------------------------------------------------------------
// host side (C-ABI)
// ----------------------------------------------------------------------------------
namespace csb {
int fill_view_batch(ViewBatch& vb, int num_views, const csb_view* views, int P, int width, int height) {
    if (num_views < 1 || num_views > CSB_MAX_VIEWS || views == nullptr || P < 0 || width < 1 || height < 1)
        return CSB_ERR_BAD_ARG;
    vb.V = num_views;
    vb.P = P;
    vb.W = width;
    vb.H = height;
    vb.gx = (width + CSB_TILE_X - 1) / CSB_TILE_X;
    vb.gy = (height + CSB_TILE_Y - 1) / CSB_TILE_Y;
    vb.tiles = vb.gx * vb.gy;
    vb.cull_exact = g_cull_exact;
    for (int v = 0; v < num_views; ++v) {
        // rasterizer_impl.cu:226-227
        vb.focal_y[v] = height / (2.0f * views[v].tan_fovy);
        vb.focal_x[v] = width / (2.0f * views[v].tan_fovx);
        vb.tan_fovx[v] = views[v].tan_fovx;
        vb.tan_fovy[v] = views[v].tan_fovy;
        vb.viewmatrix[v] = views[v].viewmatrix;
        vb.projmatrix[v] = views[v].projmatrix;
        vb.campos[v] = views[v].campos;
        vb.background[v] = views[v].background;
    }
    return 0;
}
}  // namespace csb

using namespace csb;

// renderCUDA forward: the multi-pixel-per-lane kernels (render_lanes.cuh) unless the debugging knob asks for the first version
static int launch_render_forward(cudaStream_t stream, dim3 tgrid, const ViewBatch& vb, const uint2* ranges,
                                 const uint32_t* point_list, const float2* means2D, const float4* conic_opacity,
                                 const float4* rgb_depth, float* final_T, uint32_t* n_contrib, float* out_color,
                                 float* out_depth, bool long_lists, const uint32_t* tile_order) {
    // 15 = by path (measured on B200, profiles/r2_summary.md): cloth-sized scenes on the per-tile binning path (small
    // footprints, ~30 entries per tile) run fastest with one pixel pair per lane (8x8 pixels per warp: finer culling), the
    // global sort path (C3 / C4: footprints of many tiles, thousands of entries per tile) with two pairs per lane
    int fv = g_render_variant & 15;
    if (fv == 15) fv = long_lists ? 4 : 3;
    switch (fv) {
        case 0:
            CSB_LAUNCH("render_forward", stream, launch_pdl(render_forward_kernel, tgrid, dim3(CSB_TILE_PIX), 0, stream, vb, ranges, point_list, means2D,
                                                            conic_opacity, rgb_depth, final_T, n_contrib, out_color, out_depth));
            break;
        case 3:
            CSB_LAUNCH("render_forward", stream, launch_pdl(render_forward_packed_kernel<1>, dim3(tgrid.x * tgrid.y * tgrid.z), dim3(LaneTiling<1>::THREADS), 0, stream, vb, ranges, point_list,
                                                            means2D, conic_opacity, rgb_depth, final_T, n_contrib, out_color, out_depth, tile_order));
            break;
        case 4:
            CSB_LAUNCH("render_forward", stream, launch_pdl(render_forward_packed_kernel<2>, dim3(tgrid.x * tgrid.y * tgrid.z), dim3(LaneTiling<2>::THREADS), 0, stream, vb, ranges, point_list,
                                                            means2D, conic_opacity, rgb_depth, final_T, n_contrib, out_color, out_depth, tile_order));
            break;
        case 2:
            CSB_LAUNCH("render_forward", stream, launch_pdl(render_forward_lanes_kernel<2>, tgrid, dim3(LaneTiling<2>::THREADS), 0, stream, vb, ranges, point_list,
                                                            means2D, conic_opacity, rgb_depth, final_T, n_contrib, out_color, out_depth));
            break;
        default:
            CSB_LAUNCH("render_forward", stream, launch_pdl(render_forward_lanes_kernel<1>, tgrid, dim3(LaneTiling<1>::THREADS), 0, stream, vb, ranges, point_list,
                                                            means2D, conic_opacity, rgb_depth, final_T, n_contrib, out_color, out_depth));
    }
    return 0;
}

extern "C" {

size_t csb_rasterize_geom_bytes(int num_views, int P) { return GeomState::carve(nullptr, num_views, P).bytes; }
size_t csb_rasterize_image_bytes(int num_views, int width, int height) {
    return ImageState::carve(nullptr, num_views, width, height).bytes;
}
size_t csb_rasterize_binning_bytes(int num_views, int P, int width, int height, size_t capacity) {
    return BinState::carve(nullptr, num_views, P, width, height, capacity).bytes;
}

int csb_rasterize_forward(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, int prefiltered, void* geom_ws,
                          size_t geom_bytes, void* image_ws, size_t image_bytes, void* binning_ws,
                          size_t binning_bytes, size_t binning_capacity, float* out_color, float* out_depth,
                          int* radii, uint32_t* num_rendered_dev, uint32_t* num_rendered_host, int flags,
                          void* stream_) {
    (void)prefiltered;  // the reference only uses it to trap on a culled point (auxiliary.h:155-159)
    cudaStream_t stream = (cudaStream_t)stream_;
    ViewBatch vb;
    int rc = fill_view_batch(vb, num_views, views, P, width, height);
    if (rc) return rc;
    if (P > 0) {  // P == 0 renders the background only (rasterize_points.cu:82)
        if (means3D == nullptr || opacities == nullptr || radii == nullptr) return CSB_ERR_BAD_ARG;
        if (shs == nullptr && colors_precomp == nullptr) return CSB_ERR_BAD_ARG;
        if (cov3D_precomp == nullptr && (scales == nullptr || rotations == nullptr)) return CSB_ERR_BAD_ARG;
        if (shs != nullptr && (M < (D + 1) * (D + 1) || D < 0 || D > 3)) return CSB_ERR_BAD_ARG;
    }
    if (!out_color || !out_depth || !num_rendered_dev) return CSB_ERR_BAD_ARG;
    if ((size_t)num_views * (size_t)P >= (1ull << 31) || binning_capacity >= (1ull << 30)) return CSB_ERR_UNSUPPORTED;

    GeomState g = GeomState::carve(geom_ws, num_views, P);
    ImageState im = ImageState::carve(image_ws, num_views, width, height);
    BinState b = BinState::carve(binning_ws, num_views, P, width, height, binning_capacity);
    if (geom_bytes < g.bytes || image_bytes < im.bytes || binning_bytes < b.bytes) return CSB_ERR_WORKSPACE;
    if (!geom_ws || !image_ws || (!binning_ws && binning_capacity > 0)) return CSB_ERR_BAD_ARG;

    const size_t n = (size_t)num_views * P;
    const size_t total_tiles = (size_t)num_views * vb.tiles;
    const bool early_count = (flags & 2) && num_rendered_host && n > 0;
    CSB_CUDA_TRY(cudaMemsetAsync(g.control, 0, g.control_words * sizeof(uint32_t), stream));
    CSB_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, total_tiles * sizeof(uint2), stream));
    CSB_CUDA_TRY(cudaMemsetAsync(im.tile_order, 0, TILE_ORDER_HEAD * sizeof(uint32_t), stream));  // "no order" + histogram + cursors
    // all memsets first: the kernels behind them form one programmatic-dependent-launch chain
    if (n > 0 && binning_capacity > 0) {
        // global path: look-back words of the tile sort; per-tile path (flags bit2): the status words of its scan
        const size_t words = (flags & 4) ? (total_tiles + 255) / 256 : b.lookback_words;
        if (words > b.lookback_words) return CSB_ERR_UNSUPPORTED;
        CSB_CUDA_TRY(cudaMemsetAsync(b.lookback, 0, words * sizeof(uint32_t), stream));
    }
    dim3 tgrid(vb.gx, vb.gy, num_views);
    if (n > 0) {
        const unsigned blocks = (unsigned)((n + 255) / 256);
        const size_t rec_smem = (size_t)256 * (M * 3 + 1) * sizeof(float);
        const int stage_sh = (shs != nullptr && colors_precomp == nullptr && rec_smem <= 56 * 1024) ? 1 : 0;
        {
            static unsigned long long configured = 0;
            CSB_CUDA_TRY(opt_in_dynamic_smem(preprocess_kernel, 56 * 1024, &configured));
        }
        CSB_LAUNCH("preprocess", stream, launch_pdl(preprocess_kernel, dim3(blocks), dim3(256), stage_sh ? rec_smem : 0, stream, vb, D, M, means3D, scales, scale_modifier, rotations, opacities,
                                                      shs, cov3D_precomp, colors_precomp, radii, g, stage_sh,
                                                      ((flags & 4) && binning_capacity > 0) ? im.ranges : (uint2*)nullptr));
        if ((flags & 4) && binning_capacity > 0) {
            // short tile lists: count -> scan -> fill -> per-tile sort (see bin_entries_kernel)
            if (total_tiles > 256u * 4096u) return CSB_ERR_UNSUPPORTED;
            const bool side = sorted_side(total_tiles) != 0;
            uint32_t* fill_d = side ? b.tiles_a : b.tiles_b;  // unsorted fill goes to the side the consumers do not read
            uint32_t* fill_v = side ? b.vals_a : b.vals_b;
            uint32_t* st = side ? b.tiles_b : b.tiles_a;
            uint32_t* sv = side ? b.vals_b : b.vals_a;
            uint32_t* mapped = early_count ? num_rendered_host : nullptr;
            const uint32_t sblocks_scan = (uint32_t)((total_tiles + 255) / 256);
            uint32_t* groups = b.lookback;  // status words of the chained scan: free in this path; zeroed above
            // (the per-tile counts were made by preprocess_kernel)
            // (a single 1024-thread block scanning all tiles of a small frame was measured slower than this chained scan over
            // 40 blocks: 19 vs 11 us)
            CSB_LAUNCH("bin_ranges", stream, launch_pdl(bin_ranges_kernel, dim3(sblocks_scan), dim3(256), 0, stream, im.ranges, groups,
                                                         (uint32_t)total_tiles, binning_capacity, g.control, num_rendered_dev, mapped, im.tile_order));
            CSB_LAUNCH("bin_fill", stream, launch_pdl(bin_entries_kernel<true>, dim3(blocks + sblocks_scan), dim3(256), 0, stream, vb, g, im.ranges, groups, fill_d, fill_v,
                                                       im.tile_order));
            const unsigned sblocks = (unsigned)std::min<size_t>(total_tiles, 148 * 16);
            CSB_LAUNCH("bin_sort", stream, launch_pdl(bin_sort_kernel, dim3(sblocks), dim3(256), 0, stream, im.ranges, (const uint32_t*)g.control,
                                                       (uint32_t)total_tiles, (const uint32_t*)fill_d, (const uint32_t*)fill_v, st, sv,
                                                       (flags & 1) ? b.keys : nullptr, (const uint32_t*)im.tile_order));
            rc = launch_render_forward(stream, tgrid, vb, im.ranges, sv, g.means2D, g.conic_opacity, g.rgb_depth, im.final_T, im.n_contrib, out_color, out_depth, false, im.tile_order);
            if (rc) return rc;
            if (num_rendered_host && !early_count)
                CSB_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, num_rendered_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                             stream));
            return 0;
        }
        // depth order of the (view, Gaussian) entries
        uint32_t* scan_state = g.control + CTRL_SCAN_STATE;
        uint32_t* depth_lookback = scan_state + 2 * (blocks + 1);
        int side = 0;
        rc = radix_sort_pairs(g.depth_keys, g.keys_b, g.order_a, g.order_b, g.control + CTRL_N_ENTRIES, n, 32,
                              g.control + CTRL_HIST_DEPTH, g.control + CTRL_TICKET_DEPTH, depth_lookback, &side,
                              stream, /*hist_ready=*/true, "sort_depth");
        if (rc) return rc;
        const uint32_t* order = side ? g.order_b : g.order_a;
        const int tile_passes = (higher_msb((uint32_t)total_tiles) + 7) / 8;
        const bool tile_hist_fused = tile_passes <= 3;  // emit_kernel keeps 3 x 256 bins
        CSB_LAUNCH("emit", stream, launch_pdl(emit_kernel, dim3(blocks), dim3(256), 0, stream, vb, order, g, (uint64_t*)scan_state, b.tiles_a, b.vals_a,
                                                binning_capacity, num_rendered_dev, early_count ? num_rendered_host : nullptr,
                                                tile_hist_fused ? tile_passes : 0));
        if (binning_capacity > 0) {
            const int tile_bits = higher_msb((uint32_t)total_tiles);
            int tside = 0;
            rc = radix_sort_pairs(b.tiles_a, b.tiles_b, b.vals_a, b.vals_b, g.control + CTRL_NUM_RENDERED,
                                  binning_capacity, tile_bits, g.control + CTRL_HIST_TILE,
                                  g.control + CTRL_TICKET_TILE, b.lookback, &tside, stream, /*hist_ready=*/tile_hist_fused, "sort_tile");
            if (rc) return rc;
            // the final order always lives in *_a for the consumers: copy back if it ended in *_b
            const uint32_t* st = tside ? b.tiles_b : b.tiles_a;
            const uint32_t* sv = tside ? b.vals_b : b.vals_a;
            int rblocks = (int)((binning_capacity + 255) / 256);
            if (rblocks > 148 * 16) rblocks = 148 * 16;
            CSB_LAUNCH("tile_ranges", stream, launch_pdl(tile_ranges_kernel, dim3(rblocks), dim3(256), 0, stream, g.control + CTRL_NUM_RENDERED, st, sv, g.rgb_depth,
                                                            im.ranges, (flags & 1) ? b.keys : nullptr,
                                                            (uint32_t)vb.tiles));
            {   // longest list first (build_tile_order): the big tiles of a view set start first, the small ones fill the tail
                const unsigned oblocks = (unsigned)((total_tiles + 255) / 256);
                CSB_LAUNCH("tile_order", stream, launch_pdl(tile_order_hist_kernel, dim3(oblocks), dim3(256), 0, stream, (const uint2*)im.ranges, (uint32_t)total_tiles, im.tile_order));
                CSB_LAUNCH("tile_order", stream, launch_pdl(tile_order_place_kernel, dim3(oblocks), dim3(256), 0, stream, (const uint2*)im.ranges, (uint32_t)total_tiles, im.tile_order));
            }
            rc = launch_render_forward(stream, tgrid, vb, im.ranges, sv, g.means2D, g.conic_opacity, g.rgb_depth, im.final_T, im.n_contrib, out_color, out_depth, true, im.tile_order);
            if (rc) return rc;
            if (tside != sorted_side(total_tiles)) return CSB_ERR_UNSUPPORTED;  // consumers recompute the side
        }
    } else {
        CSB_CUDA_TRY(cudaMemsetAsync(num_rendered_dev, 0, sizeof(uint32_t), stream));
        rc = launch_render_forward(stream, tgrid, vb, im.ranges, b.vals_a, g.means2D, g.conic_opacity, g.rgb_depth, im.final_T, im.n_contrib, out_color, out_depth, true, im.tile_order);
        if (rc) return rc;
    }
    if (num_rendered_host && !early_count)
        CSB_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, num_rendered_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                     stream));
    return 0;
}

int csb_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     void* stream_) {
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return CSB_ERR_BAD_ARG;
    if (P == 0) return 0;
    CSB_LAUNCH("mark_visible", (cudaStream_t)stream_, mark_visible_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, means3D, viewmatrix, present));
    return 0;
}

}  // extern "C"
