// Hand-written onesweep LSD radix sort of (u32 key, u32 value) pairs for sm_100a.
//
// Replaces cub::DeviceRadixSort::SortPairs at RAST/cuda_rasterizer/rasterizer_impl.cu:304-309
// (and KNN/simple_knn.cu:210-213).  The rasterizer does not sort 64-bit tile|depth keys in one
// go: Gaussians are ordered by depth once (P elements, 32 key bits), instances are emitted in
// that order and then only the tile bits (12-18) are sorted.  Because every pass is stable the
// final order equals the reference's single 64-bit sort, at a fraction of the traffic
// ("binning").
//
// One upfront histogram kernel counts every pass' digits; each pass is ONE kernel: a tile of
// 4096 pairs is ranked with warp match-any + per-warp digit counters, the tile's digit counts are
// published and the global prefix is obtained by decoupled look-back over earlier tiles.
#include <stdlib.h>

#include "common.cuh"

namespace csb {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;

constexpr uint32_t LB_FLAG_AGG = 1u << 30;
constexpr uint32_t LB_FLAG_INC = 2u << 30;
constexpr uint32_t LB_VALUE_MASK = (1u << 30) - 1;

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- histogram of all passes -------------------------------------------------------
__global__ void __launch_bounds__(SORT_THREADS)
sort_histogram_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ n_dev, int passes,
                      uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[4 * RADIX];
    const uint32_t n = *n_dev;
    for (int i = threadIdx.x; i < passes * RADIX; i += SORT_THREADS) sh[i] = 0;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * SORT_THREADS;
    const uint32_t lane = threadIdx.x & 31;
    // warp-uniform trip count: warp_hist_add is warp-collective
    for (size_t base = (size_t)blockIdx.x * SORT_THREADS + (threadIdx.x & ~31u); base < n; base += stride) {
        const size_t i = base + lane;
        const bool ok = i < n;
        warp_hist_add(sh, passes, ok ? keys[i] : 0u, ok, (int)lane);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RADIX; i += SORT_THREADS) {
        const uint32_t c = sh[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// ---- one onesweep pass -------------------------------------------------------------
template <int SORT_ITEMS, int MINB>  // pairs per thread: 16 (4096-pair tiles) or 4 (1024-pair tiles, small inputs); resident blocks asked for
__global__ void __launch_bounds__(SORT_THREADS, MINB)
sort_onesweep_kernel(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                     const uint32_t* __restrict__ n_dev, int shift, const uint32_t* __restrict__ hist,
                     uint32_t* __restrict__ ticket, uint32_t* __restrict__ lookback) {
    __shared__ uint32_t warp_cnt[SORT_WARPS][RADIX];  // per-warp digit counters -> exclusive warp offsets
    __shared__ uint32_t digit_start[RADIX];           // first slot of a digit inside the sorted tile
    __shared__ uint32_t global_base[RADIX];           // destination of the digit's run minus its first tile slot
    __shared__ uint32_t scan_tmp[SORT_WARPS];
    constexpr int TILE = SORT_ITEMS * SORT_THREADS;
    __shared__ uint32_t s_keys[TILE];
    __shared__ uint32_t s_vals[TILE];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    grid_dep_wait();
    grid_dep_launch();
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t n = *n_dev;
    const size_t tile_start = (size_t)tile * TILE;
    if (tile_start >= n) return;
    const uint32_t n_valid = (uint32_t)min((size_t)TILE, (size_t)n - tile_start);

    // a pass in which every key has the same digit is the identity permutation: plain copy
    // (depth exponents and unused tile bits produce such passes)
    {
        // hist[d] == n for one d  <=>  max over d == n.  Checked by every thread on its own digit.
        const int single = __syncthreads_or(hist[tid] == n);
        if (single) {
            for (uint32_t i = tid; i < n_valid; i += SORT_THREADS) {
                keys_out[tile_start + i] = keys_in[tile_start + i];
                vals_out[tile_start + i] = vals_in[tile_start + i];
            }
            return;
        }
    }

    // warp-striped load: item i of lane l is element warp*(32*SORT_ITEMS) + i*32 + l of the tile
    uint32_t key[SORT_ITEMS];
    uint16_t rank[SORT_ITEMS];
    const uint32_t wbase = warp * (SORT_ITEMS * 32) + lane;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t e = wbase + i * 32;
        const bool ok = e < n_valid;
        key[i] = ok ? keys_in[tile_start + e] : 0xFFFFFFFFu;
    }
    // stable ranking inside the warp
    const uint32_t lt_mask = (1u << lane) - 1;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t d = (key[i] >> shift) & (RADIX - 1);
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t pre = warp_cnt[warp][d];
        __syncwarp();
        rank[i] = (uint16_t)(pre + __popc(peers & lt_mask));
        if ((peers & lt_mask) == 0) warp_cnt[warp][d] = pre + __popc(peers);
        __syncwarp();
    }
    __syncthreads();

    // thread d owns digit d: exclusive scan of the warps' counts, tile total
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) {
        const uint32_t c = warp_cnt[w][tid];
        warp_cnt[w][tid] = count;
        count += c;
    }
    // publish this tile's digit count, then look back for the prefix over earlier tiles
    uint32_t* my_status = lookback + (size_t)tile * RADIX + tid;
    if (tile == 0) {
        st_volatile_u32(my_status, LB_FLAG_INC | count);
    } else {
        st_volatile_u32(my_status, LB_FLAG_AGG | count);
    }
    // exclusive scan over digits: of the global histogram (digit base) and of the tile counts
    uint32_t hsum = hist[tid], tsum = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xFFFFFFFFu, hsum, o);
        const uint32_t b = __shfl_up_sync(0xFFFFFFFFu, tsum, o);
        if (lane >= o) { hsum += a; tsum += b; }
    }
    if (lane == 31) scan_tmp[warp] = hsum;
    __syncthreads();
    uint32_t hoff = 0;
    for (int w = 0; w < warp; ++w) hoff += scan_tmp[w];
    __syncthreads();
    if (lane == 31) scan_tmp[warp] = tsum;
    __syncthreads();
    uint32_t toff = 0;
    for (int w = 0; w < warp; ++w) toff += scan_tmp[w];
    const uint32_t hist_excl = hoff + hsum - hist[tid];
    const uint32_t tile_excl = toff + tsum - count;

    uint32_t prefix = 0;
    if (tile > 0) {
        // Look-back, LB_BATCH predecessors per round trip: the status loads of a batch are independent and in
        // flight together (a one-at-a-time walk costs an L2 latency per predecessor, and with ~700 tiles
        // resident the first wave walks hundreds of them).
        constexpr int LB_BATCH = 8;
        int t = (int)tile - 1;
        bool done = false;
        while (!done) {
            uint32_t s[LB_BATCH];
#pragma unroll
            for (int b = 0; b < LB_BATCH; ++b)
                s[b] = (t - b) >= 0 ? ld_volatile_u32(lookback + (size_t)(t - b) * RADIX + tid) : LB_FLAG_INC;
            int used = 0;
#pragma unroll
            for (int b = 0; b < LB_BATCH; ++b) {
                if (!done && used == b) {
                    const uint32_t f = s[b] & ~LB_VALUE_MASK;
                    if (f != 0) {  // published (an unpublished predecessor holds an earlier ticket: it runs)
                        prefix += s[b] & LB_VALUE_MASK;
                        used = b + 1;
                        if (f == LB_FLAG_INC) done = true;
                    }
                }
            }
            t -= used;
        }
        st_volatile_u32(my_status, LB_FLAG_INC | (prefix + count));
    }
    digit_start[tid] = tile_excl;
    global_base[tid] = hist_excl + prefix - tile_excl;  // destination of tile slot p of this digit = base + p (wraps)
    __syncthreads();

    // scatter into tile-sorted order in shared memory (the values are only loaded now: 16 fewer live
    // registers while ranking and looking back, the loads overlap other blocks' work)
    uint32_t val[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t e = wbase + i * 32;
        val[i] = e < n_valid ? vals_in[tile_start + e] : 0u;
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t d = (key[i] >> shift) & (RADIX - 1);
        const uint32_t pos = digit_start[d] + warp_cnt[warp][d] + rank[i];
        s_keys[pos] = key[i];
        s_vals[pos] = val[i];
    }
    __syncthreads();
    // coalesced runs per digit to their final place
    for (uint32_t p = tid; p < n_valid; p += SORT_THREADS) {
        const uint32_t k = s_keys[p];
        const uint32_t d = (k >> shift) & (RADIX - 1);
        const uint32_t dst = global_base[d] + p;
        keys_out[dst] = k;
        vals_out[dst] = s_vals[p];
    }
}

int radix_sort_pairs(uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                     const uint32_t* n_dev, size_t n_max, int end_bit, uint32_t* hist, uint32_t* tickets,
                     uint32_t* lookback, int* side_out, cudaStream_t stream, bool hist_ready, const char* launch_name) {
    if (end_bit < 1) end_bit = 1;
    if (end_bit > 32) return CSB_ERR_BAD_ARG;
    const int passes = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
    *side_out = passes & 1;
    if (n_max == 0) { *side_out = 0; return 0; }
    const size_t tile = sort_tile_size(n_max);
    const size_t tiles = (n_max + tile - 1) / tile;
    // Big inputs (the tile sort of C3 / C4: tens of millions of pairs) run with four resident blocks per SM (64 registers,
    // a 32-byte spill) instead of three: 486 -> 456 us for C3's two tile passes, 1745 -> 1642 us at C4; the small depth
    // sorts (<= a few million keys, latency-bound look-back chains) are faster with three.  CSB_SORT_MINB overrides (A/B).
    static const int forced = [] { const char* e = getenv("CSB_SORT_MINB"); return e ? atoi(e) : 0; }();
    const int minb = forced ? forced : (n_max >= ((size_t)8 << 20) ? 4 : 3);
    auto kernel = tile == CSB_SORT_TILE ? (minb >= 4 ? sort_onesweep_kernel<CSB_SORT_TILE / SORT_THREADS, 4>
                                                     : sort_onesweep_kernel<CSB_SORT_TILE / SORT_THREADS, 3>)
                                        : sort_onesweep_kernel<CSB_SORT_TILE_SMALL / SORT_THREADS, 3>;
    int hist_blocks = (int)((n_max + SORT_THREADS * 16 - 1) / (SORT_THREADS * 16));
    if (hist_blocks > 148 * 8) hist_blocks = 148 * 8;
    if (hist_blocks < 1) hist_blocks = 1;
    if (!hist_ready)
        CSB_LAUNCH("sort_histogram", stream, sort_histogram_kernel<<<hist_blocks, SORT_THREADS, 0, stream>>>(keys_a, n_dev, passes, hist));
    uint32_t *kin = keys_a, *kout = keys_b, *vin = vals_a, *vout = vals_b;
    for (int p = 0; p < passes; ++p) {
        CSB_LAUNCH(launch_name ? launch_name : "sort_onesweep", stream, launch_pdl(kernel, dim3((unsigned)tiles), dim3(SORT_THREADS), 0, stream, 
            kin, kout, vin, vout, n_dev, p * RADIX_BITS, hist + p * RADIX, tickets + p,
            lookback + (size_t)p * (tiles + 1) * RADIX));
        uint32_t* t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    return 0;
}

}  // namespace csb
