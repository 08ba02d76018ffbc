// meshnet GNN forward (encode - process - decode) for sm_100a: bf16 tcgen05 MLPs with TMEM accumulators.
//
// Replaces EncodeProcessDecode.forward (meshnet/graph_network.py:387-408) and what it runs per call:
// Encoder (:98-111), 15 x InteractionNetwork (message :178-199 = edge MLP over cat[x[dst], x[src], e];
// aggregate = sum at dst :136; update :201-222 = node MLP over cat[agg, x]; residuals :176) and the
// Decoder (:323-332), i.e. ~300 cuBLAS / index_select / scatter_add_ / layer_norm launches of PyTorch.
//
// One kernel template, four roles.  A CTA owns tiles of 128 rows (edges or nodes).  The role's weights
// (a "unit" image: bf16, already laid out in the 128-byte-swizzled K-major form tcgen05.mma reads)
// are bulk-copied (cp.async.bulk, TMA engine) into shared memory once per CTA and stay resident while
// the CTA walks its tiles.  Per tile:
//     stage A operand (gather) -> tcgen05.mma chain -> TMEM -> registers (+bias, ReLU) -> bf16 A operand
//     of the next layer -> ... -> LayerNorm in registers -> role specific output.
//   ENC   first Linear (K = 8 / 4) on the CUDA cores in fp32, two tensor-core Linears, LN
//   EDGE  gather x[dst] | x[src] | e0 (K = 384), three Linears, LN, segmented sum over the dst-sorted rows
//         of the tile, one red.add per (segment, column) into agg
//   NODE  agg | x (K = 256), three Linears, LN, x += .  (fp32 master copy + bf16 copy for the gathers)
//   DEC   two tensor-core Linears, last Linear (N = 3) on the CUDA cores in fp32
// The reference never replaces the edge latents by the edge MLP output: e_m = 2^m e_0 (SURVEY.md fact
// 0.3).  e_0 is therefore stored once (bf16, in dst-sorted edge order) and the factor 2^m is folded into
// the edge block of the first edge Linear when the weights are packed (exact in bf16).

#include <type_traits>

#include "gnn_common.cuh"

namespace csb {
namespace gnn {

// Debugging aid (csb_debug_gnn_trace): thread 0 of ONE chosen CTA stamps %globaltimer at the stations of its tile walk
// into a device buffer: word 0 = count, then (time_ns << 8 | station).  A null buffer (the default) costs one load per station.
__device__ unsigned long long* g_trace_buf = nullptr;
__device__ int g_trace_block = 0;
__device__ __forceinline__ void trace_pt(int station) {
    if (threadIdx.x == 0 && (int)blockIdx.x == g_trace_block) {
        unsigned long long* p = g_trace_buf;
        if (p) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            const unsigned long long n = atomicAdd(p, 1ull);
            if (n < 8000) p[1 + n] = (t << 8) | (unsigned long long)station;
        }
    }
}

// W [128, K] row-major fp32 -> chunks of [128 rows x 64 k] bf16, swizzle-128B K-major
__device__ void pack_linear(const float* __restrict__ W, int K, int col_scale_from, float scale, uint8_t* dst) {
    for (int i = threadIdx.x; i < LAT * K; i += blockDim.x) {
        const int n = i / K, k = i - n * K;
        float w = W[i];
        if (k >= col_scale_from) w *= scale;
        const uint32_t off = (uint32_t)(k >> 6) * CHUNK_BYTES + swz_off(n, (k & 63) >> 3) + (k & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(dst + off) = __float2bfloat16_rn(w);
    }
}
__device__ void copy_f32(const float* __restrict__ s, float* d, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
}

__global__ void __launch_bounds__(256) gnn_pack_kernel(const float* __restrict__ flat, uint8_t* __restrict__ packed,
                                                        const __grid_constant__ PackTable tab) {
    const UnitDesc& u = tab.u[blockIdx.x];
    const float* W0 = flat + u.src;
    const float* b0 = W0 + (size_t)LAT * u.k0;
    const float* W1 = b0 + LAT;
    const float* b1 = W1 + LAT * LAT;
    const float* W2 = b1 + LAT;
    const float* b2 = W2 + (size_t)u.n2 * LAT;
    const float* lng = b2 + u.n2;
    const float* lnb = lng + LAT;
    uint8_t* d = packed + u.dst;
    float* f = reinterpret_cast<float*>(d + (size_t)unit_chunks(u.mode) * CHUNK_BYTES);
    if (u.mode == EDGE || u.mode == NODE) {
        const int c0 = u.k0 / 64;
        pack_linear(W0, u.k0, u.mode == EDGE ? 2 * LAT : u.k0, u.e_scale, d);
        pack_linear(W1, LAT, LAT, 1.f, d + (size_t)c0 * CHUNK_BYTES);
        pack_linear(W2, LAT, LAT, 1.f, d + (size_t)(c0 + 2) * CHUNK_BYTES);
        copy_f32(b0, f, LAT); copy_f32(b1, f + LAT, LAT); copy_f32(b2, f + 2 * LAT, LAT);
        copy_f32(lng, f + 3 * LAT, LAT); copy_f32(lnb, f + 4 * LAT, LAT);
    } else if (u.mode == ENC) {
        pack_linear(W1, LAT, LAT, 1.f, d);
        pack_linear(W2, LAT, LAT, 1.f, d + 2 * CHUNK_BYTES);
        copy_f32(b0, f, LAT); copy_f32(b1, f + LAT, LAT); copy_f32(b2, f + 2 * LAT, LAT);
        copy_f32(lng, f + 3 * LAT, LAT); copy_f32(lnb, f + 4 * LAT, LAT);
        copy_f32(W0, f + 5 * LAT, LAT * u.k0);
    } else {
        pack_linear(W0, LAT, LAT, 1.f, d);
        pack_linear(W1, LAT, LAT, 1.f, d + 2 * CHUNK_BYTES);
        copy_f32(b0, f, LAT); copy_f32(b1, f + LAT, LAT);
        copy_f32(W2, f + 2 * LAT, u.n2 * LAT);
        copy_f32(b2, f + 2 * LAT + u.n2 * LAT, u.n2);
    }
}

// ---- kernel arguments ---------------------------------------------------------------------------
struct Args {
    const uint8_t* unit;  // packed unit image
    int rows;             // E (ENC edges, EDGE) or N
    int kin, nout;        // ENC: input features; DEC: output features
    int N;
    const int* dst_sorted;           // EDGE [E]
    const int* src_sorted;           // EDGE [E]
    const __nv_bfloat16* x_bf16;     // EDGE / DEC gathers, NODE input  [N,128]
    const __nv_bfloat16* e0;         // EDGE [E,128] (dst-sorted order)
    float* agg;                      // EDGE out (atomic), NODE in (then zeroed)  [N,128]
    float* x_f32;                    // NODE in/out, ENC(node) out  [N,128]
    __nv_bfloat16* out_bf16;         // NODE / ENC out  [rows,128]
    const float* in;                 // ENC input [rows_src, kin]
    const int* perm;                 // ENC: source row of output row i (NULL = identity)
    float* y;                        // DEC out [N, nout]
    __nv_bfloat16* agg_save;         // NODE (training): bf16 copy of the aggregated messages it consumed [N,128]
};

template <int MODE> struct Cfg {
    static constexpr int K1_CHUNKS = MODE == EDGE ? 6 : MODE == NODE ? 4 : 2;  // K blocks of the first TC layer
    static constexpr int A_SLOTS = MODE == EDGE ? 3 : MODE == NODE ? 4 : 2;    // resident A chunks
    static constexpr int TC_LAYERS = (MODE == EDGE || MODE == NODE) ? 3 : 2;
    static constexpr int W_CHUNKS = MODE == EDGE ? 10 : MODE == NODE ? 8 : 4;
};

constexpr int SCRATCH_FLOATS = 2048;  // LayerNorm / decoder partials
static size_t smem_bytes(int mode, int kin, int nout) {
    const int a_slots = mode == EDGE ? 3 : mode == NODE ? 4 : 2;
    size_t b = 1024;  // alignment slack
    b += (size_t)unit_chunks(mode) * CHUNK_BYTES;
    b += (size_t)a_slots * CHUNK_BYTES;
    b += ((size_t)unit_f32(mode, kin, nout) * 4 + 127) / 128 * 128;
    b += SCRATCH_FLOATS * 4 + 2 * TILE_M * 4 + 64;
    return b;
}

// What a CTA carries from one role / tile walk to the next: the weight barrier's phase (the unit image is bulk-copied once
// per walk), the MMA barrier's parity.  Identical in all threads.
struct WalkState {
    uint32_t w_phase = 0, mma_parity = 0;
    bool weights_ready = true;
};

template <int MODE> struct Carve {
    using C = Cfg<MODE>;
    uint8_t *sW, *sA;
    float *sF, *sRed;
    int *sDst, *sSrc;
    __device__ Carve(uint8_t* smem, int nf32) {
        sW = smem;                                               // weight chunks
        sA = sW + (size_t)C::W_CHUNKS * CHUNK_BYTES;            // A operand chunks (also scatter staging)
        sF = reinterpret_cast<float*>(sA + (size_t)C::A_SLOTS * CHUNK_BYTES);  // fp32 parameters
        sRed = sF + ((nf32 + 31) & ~31);                         // SCRATCH_FLOATS
        sDst = reinterpret_cast<int*>(sRed + SCRATCH_FLOATS);    // [128]
        sSrc = sDst + TILE_M;                                    // [128]
    }
};

// Loads that see what OTHER CTAs of the same (persistent) kernel wrote after a grid barrier: L2 only, no L1 line that an
// earlier phase may have left behind.  Between kernel launches (COH = false) plain loads are coherent.
template <bool COH> __device__ __forceinline__ uint4 ld_act(const uint4* p) { return COH ? __ldcg(p) : *p; }
template <bool COH> __device__ __forceinline__ float4 ld_act(const float4* p) { return COH ? __ldcg(p) : *p; }

// the role's resident weights: one bulk copy per 16 KB chunk + the fp32 block (one thread)
template <int MODE>
__device__ __forceinline__ void issue_weights(const Args& a, uint8_t* smem, uint32_t bar_w, WalkState& ws) {
    using C = Cfg<MODE>;
    const int nf32 = unit_f32(MODE, a.kin, a.nout);
    Carve<MODE> cv(smem, nf32);
    if (threadIdx.x == 0) {
        const uint32_t wbytes = (uint32_t)C::W_CHUNKS * CHUNK_BYTES;
        const uint32_t fbytes = (uint32_t)nf32 * 4;
        mbar_expect_tx(bar_w, wbytes + fbytes);
        for (int c = 0; c < C::W_CHUNKS; ++c)
            bulk_g2s(smem_u32(cv.sW + (size_t)c * CHUNK_BYTES), a.unit + (size_t)c * CHUNK_BYTES, CHUNK_BYTES, bar_w);
        bulk_g2s(smem_u32(cv.sF), a.unit + wbytes, fbytes, bar_w);
    }
    ws.weights_ready = false;
}

// The tile walk of a role: tiles tile0, tile0 + tile_step, ... of a.rows rows.
// Row indices of an EDGE tile, fetched by threads 0..127 BEFORE the walk waits for its predecessor (the graph is constant:
// the load's latency hides behind the dependency wait instead of heading the tile's chain).
struct EdgeRows {
    int d = -1, s = -1;
    bool valid = false;
};
__device__ __forceinline__ EdgeRows load_edge_rows(const Args& a, int tile) {
    EdgeRows r;
    if (threadIdx.x < TILE_M) {
        const int p = tile * TILE_M + threadIdx.x;
        if (p < a.rows) {
            r.d = a.dst_sorted[p];
            r.s = a.src_sorted[p];
            if ((unsigned)r.d >= (unsigned)a.N || (unsigned)r.s >= (unsigned)a.N) r.d = r.s = -1;
        }
    }
    r.valid = true;
    return r;
}

template <int MODE, bool COH, bool TRACE>
__device__ __forceinline__ void mlp_tiles(const Args& a, uint8_t* smem, const uint32_t tmem, const uint32_t bar_w,
                                          const uint32_t bar_mma, WalkState& ws, const int tile0, const int tile_step,
                                          const EdgeRows pre = EdgeRows()) {
    using C = Cfg<MODE>;
    const int nf32 = unit_f32(MODE, a.kin, a.nout);
    Carve<MODE> cv(smem, nf32);
    uint8_t* sW = cv.sW;
    uint8_t* sA = cv.sA;
    float* sF = cv.sF;
    float* sRed = cv.sRed;
    int* sDst = cv.sDst;
    int* sSrc = cv.sSrc;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int warp_u = __shfl_sync(0xFFFFFFFFu, warp, 0);  // the same value, known to be warp-uniform
    const int row = (warp & 3) * 32 + lane;  // TMEM lane == tile row owned in the epilogues
    const int half = warp >> 2;              // columns [64*half, 64*half + 64)
    bool weights_ready = ws.weights_ready;
    uint32_t mma_parity = ws.mma_parity;
    const uint32_t sA_u32 = smem_u32(sA), sW_u32 = smem_u32(sW);
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(half * 64);

    const int num_tiles = (a.rows + TILE_M - 1) / TILE_M;
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        const int row0 = tile * TILE_M;
        const int grow = row0 + row;  // global row of this thread's epilogue row
        const bool rvalid = grow < a.rows;
        if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(1);

        // ------------------------------------------------------------------ stage the first A operand
        if (MODE == EDGE) {
            const EdgeRows er = (pre.valid && tile == tile0) ? pre : load_edge_rows(a, tile);
            if (tid < TILE_M) {
                sDst[tid] = er.d;
                sSrc[tid] = er.s;
            }
            __syncthreads();
        }
        int wchunk = 0;  // next weight chunk to multiply with
        if (MODE == ENC) {
            // first Linear on the CUDA cores (fp32), K = kin
            if (!weights_ready) { mbar_wait(bar_w, ws.w_phase); ws.w_phase ^= 1; weights_ready = true; }
            const float* W0 = sF + 5 * LAT;
            const float* b0 = sF;
            float xin[16];
            const int kin = a.kin;
            const long long srow = rvalid ? (a.perm ? (long long)a.perm[grow] : (long long)grow) : -1;
#pragma unroll
            for (int k = 0; k < 16; ++k) xin[k] = (srow >= 0 && k < kin) ? a.in[srow * kin + k] : 0.f;
#pragma unroll 1
            for (int c = 0; c < 8; ++c) {
                float h[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int o = half * 64 + c * 8 + j;
                    float acc = b0[o];
#pragma unroll
                    for (int k = 0; k < 16; ++k)
                        if (k < kin) acc = fmaf(W0[o * kin + k], xin[k], acc);
                    h[j] = fmaxf(acc, 0.f);
                }
                uint4 pk = make_uint4(pack_bf16(h[0], h[1]), pack_bf16(h[2], h[3]), pack_bf16(h[4], h[5]),
                                      pack_bf16(h[6], h[7]));
                *reinterpret_cast<uint4*>(sA + (size_t)half * CHUNK_BYTES + swz_off(row, c)) = pk;
            }
        } else if (MODE == DEC) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                const int n = row0 + r;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (n < a.rows) v = *reinterpret_cast<const uint4*>(a.x_bf16 + (size_t)n * LAT + kb * 64 + c * 8);
                *reinterpret_cast<uint4*>(sA + (size_t)kb * CHUNK_BYTES + swz_off(r, c)) = v;
            }
        } else if (MODE == NODE) {
            // all 24 loads of the thread in flight before the first use (one memory latency per tile)
            float4 fa[16];
            uint4 xb[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                const int n = row0 + r;
                fa[2 * it] = fa[2 * it + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
                xb[it] = make_uint4(0, 0, 0, 0);
                if (n < a.rows) {
                    const float4* p = reinterpret_cast<const float4*>(a.agg + (size_t)n * LAT + kb * 64 + c * 8);
                    fa[2 * it] = ld_act<COH>(p);
                    fa[2 * it + 1] = ld_act<COH>(p + 1);
                    xb[it] = *reinterpret_cast<const uint4*>(a.x_bf16 + (size_t)n * LAT + kb * 64 + c * 8);
                }
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                const int n = row0 + r;
                const float4 f0 = fa[2 * it], f1 = fa[2 * it + 1];
                const uint4 v = make_uint4(pack_bf16(f0.x, f0.y), pack_bf16(f0.z, f0.w), pack_bf16(f1.x, f1.y),
                                           pack_bf16(f1.z, f1.w));
                *reinterpret_cast<uint4*>(sA + (size_t)kb * CHUNK_BYTES + swz_off(r, c)) = v;
                *reinterpret_cast<uint4*>(sA + (size_t)(kb + 2) * CHUNK_BYTES + swz_off(r, c)) = xb[it];
                if (n < a.rows && a.agg_save)  // the exact operand, kept for the backward's recomputation
                    *reinterpret_cast<uint4*>(a.agg_save + (size_t)n * LAT + kb * 64 + c * 8) = v;
                if (n < a.rows) {  // agg is ready for the next step's scatter
                    float4* p = reinterpret_cast<float4*>(a.agg + (size_t)n * LAT + kb * 64 + c * 8);
                    p[0] = make_float4(0.f, 0.f, 0.f, 0.f);
                    p[1] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            // pull the next tile of this CTA towards L2 while this one is multiplied
            {
                const int nt = tile + tile_step;
                if (nt < num_tiles) {
                    const int r = tid >> 1, n = nt * TILE_M + r;
                    if (n < a.rows) {
                        const char* pa = reinterpret_cast<const char*>(a.agg + (size_t)n * LAT) + (tid & 1) * 256;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pa));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + 128));
                        const char* px = reinterpret_cast<const char*>(a.x_bf16 + (size_t)n * LAT) + (tid & 1) * 128;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(px));
                        const char* pf = reinterpret_cast<const char*>(a.x_f32 + (size_t)n * LAT) + (tid & 1) * 256;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pf));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + 128));
                    }
                }
            }
        }
        if (MODE == EDGE) {
            // next tile of this CTA: its rows' x[dst], x[src] and e0 lines towards L2
            const int nt = tile + tile_step;
            if (nt < num_tiles) {
                const int r = tid >> 1, p = nt * TILE_M + r;
                if (p < a.rows) {
                    const int d = a.dst_sorted[p], sr = a.src_sorted[p];
                    if ((unsigned)d < (unsigned)a.N && (unsigned)sr < (unsigned)a.N) {
                        const int hlf = (tid & 1) * 128;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(a.x_bf16 + (size_t)d * LAT) + hlf));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(a.x_bf16 + (size_t)sr * LAT) + hlf));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(a.e0 + (size_t)p * LAT) + hlf));
                    }
                }
            }
        }

        // ------------------------------------------------------------------ tensor-core layers
        uint4 gv[MODE == EDGE ? 12 : 1];
        auto edge_gather = [&](int ph) {
            if (MODE != EDGE) return;
#pragma unroll
            for (int it = 0; it < (MODE == EDGE ? 12 : 1); ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = 3 * ph + (idx >> 10);
                const int d = sDst[r];
                gv[it] = make_uint4(0, 0, 0, 0);
                if (d >= 0) {
                    const __nv_bfloat16* src;
                    if (kb < 2) src = a.x_bf16 + (size_t)d * LAT + kb * 64;
                    else if (kb < 4) src = a.x_bf16 + (size_t)sSrc[r] * LAT + (kb - 2) * 64;
                    else src = a.e0 + (size_t)(row0 + r) * LAT + (kb - 4) * 64;
                    gv[it] = ld_act<COH>(reinterpret_cast<const uint4*>(src + c * 8));
                }
            }
        };
#pragma unroll 1
        for (int layer = 0; layer < C::TC_LAYERS; ++layer) {
            const bool first = layer == 0;
            const int phases = (MODE == EDGE && first) ? 2 : 1;
#pragma unroll 1
            for (int ph = 0; ph < phases; ++ph) {
                int nchunks = first ? (MODE == EDGE ? 3 : C::K1_CHUNKS) : 2;
                if (MODE == EDGE && first) {
                    // 3 K blocks of [x[dst] | x[src] | e0] (blocks 3*ph .. 3*ph+2): phase 0 is fetched here, phase 1 was
                    // fetched into gv[] while phase 0's MMAs ran (below)
                    if (ph == 0) edge_gather(0);
#pragma unroll
                    for (int it = 0; it < 12; ++it) {
                        const int idx = it * THREADS + tid;
                        const int c = idx & 7, r = (idx >> 3) & 127, kl = idx >> 10;
                        *reinterpret_cast<uint4*>(sA + (size_t)kl * CHUNK_BYTES + swz_off(r, c)) = gv[it];
                    }
                }
                // the resident unit image (bulk copy) must have landed before anybody touches it
                if (!weights_ready) { mbar_wait(bar_w, ws.w_phase); ws.w_phase ^= 1; weights_ready = true; }
                // generic-proxy writes of the A operand -> visible to the tensor core (async proxy)
                fence_proxy_async();
                tc_fence_before();
                __syncthreads();
                if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(2);
                // One elected lane of warp 0 issues the chain (a warp-uniform branch: the descriptor arithmetic stays in the
                // uniform datapath instead of being moved there value by value behind a divergent `tid == 0`); the chunk loop
                // is unrolled to its compile-time maximum.  Measured (csb_debug_gnn_trace): 12 MMAs took 1.06 us to issue.
                if (warp_u == 0) {
                    if (elect_one()) {
                        tc_fence_after();
                        const uint32_t a0 = sA_u32, b0 = sW_u32 + (uint32_t)wchunk * CHUNK_BYTES;
#pragma unroll
                        for (int s = 0; s < C::K1_CHUNKS; ++s) {
                            if (s < nchunks) {
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const uint64_t ad = umma_desc(a0 + (uint32_t)s * CHUNK_BYTES + k * 32);
                                    const uint64_t bd = umma_desc(b0 + (uint32_t)s * CHUNK_BYTES + k * 32);
                                    umma(tmem, ad, bd, (ph > 0 || s > 0 || k > 0) ? 1u : 0u);
                                }
                            }
                        }
                        umma_commit(bar_mma);
                    }
                    __syncwarp();
                }
                wchunk += nchunks;
                if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(3);
                if (MODE == EDGE && first && ph == 0) edge_gather(1);  // in flight while phase 0 is multiplied
                mbar_wait(bar_mma, mma_parity);
                mma_parity ^= 1;
                tc_fence_after();
                if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(4);
            }
            // ---- epilogue of this layer: TMEM -> registers
            float v[64];
            tmem_ld32(taddr, v);
            tmem_ld32(taddr + 32, v + 32);
            const int bias_idx = (MODE == ENC) ? layer + 1 : layer;
            const float* bias = sF + bias_idx * LAT + half * 64;
            const bool last = layer == C::TC_LAYERS - 1;
            if (!last) {
                // + bias, ReLU, bf16 -> A operand of the next layer (K block `half`)
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    float h[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) h[j] = fmaxf(v[c * 8 + j] + bias[c * 8 + j], 0.f);
                    uint4 pk = make_uint4(pack_bf16(h[0], h[1]), pack_bf16(h[2], h[3]), pack_bf16(h[4], h[5]),
                                          pack_bf16(h[6], h[7]));
                    *reinterpret_cast<uint4*>(sA + (size_t)half * CHUNK_BYTES + swz_off(row, c)) = pk;
                }
                if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(5);
                continue;
            }
            if (MODE == DEC) {
                // ReLU(h2) stays fp32; last Linear (nout x 128) on the CUDA cores
                const float* W2 = sF + 2 * LAT;
                const float* b2 = W2 + a.nout * LAT;
#pragma unroll
                for (int j = 0; j < 64; ++j) v[j] = fmaxf(v[j] + bias[j], 0.f);
                for (int o = 0; o < a.nout; ++o) {
                    const float* w = W2 + o * LAT + half * 64;
                    float acc = 0.f;
#pragma unroll
                    for (int j = 0; j < 64; ++j) acc = fmaf(v[j], w[j], acc);
                    sRed[(row * 2 + half) * 8 + o] = acc;
                }
                __syncthreads();
                if (half == 0 && rvalid)
                    for (int o = 0; o < a.nout; ++o)
                        a.y[(size_t)grow * a.nout + o] = sRed[(row * 2) * 8 + o] + sRed[(row * 2 + 1) * 8 + o] + b2[o];
                tc_fence_before();
                __syncthreads();
                continue;
            }
            // ---- + bias, LayerNorm (two pass, fp32)
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) { v[j] += bias[j]; s += v[j]; }
            sRed[row * 2 + half] = s;
            __syncthreads();
            const float mean = (sRed[row * 2] + sRed[row * 2 + 1]) * (1.0f / LAT);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) { const float dlt = v[j] - mean; q = fmaf(dlt, dlt, q); }
            sRed[256 + row * 2 + half] = q;
            __syncthreads();
            const float var = (sRed[256 + row * 2] + sRed[256 + row * 2 + 1]) * (1.0f / LAT);
            const float rstd = 1.0f / sqrtf(var + 1e-5f);
            const float* lg = sF + 3 * LAT + half * 64;
            const float* lb = sF + 4 * LAT + half * 64;
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = (v[j] - mean) * rstd * lg[j] + lb[j];
            if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(6);

            if (MODE == ENC) {
                if (rvalid) {
                    uint4* ob = reinterpret_cast<uint4*>(a.out_bf16 + (size_t)grow * LAT + half * 64);
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        ob[c] = make_uint4(pack_bf16(v[c * 8], v[c * 8 + 1]), pack_bf16(v[c * 8 + 2], v[c * 8 + 3]),
                                           pack_bf16(v[c * 8 + 4], v[c * 8 + 5]), pack_bf16(v[c * 8 + 6], v[c * 8 + 7]));
                    if (a.x_f32) {
                        float4* of = reinterpret_cast<float4*>(a.x_f32 + (size_t)grow * LAT + half * 64);
#pragma unroll
                        for (int c = 0; c < 16; ++c) of[c] = make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
                    }
                }
            } else if (MODE == NODE) {
                // stage LN(...) row-per-thread (swizzled float4 chunks, conflict free), then update the
                // fp32 master copy and the bf16 copy with fully coalesced row accesses
                float* stg = reinterpret_cast<float*>(sA);  // 128 x 128 fp32 = the idle A region
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const int chunk = (half * 16 + c) ^ (row & 7);
                    *reinterpret_cast<float4*>(stg + row * LAT + chunk * 4) =
                        make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
                }
                __syncthreads();
                // the warp's 16 rows of the fp32 master: every load in flight before the first use (v[] is dead here; one
                // at a time this loop was 4 x a global-load latency = 4.5 us of a 15 us tile, csb_debug_gnn_trace)
                float4 xo[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int n = row0 + warp * 16 + i;
                    xo[i] = n < a.rows ? *(reinterpret_cast<const float4*>(a.x_f32 + (size_t)n * LAT) + lane)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int r = warp * 16 + i;
                    const int n = row0 + r;
                    if (n >= a.rows) break;  // warp-uniform
                    const float4 u = *reinterpret_cast<const float4*>(stg + r * LAT + ((lane ^ (r & 7)) * 4));
                    float4 o = xo[i];
                    o.x += u.x; o.y += u.y; o.z += u.z; o.w += u.w;
                    *(reinterpret_cast<float4*>(a.x_f32 + (size_t)n * LAT) + lane) = o;
                    *(reinterpret_cast<uint2*>(a.out_bf16 + (size_t)n * LAT) + lane) =
                        make_uint2(pack_bf16(o.x, o.y), pack_bf16(o.z, o.w));
                }
            } else if (MODE == EDGE) {
                // segmented sum over the warp's 32 dst-sorted rows, 32 columns per round, staged in the
                // (now idle) A region; lane j owns one column and adds once per segment.
                // The segment structure of the warp's 32 rows is one ballot (bit i: row i is the last of its dst run) and
                // the walk below is straight-line code: 32 independent shared loads, a running sum, a predicated red.add at
                // the run ends.  (A data-dependent branch per row made this loop 4.3 us of a 20 us tile: every iteration
                // waited for its own two shared loads, csb_debug_gnn_trace.)  Same summation order as before.
                float* st = reinterpret_cast<float*>(sA) + warp * (32 * 33);
                const int* wd = sDst + (warp & 3) * 32;
                const int my_d = wd[lane];
                const int nx_d = lane < 31 ? wd[lane + 1] : -2;
                const uint32_t ends = __ballot_sync(0xFFFFFFFFu, my_d != nx_d);
#pragma unroll
                for (int rd = 0; rd < 2; ++rd) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) st[lane * 33 + j] = v[rd * 32 + j];
                    __syncwarp();
                    float* out_col = a.agg + (half * 64 + rd * 32 + lane);
                    float xv[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) xv[i] = st[i * 33 + lane];
                    float acc = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        acc += xv[i];
                        if ((ends >> i) & 1u) {  // warp-uniform
                            const int d = __shfl_sync(0xFFFFFFFFu, my_d, i);
                            if (d >= 0) atomicAdd(out_col + (size_t)d * LAT, acc);
                            acc = 0.f;
                        }
                    }
                    __syncwarp();
                }
            }
            // the next tile overwrites sA / TMEM: order this tile's TMEM reads and staging before it
            tc_fence_before();
            __syncthreads();
            if (TRACE && (MODE == EDGE || MODE == NODE)) trace_pt(7);
        }
    }
    ws.weights_ready = weights_ready;
    ws.mma_parity = mma_parity;
}

struct CtaSetup {
    uint8_t* smem;
    uint32_t tmem, bar_w, bar_mma;
};
// mbarriers + TMEM of a CTA of any role; the control words sit behind the largest role's carve-up
__device__ __forceinline__ CtaSetup cta_setup(uint8_t* smem_raw, size_t ctrl_offset) {
    CtaSetup cs;
    cs.smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(cs.smem + ctrl_offset);  // [0] weights, [1] mma
    uint32_t* sTmem = reinterpret_cast<uint32_t*>(bars + 2);
    cs.bar_w = smem_u32(&bars[0]);
    cs.bar_mma = smem_u32(&bars[1]);
    if (threadIdx.x == 0) {
        mbar_init(cs.bar_w, 1);
        mbar_init(cs.bar_mma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) tmem_alloc(smem_u32(sTmem), TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cs.tmem = *sTmem;
    return cs;
}
template <int MODE> __device__ __forceinline__ size_t ctrl_offset_of(int kin, int nout) {
    const int nf32 = unit_f32(MODE, kin, nout);
    return (size_t)(Cfg<MODE>::W_CHUNKS + Cfg<MODE>::A_SLOTS) * CHUNK_BYTES + (size_t)((nf32 + 31) & ~31) * 4 +
           SCRATCH_FLOATS * 4 + 2 * TILE_M * 4;
}

template <int MODE, bool TRACE = false>
__global__ void __launch_bounds__(THREADS, (MODE == ENC || MODE == DEC) ? 2 : 1) gnn_mlp_kernel(const __grid_constant__ Args a) {
    extern __shared__ uint8_t smem_raw[];
    const CtaSetup cs = cta_setup(smem_raw, ctrl_offset_of<MODE>(a.kin, a.nout));
    WalkState ws;
    issue_weights<MODE>(a, cs.smem, cs.bar_w, ws);
    // Programmatic dependent launch: everything above (barriers, TMEM, the bulk copy of the constant weight image)
    // overlaps the tail of the previous kernel of the forward; its activations are only touched below.
    grid_dep_trigger();
    EdgeRows pre;
    if (MODE == EDGE) pre = load_edge_rows(a, blockIdx.x);
    grid_dep_wait();
    mlp_tiles<MODE, false, TRACE>(a, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, blockIdx.x, gridDim.x, pre);
    if (!ws.weights_ready && threadIdx.x == 0) mbar_wait(cs.bar_w, ws.w_phase);  // never leave a bulk copy in flight
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(cs.tmem, TMEM_COLS);
}

// ---- small graphs: the whole forward as ONE persistent kernel ----------------------------------------------------
// A graph whose edge and node tiles all fit on the GPU at once (C1: 20x20 cloth = 18 + 4 tiles on 148 SMs) spends its
// forward in launch boundaries: 33 launches of ~10 us of work each.  Here every tile gets its own CTA for the whole
// forward: edge CTA t owns edge tile t through encoder and all message-passing steps, node CTA t owns node tile t through
// encoder, all steps and the decoder.  The roles meet at two grid-wide counters per step in global memory (edge phase done
// -> node phase; node phase done -> next edge phase); the weights of the NEXT walk are bulk-copied while the CTA waits for
// the other role, so no weight load is exposed.  Data other CTAs produce (x for the edge gathers, agg for the node role)
// is read with L2-only loads (ld_act<true>).  Arithmetic is mlp_tiles' -- the same as the multi-launch path.
struct SmallArgs {
    Args enc_node, enc_edge, edge, node, dec;   // edge / node: `unit` is replaced per step
    const uint8_t* packed;
    uint32_t edge_unit[32], node_unit[32];      // byte offsets of the steps' unit images
    int steps, edge_tiles, node_tiles;
    uint32_t* ctrl;                              // [0] encoders done, [1 + m] edge phase m, [33 + m] node phase m, [65] timeout flag
    unsigned long long timeout_ns;
};
constexpr int SMALL_CTRL_WORDS = 128;

__device__ __forceinline__ unsigned long long gnn_now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// all threads of the CTA have finished their global writes of the phase -> one release increment
__device__ __forceinline__ void grid_arrive(uint32_t* ctr) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
    }
}
// false: the other role did not get there within the timeout (the kernel then winds down instead of hanging the GPU)
__device__ __forceinline__ bool grid_wait(uint32_t* ctr, uint32_t target, const SmallArgs& s) {
    __shared__ int ok_flag;
    if (threadIdx.x == 0) {
        int ok = 1;
        const unsigned long long t0 = gnn_now_ns();
        uint32_t spins = 0, v;
        while (true) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
            if (v >= target) break;
            if ((++spins & 255u) == 0 && (gnn_now_ns() - t0 > s.timeout_ns || *(volatile uint32_t*)(s.ctrl + 65))) {
                *(volatile uint32_t*)(s.ctrl + 65) = 1;
                ok = 0;
                break;
            }
        }
        ok_flag = ok;
    }
    __syncthreads();
    const bool ok = ok_flag != 0;
    __syncthreads();
    return ok;
}

template <bool TRACE>
__global__ void __launch_bounds__(THREADS, 1) gnn_small_kernel(const __grid_constant__ SmallArgs s) {
    extern __shared__ uint8_t smem_raw[];
    // control words behind the largest carve-up (EDGE): the same place for every role of this kernel
    const CtaSetup cs = cta_setup(smem_raw, ctrl_offset_of<EDGE>(0, LAT) > ctrl_offset_of<NODE>(0, LAT)
                                                ? ctrl_offset_of<EDGE>(0, LAT) : ctrl_offset_of<NODE>(0, LAT));
    WalkState ws;
    const bool is_edge = (int)blockIdx.x < s.edge_tiles;
    const int tile = is_edge ? blockIdx.x : blockIdx.x - s.edge_tiles;
    const uint32_t everyone = (uint32_t)(s.edge_tiles + s.node_tiles);
    // generic-proxy accesses of this CTA to shared memory (epilogues, staging) before the next bulk copy lands on it
    auto next_weights = [&](auto mode_tag, const Args& a) {
        __syncthreads();
        fence_proxy_async();
        issue_weights<decltype(mode_tag)::value>(a, cs.smem, cs.bar_w, ws);
    };
    bool ok = true;
    if (is_edge) {
        issue_weights<ENC>(s.enc_edge, cs.smem, cs.bar_w, ws);
        mlp_tiles<ENC, false, false>(s.enc_edge, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, tile, 1 << 30);
        Args e = s.edge;
        e.unit = s.packed + s.edge_unit[0];
        const EdgeRows rows = load_edge_rows(e, tile);  // the CTA's tile never changes: its row indices are read once
        next_weights(std::integral_constant<int, EDGE>{}, e);
        grid_arrive(s.ctrl + 0);
        for (int m = 0; m < s.steps && ok; ++m) {
            ok = m == 0 ? grid_wait(s.ctrl + 0, everyone, s) : grid_wait(s.ctrl + 33 + (m - 1), (uint32_t)s.node_tiles, s);
            if (!ok) break;
            if (TRACE) trace_pt(8);
            mlp_tiles<EDGE, true, TRACE>(e, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, tile, 1 << 30, rows);
            if (m + 1 < s.steps) {
                e.unit = s.packed + s.edge_unit[m + 1];
                next_weights(std::integral_constant<int, EDGE>{}, e);
            }
            grid_arrive(s.ctrl + 1 + m);
            if (TRACE) trace_pt(9);
        }
    } else {
        issue_weights<ENC>(s.enc_node, cs.smem, cs.bar_w, ws);
        mlp_tiles<ENC, false, false>(s.enc_node, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, tile, 1 << 30);
        Args n = s.node;
        n.unit = s.packed + s.node_unit[0];
        next_weights(std::integral_constant<int, NODE>{}, n);
        grid_arrive(s.ctrl + 0);
        for (int m = 0; m < s.steps && ok; ++m) {
            ok = grid_wait(s.ctrl + 1 + m, (uint32_t)s.edge_tiles, s);
            if (!ok) break;
            if (TRACE) trace_pt(8);
            mlp_tiles<NODE, true, TRACE>(n, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, tile, 1 << 30);
            if (m + 1 < s.steps) {
                n.unit = s.packed + s.node_unit[m + 1];
                next_weights(std::integral_constant<int, NODE>{}, n);
            } else {
                next_weights(std::integral_constant<int, DEC>{}, s.dec);
            }
            grid_arrive(s.ctrl + 33 + m);
            if (TRACE) trace_pt(9);
        }
        // the decoder reads the rows this CTA has just written itself
        if (ok) mlp_tiles<DEC, false, false>(s.dec, cs.smem, cs.tmem, cs.bar_w, cs.bar_mma, ws, tile, 1 << 30);
    }
    if (!ws.weights_ready && threadIdx.x == 0) mbar_wait(cs.bar_w, ws.w_phase);  // never leave a bulk copy in flight
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(cs.tmem, TMEM_COLS);
}

// ---- graph preparation: dst-sorted edge order ---------------------------------------------------
__global__ void gnn_edge_keys_kernel(int E, const long long* __restrict__ edge_index, uint32_t* keys, uint32_t* vals,
                                     uint32_t* n_dev) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *n_dev = (uint32_t)E;
    if (i >= E) return;
    keys[i] = (uint32_t)edge_index[(size_t)E + i];  // dst = edge_index[1]
    vals[i] = (uint32_t)i;
}
__global__ void gnn_edge_finish_kernel(int E, const long long* __restrict__ edge_index, const uint32_t* __restrict__ keys,
                                       const uint32_t* __restrict__ vals, int* dst_sorted, int* src_sorted, int* perm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= E) return;
    const uint32_t e = vals[i];
    dst_sorted[i] = (int)keys[i];
    src_sorted[i] = (int)edge_index[e];
    perm[i] = (int)e;
}

struct Workspace {
    float* x_f32;
    __nv_bfloat16* x_bf16;
    float* agg;
    __nv_bfloat16* e0;
    uint32_t* ctrl;  // grid counters of gnn_small_kernel
    size_t bytes;
    static Workspace carve(void* ws, int N, int E) {
        Workspace w;
        Carver c(ws);
        w.x_f32 = c.take<float>((size_t)N * LAT);
        w.x_bf16 = c.take<__nv_bfloat16>((size_t)N * LAT);
        w.agg = c.take<float>((size_t)N * LAT);
        w.e0 = c.take<__nv_bfloat16>((size_t)E * LAT);
        w.ctrl = c.take<uint32_t>(SMALL_CTRL_WORDS);
        w.bytes = c.total();
        return w;
    }
};

struct GraphWs {
    uint32_t *keys_a, *keys_b, *vals_a, *vals_b, *control;
    size_t control_words, bytes;
    static GraphWs carve(void* ws, int E) {
        GraphWs g;
        Carver c(ws);
        g.keys_a = c.take<uint32_t>(E);
        g.keys_b = c.take<uint32_t>(E);
        g.vals_a = c.take<uint32_t>(E);
        g.vals_b = c.take<uint32_t>(E);
        g.control_words = 8 + 4 * 256 + sort_lookback_words((size_t)E, 4);
        g.control = c.take<uint32_t>(g.control_words);
        g.bytes = c.total();
        return g;
    }
};

static bool g_trace_on = false;  // csb_debug_gnn_trace: the instantiations with %globaltimer stations (EDGE / NODE roles)

template <int MODE>
static int launch_role(const Args& a, int kin, int nout, cudaStream_t stream, const char* name) {
    static unsigned long long configured = 0, configured_t = 0;  // per-device opt-in (common.cuh)
    const size_t smem = smem_bytes(MODE, kin, nout);
    const bool traced = g_trace_on && (MODE == EDGE || MODE == NODE);
    if (traced) CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_mlp_kernel<MODE, true>, 232448 - 1024, &configured_t));
    CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_mlp_kernel<MODE>, 232448 - 1024, &configured));
    if (smem > 232448 - 1024) return CSB_ERR_UNSUPPORTED;
    const int tiles = (a.rows + TILE_M - 1) / TILE_M;
    if (tiles == 0) return 0;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // ENC / DEC: ~108 KB of shared memory and 128 registers: two CTAs per SM overlap each other's serial tile chains
    const int per_sm = (MODE == ENC || MODE == DEC) ? 2 : 1;
    const int grid = tiles < sms * per_sm ? tiles : sms * per_sm;
    if (traced) CSB_LAUNCH(name, stream, launch_pdl(gnn_mlp_kernel<MODE, true>, dim3(grid), dim3(THREADS), smem, stream, a));
    else CSB_LAUNCH(name, stream, launch_pdl(gnn_mlp_kernel<MODE>, dim3(grid), dim3(THREADS), smem, stream, a));
    return 0;
}

// ---- EDGE role, two tiles in flight per CTA -------------------------------------------------------
// The generic kernel above walks a tile as one serial chain (gather -> MMA -> epilogue -> ... -> scatter): the tensor
// pipe idles during gathers and epilogues, the load/store units during the MMAs.  Here the CTA is two independent
// groups of 128 threads that share the resident weight image (160 KB): each group owns its own tile, 2 K-blocks of A
// operand (32 KB), 128 TMEM columns, an mbarrier and a named barrier, so one group's gathers / epilogues overlap the
// other's MMAs -- two "virtual CTAs" per SM, which the weights' footprint forbids as real ones.  To fit 227 KB the
// row indices live in registers (a warp gathers and scatters its own 32 rows: indices by shuffle), a thread owns a
// whole row (LayerNorm without shared-memory partials; same summation order as the generic kernel) and the first
// Linear's K = 384 is fed in three phases of two K-blocks whose loads are issued before the previous phase's MMAs are
// waited for.  Arithmetic and roundings are those of gnn_mlp_kernel<EDGE>.
constexpr int E2_A_CHUNKS = 2;
constexpr size_t E2_SMEM = (size_t)10 * CHUNK_BYTES + 2 * E2_A_CHUNKS * CHUNK_BYTES + 5 * LAT * 4 + 64;
static_assert(E2_SMEM <= 232448, "edge kernel shared memory");

__device__ __forceinline__ void group_sync(int grp) { asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory"); }

__global__ void __launch_bounds__(THREADS, 1) gnn_edge2_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sW = smem;
    uint8_t* sAall = sW + (size_t)10 * CHUNK_BYTES;
    float* sF = reinterpret_cast<float*>(sAall + (size_t)2 * E2_A_CHUNKS * CHUNK_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sF + 5 * LAT);  // [0] weights, [1], [2] MMA of group 0 / 1
    uint32_t* sTmem = reinterpret_cast<uint32_t*>(bars + 3);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = tid >> 7, gt = tid & 127, wg = warp & 3;
    const int row = gt;  // tile row == TMEM lane owned by this thread
    uint8_t* sA = sAall + (size_t)grp * E2_A_CHUNKS * CHUNK_BYTES;
    const uint32_t bar_w = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1 + grp]);

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();  // swizzle-128B operands need the 1024-byte aligned base
        mbar_init(bar_w, 1);
        mbar_init(smem_u32(&bars[1]), 1);
        mbar_init(smem_u32(&bars[2]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(sTmem), 2 * TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *sTmem + (uint32_t)(grp * TMEM_COLS);  // this group's accumulator columns
    if (tid == 0) {
        const uint32_t wbytes = 10u * CHUNK_BYTES, fbytes = 5u * LAT * 4u;
        mbar_expect_tx(bar_w, wbytes + fbytes);
        for (int c = 0; c < 10; ++c)
            bulk_g2s(smem_u32(sW + (size_t)c * CHUNK_BYTES), a.unit + (size_t)c * CHUNK_BYTES, CHUNK_BYTES, bar_w);
        bulk_g2s(smem_u32(sF), a.unit + wbytes, fbytes, bar_w);
    }
    grid_dep_trigger();
    grid_dep_wait();
    bool weights_ready = false;
    uint32_t mma_parity = 0;
    const uint32_t sA_u32 = smem_u32(sA), sW_u32 = smem_u32(sW);
    const uint32_t taddr = tmem + ((uint32_t)(wg * 32) << 16);
    const unsigned full = 0xffffffffu;

    const int num_tiles = (a.rows + TILE_M - 1) / TILE_M;
    const int stride = 2 * (int)gridDim.x;
    int tile = (int)blockIdx.x + grp * (int)gridDim.x;
    // row indices of the first tile
    int d = -1, sidx = -1;
    if (tile < num_tiles) {
        const int p = tile * TILE_M + row;
        if (p < a.rows) {
            d = a.dst_sorted[p];
            sidx = a.src_sorted[p];
            if ((unsigned)d >= (unsigned)a.N || (unsigned)sidx >= (unsigned)a.N) d = sidx = -1;
        }
    }
    for (; tile < num_tiles; tile += stride) {
        const int row0 = tile * TILE_M;
        // indices of this group's next tile: loaded now, used for the L2 prefetch below and as (d, sidx) next round
        int dn = -1, sn = -1;
        const int pn = (tile + stride) * TILE_M + row;
        if (tile + stride < num_tiles && pn < a.rows) {
            dn = a.dst_sorted[pn];
            sn = a.src_sorted[pn];
            if ((unsigned)dn >= (unsigned)a.N || (unsigned)sn >= (unsigned)a.N) dn = sn = -1;
        }

        // ---- first Linear, K = 384 as three phases of two K-blocks: x[dst] | x[src] | e0
        // warp wg gathers its own 32 rows: item (it, lane) = 16-byte piece (lane & 7) of row 32 wg + 4 it + (lane >> 3)
        uint4 v[16];
        auto issue_loads = [&](int ph) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int rl = it * 4 + (lane >> 3);
                const int dd = __shfl_sync(full, d, rl);
                const int ss = __shfl_sync(full, sidx, rl);
                const __nv_bfloat16* src;
                if (ph == 0) src = a.x_bf16 + (size_t)(dd < 0 ? 0 : dd) * LAT;
                else if (ph == 1) src = a.x_bf16 + (size_t)(ss < 0 ? 0 : ss) * LAT;
                else src = a.e0 + (size_t)(row0 + wg * 32 + rl) * LAT;
                v[2 * it] = v[2 * it + 1] = make_uint4(0, 0, 0, 0);
                if (dd >= 0) {
                    v[2 * it] = *reinterpret_cast<const uint4*>(src + (lane & 7) * 8);
                    v[2 * it + 1] = *reinterpret_cast<const uint4*>(src + 64 + (lane & 7) * 8);
                }
            }
        };
        issue_loads(0);
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            if (ph > 0) {  // the previous phase's MMAs have read the A blocks
                mbar_wait(bar_mma, mma_parity);
                mma_parity ^= 1;
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int r = wg * 32 + it * 4 + (lane >> 3);
                *reinterpret_cast<uint4*>(sA + swz_off(r, lane & 7)) = v[2 * it];
                *reinterpret_cast<uint4*>(sA + CHUNK_BYTES + swz_off(r, lane & 7)) = v[2 * it + 1];
            }
            if (!weights_ready) { mbar_wait(bar_w, 0); weights_ready = true; }
            fence_proxy_async();
            tc_fence_before();
            group_sync(grp);
            if (gt == 0) {
                tc_fence_after();
                // descriptors are rebuilt from the two base addresses at every issue (a few ALU ops): left to the
                // compiler, all 40 of them become loop invariants held in registers and spill
                uint32_t aw = sA_u32, ww = sW_u32;
                asm volatile("" : "+r"(aw), "+r"(ww));
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma(tmem, umma_desc(aw + (uint32_t)s * CHUNK_BYTES + k * 32),
                             umma_desc(ww + (uint32_t)(2 * ph + s) * CHUNK_BYTES + k * 32),
                             (ph > 0 || s > 0 || k > 0) ? 1u : 0u);
                umma_commit(bar_mma);
            }
            if (ph < 2) issue_loads(ph + 1);  // in flight while the tensor core works
        }
        // next tile's rows towards L2 (each thread its own row: x[dst], x[src], e0 -- two 128-byte lines each)
        if (dn >= 0) {
            const char* pd = reinterpret_cast<const char*>(a.x_bf16 + (size_t)dn * LAT);
            const char* ps = reinterpret_cast<const char*>(a.x_bf16 + (size_t)sn * LAT);
            const char* pe = reinterpret_cast<const char*>(a.e0 + (size_t)pn * LAT);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pd));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pd + 128));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(ps));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(ps + 128));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pe));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pe + 128));
        }
        mbar_wait(bar_mma, mma_parity);
        mma_parity ^= 1;
        tc_fence_after();

        // ---- layers 2 and 3: + bias, ReLU, bf16 -> A blocks -> MMA
#pragma unroll 1
        for (int layer = 0; layer < 2; ++layer) {
            const float* bias = sF + layer * LAT;
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
                float f[64];
                tmem_ld_cols<2>(taddr + h * 64, f);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    float q[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) q[j] = fmaxf(f[c * 8 + j] + bias[h * 64 + c * 8 + j], 0.f);
                    *reinterpret_cast<uint4*>(sA + (size_t)h * CHUNK_BYTES + swz_off(row, c)) =
                        make_uint4(pack_bf16(q[0], q[1]), pack_bf16(q[2], q[3]), pack_bf16(q[4], q[5]), pack_bf16(q[6], q[7]));
                }
            }
            fence_proxy_async();
            tc_fence_before();
            group_sync(grp);
            if (gt == 0) {
                tc_fence_after();
                uint32_t aw = sA_u32, ww = sW_u32 + (uint32_t)(6 + 2 * layer) * CHUNK_BYTES;
                asm volatile("" : "+r"(aw), "+r"(ww));
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma(tmem, umma_desc(aw + (uint32_t)s * CHUNK_BYTES + k * 32),
                             umma_desc(ww + (uint32_t)s * CHUNK_BYTES + k * 32), (s > 0 || k > 0) ? 1u : 0u);
                umma_commit(bar_mma);
            }
            mbar_wait(bar_mma, mma_parity);
            mma_parity ^= 1;
            tc_fence_after();
        }

        // ---- + bias, LayerNorm over the thread's row, all 128 values in registers.  Four interleaved partial sums per
        //      statistic (a 128-long dependent FADD chain is ~0.25 us of pure latency with two warps per scheduler)
        const float* b2 = sF + 2 * LAT;
        float f[LAT];
        tmem_ld_cols<4>(taddr, f);
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < LAT; ++j) { f[j] += b2[j]; s4[j & 3] += f[j]; }
        const float mean = ((s4[0] + s4[1]) + (s4[2] + s4[3])) * (1.0f / LAT);
        float q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < LAT; ++j) { const float dl = f[j] - mean; q4[j & 3] = fmaf(dl, dl, q4[j & 3]); }
        const float var = ((q4[0] + q4[1]) + (q4[2] + q4[3])) * (1.0f / LAT);
        const float rstd = 1.0f / sqrtf(var + 1e-5f);
        const float* lg = sF + 3 * LAT;
        const float* lb = sF + 4 * LAT;

        // ---- segmented sum over the warp's 32 dst-sorted rows, 32 columns per round, staged in the warp's slice of
        //      the (idle) A blocks; lane j owns a column and adds once per segment.  Bit i of `ends`: row i is the last
        //      of its segment (warp-uniform, found once per tile).
        const int dnext = __shfl_down_sync(full, d, 1);
        const unsigned ends = __ballot_sync(full, lane == 31 || dnext != d);
        float* st = reinterpret_cast<float*>(sA) + wg * (32 * 33);
#pragma unroll
        for (int rd = 0; rd < 4; ++rd) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int col = rd * 32 + j;
                st[lane * 33 + j] = (f[col] - mean) * rstd * lg[col] + lb[col];
            }
            __syncwarp();
            const int col = rd * 32 + lane;
            float acc = 0.f;
            // the column's 32 values are fetched 16 at a time BEFORE the walk that may issue atomics: behind an atomic the
            // compiler keeps every shared load in place, and each iteration then waits out its own load (csb_debug_gnn_trace:
            // the scatter was 6.1 us of a 19.5 us tile)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float xv[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) xv[i] = st[(h * 16 + i) * 33 + lane];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    acc += xv[i];
                    if ((ends >> (h * 16 + i)) & 1u) {
                        const int dd = __shfl_sync(full, d, h * 16 + i);
                        if (dd >= 0) atomicAdd(a.agg + (size_t)dd * LAT + col, acc);
                        acc = 0.f;
                    }
                }
            }
            __syncwarp();
        }
        d = dn;
        sidx = sn;
        // the next tile overwrites the A blocks / TMEM columns of this group
        tc_fence_before();
        group_sync(grp);
    }
    if (!weights_ready && tid == 0) mbar_wait(bar_w, 0);  // never leave a bulk copy in flight
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*sTmem, 2 * TMEM_COLS);
}

static int launch_edge2(const Args& a, cudaStream_t stream) {
    static unsigned long long configured = 0;
    CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_edge2_kernel, (int)E2_SMEM, &configured));
    const int tiles = (a.rows + TILE_M - 1) / TILE_M;
    if (tiles == 0) return 0;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // at most one tile per SM: nothing to overlap, the generic kernel's two-phase chain is shorter (C1: 18 tiles)
    if (tiles <= sms) return launch_role<EDGE>(a, 0, LAT, stream, "gnn_edge");
    const int grid = sms;
    CSB_LAUNCH("gnn_edge", stream, launch_pdl(gnn_edge2_kernel, dim3(grid), dim3(THREADS), E2_SMEM, stream, a));
    return 0;
}

}  // namespace gnn
}  // namespace csb

using namespace csb;
using namespace csb::gnn;

static int g_small_path = 1;  // csb_debug_gnn_small: 0 = always the multi-launch path

extern "C" {

size_t csb_gnn_param_floats(const csb_gnn_dims* d) {
    if (check_dims(d)) return 0;
    PackTable t;
    size_t f = 0;
    build_table(d, t, &f, nullptr);
    return f;
}
size_t csb_gnn_packed_bytes(const csb_gnn_dims* d) {
    if (check_dims(d)) return 0;
    PackTable t;
    size_t b = 0;
    build_table(d, t, nullptr, &b);
    return b;
}
int csb_gnn_pack(const csb_gnn_dims* d, const float* flat_params, void* packed, void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (!flat_params || !packed) return CSB_ERR_BAD_ARG;
    PackTable t;
    build_table(d, t, nullptr, nullptr);
    cudaStream_t stream = (cudaStream_t)stream_;
    CSB_LAUNCH("gnn_pack", stream, gnn_pack_kernel<<<t.n, 256, 0, stream>>>(flat_params, (uint8_t*)packed, t));
    return 0;
}

int csb_debug_gnn_trace(void* device_buffer, int block) {
    unsigned long long* p = (unsigned long long*)device_buffer;
    CSB_CUDA_TRY(cudaMemcpyToSymbol(csb::gnn::g_trace_buf, &p, sizeof(p)));
    CSB_CUDA_TRY(cudaMemcpyToSymbol(csb::gnn::g_trace_block, &block, sizeof(block)));
    csb::gnn::g_trace_on = p != nullptr;
    return 0;
}

int csb_debug_gnn_small(int on) {
    const int old = g_small_path;
    if (on == 0 || on == 1) g_small_path = on;
    return old;
}

size_t csb_gnn_graph_workspace_bytes(int E) { return GraphWs::carve(nullptr, E < 0 ? 0 : E).bytes; }
int csb_gnn_prepare_graph(int N, int E, const int64_t* edge_index, int32_t* dst_sorted, int32_t* src_sorted,
                          int32_t* perm, void* ws, size_t ws_bytes, void* stream_) {
    if (N < 0 || E < 0) return CSB_ERR_BAD_ARG;
    if (E == 0) return 0;
    if (!edge_index || !dst_sorted || !src_sorted || !perm || !ws) return CSB_ERR_BAD_ARG;
    GraphWs g = GraphWs::carve(ws, E);
    if (ws_bytes < g.bytes) return CSB_ERR_WORKSPACE;
    cudaStream_t stream = (cudaStream_t)stream_;
    CSB_CUDA_TRY(cudaMemsetAsync(g.control, 0, g.control_words * sizeof(uint32_t), stream));
    const int blocks = (E + 255) / 256;
    CSB_LAUNCH("gnn_edge_keys", stream, gnn_edge_keys_kernel<<<blocks, 256, 0, stream>>>(
        E, (const long long*)edge_index, g.keys_a, g.vals_a, g.control));
    int side = 0;
    int rc = radix_sort_pairs(g.keys_a, g.keys_b, g.vals_a, g.vals_b, g.control, (size_t)E,
                              higher_msb((uint32_t)(N > 0 ? N - 1 : 0)), g.control + 8, g.control + 4,
                              g.control + 8 + 4 * 256, &side, stream);
    if (rc) return rc;
    CSB_LAUNCH("gnn_edge_finish", stream, gnn_edge_finish_kernel<<<blocks, 256, 0, stream>>>(
        E, (const long long*)edge_index, side ? g.keys_b : g.keys_a, side ? g.vals_b : g.vals_a, dst_sorted,
        src_sorted, perm));
    return 0;
}

size_t csb_gnn_workspace_bytes(const csb_gnn_dims* d, int N, int E) {
    if (check_dims(d) || N < 0 || E < 0) return 0;
    return Workspace::carve(nullptr, N, E).bytes;
}


// The whole forward as one persistent kernel (gnn_small_kernel) when every tile can have its own CTA.  Returns 1 when it
// ran, 0 when the graph does not qualify, < 0 / cudaError_t on failure.
static int forward_small(const csb_gnn_dims* d, const PackTable& t, const uint8_t* pk, int N, int E, const float* x,
                         const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                         const int32_t* perm, float* y, float* x_f32, float* agg, __nv_bfloat16* e0,
                         __nv_bfloat16* x_bf16, uint32_t* ctrl, cudaStream_t stream) {
    const int steps = d->nmessage_passing_steps;
    const int te = (E + TILE_M - 1) / TILE_M, tn = (N + TILE_M - 1) / TILE_M;
    int dev = 0, sms = 148, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (!g_small_path || !coop || !ctrl || E <= 0 || steps < 1 || steps > 32 || te + tn > sms) return 0;
    static unsigned long long configured = 0, configured_t = 0;
    const size_t smem = smem_bytes(EDGE, 0, LAT);
    CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_small_kernel<false>, 232448 - 1024, &configured));
    if (g_trace_on) CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_small_kernel<true>, 232448 - 1024, &configured_t));
    SmallArgs s{};
    Args a{};
    a.N = N;
    a.unit = pk + t.u[0].dst; a.rows = N; a.kin = d->nnode_in; a.nout = LAT;
    a.in = x; a.perm = nullptr; a.out_bf16 = x_bf16; a.x_f32 = x_f32;
    s.enc_node = a;
    a.unit = pk + t.u[1].dst; a.rows = E; a.kin = d->nedge_in;
    a.in = edge_features; a.perm = perm; a.out_bf16 = e0; a.x_f32 = nullptr;
    s.enc_edge = a;
    Args e{};
    e.N = N; e.kin = 0; e.nout = LAT; e.rows = E;
    e.dst_sorted = dst_sorted; e.src_sorted = src_sorted; e.x_bf16 = x_bf16; e.e0 = e0; e.agg = agg;
    s.edge = e;
    Args n{};
    n.N = N; n.kin = 0; n.nout = LAT; n.rows = N;
    n.agg = agg; n.x_bf16 = x_bf16; n.x_f32 = x_f32; n.out_bf16 = x_bf16;
    s.node = n;
    Args c{};
    c.N = N; c.kin = 0; c.nout = d->nnode_out;
    c.unit = pk + t.u[t.n - 1].dst; c.rows = N; c.x_bf16 = x_bf16; c.y = y;
    s.dec = c;
    s.packed = pk;
    for (int m = 0; m < steps; ++m) {
        s.edge_unit[m] = (uint32_t)t.u[2 + 2 * m].dst;
        s.node_unit[m] = (uint32_t)t.u[3 + 2 * m].dst;
    }
    s.steps = steps; s.edge_tiles = te; s.node_tiles = tn;
    s.ctrl = ctrl;
    s.timeout_ns = 2000000000ull;  // 2 s: a forward takes well under a millisecond
    CSB_CUDA_TRY(cudaMemsetAsync(ctrl, 0, SMALL_CTRL_WORDS * sizeof(uint32_t), stream));
    void* kargs[] = {(void*)&s};
    const void* fn = g_trace_on ? (const void*)gnn_small_kernel<true> : (const void*)gnn_small_kernel<false>;
    CSB_LAUNCH("gnn_small", stream, cudaLaunchCooperativeKernel(fn, dim3(te + tn), dim3(THREADS), kargs, smem, stream));
    return 1;
}

static int forward_impl(const csb_gnn_dims* d, const void* packed, int N, int E, const float* x,
                        const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                        const int32_t* perm, float* y, float* x_latent_out, float* x_f32, float* agg,
                        __nv_bfloat16* e0, __nv_bfloat16* x_stack, size_t x_stride, __nv_bfloat16* agg_stack,
                        uint32_t* small_ctrl, cudaStream_t stream) {
    PackTable t;
    build_table(d, t, nullptr, nullptr);
    const uint8_t* pk = (const uint8_t*)packed;
    int rc;

    CSB_CUDA_TRY(cudaMemsetAsync(agg, 0, (size_t)N * LAT * sizeof(float), stream));
    if (small_ctrl && x_stride == 0 && !agg_stack) {  // inference: small graphs run as one persistent kernel
        rc = forward_small(d, t, pk, N, E, x, edge_features, dst_sorted, src_sorted, perm, y, x_f32, agg, e0, x_stack,
                           small_ctrl, stream);
        if (rc < 0 || rc > 1) return rc;
        if (rc == 1) {
            if (x_latent_out)
                CSB_CUDA_TRY(cudaMemcpyAsync(x_latent_out, x_f32, (size_t)N * LAT * sizeof(float), cudaMemcpyDeviceToDevice, stream));
            return 0;
        }
    }
    Args a{};
    a.N = N;
    // encoders
    a.unit = pk + t.u[0].dst; a.rows = N; a.kin = d->nnode_in; a.nout = LAT;
    a.in = x; a.perm = nullptr; a.out_bf16 = x_stack; a.x_f32 = x_f32;
    rc = launch_role<ENC>(a, a.kin, LAT, stream, "gnn_encode");
    if (rc) return rc;
    if (E > 0) {
        a.unit = pk + t.u[1].dst; a.rows = E; a.kin = d->nedge_in;
        a.in = edge_features; a.perm = perm; a.out_bf16 = e0; a.x_f32 = nullptr;
        rc = launch_role<ENC>(a, a.kin, LAT, stream, "gnn_encode");
        if (rc) return rc;
    }
    // processor
    __nv_bfloat16* xcur = x_stack;
    for (int m = 0; m < d->nmessage_passing_steps; ++m) {
        Args e{};
        e.N = N; e.kin = 0; e.nout = LAT;
        e.unit = pk + t.u[2 + 2 * m].dst; e.rows = E;
        e.dst_sorted = dst_sorted; e.src_sorted = src_sorted; e.x_bf16 = xcur; e.e0 = e0; e.agg = agg;
        rc = launch_edge2(e, stream);
        if (rc) return rc;
        Args n{};
        n.N = N; n.kin = 0; n.nout = LAT;
        n.unit = pk + t.u[3 + 2 * m].dst; n.rows = N;
        n.agg = agg; n.x_bf16 = xcur; n.x_f32 = x_f32; n.out_bf16 = xcur + x_stride;
        n.agg_save = agg_stack ? agg_stack + (size_t)m * N * LAT : nullptr;
        rc = launch_role<NODE>(n, 0, LAT, stream, "gnn_node");
        if (rc) return rc;
        xcur += x_stride;
    }
    if (x_latent_out)
        CSB_CUDA_TRY(cudaMemcpyAsync(x_latent_out, x_f32, (size_t)N * LAT * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    // decoder
    Args c{};
    c.N = N; c.kin = 0; c.nout = d->nnode_out;
    c.unit = pk + t.u[t.n - 1].dst; c.rows = N; c.x_bf16 = xcur; c.y = y;
    return launch_role<DEC>(c, 0, d->nnode_out, stream, "gnn_decode");
}

int csb_gnn_forward(const csb_gnn_dims* d, const void* packed, int N, int E, const float* x,
                    const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                    const int32_t* perm, float* y, float* x_latent_out, void* ws, size_t ws_bytes, void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (N < 0 || E < 0) return CSB_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!packed || !x || !y || !ws) return CSB_ERR_BAD_ARG;
    if (E > 0 && (!edge_features || !dst_sorted || !src_sorted || !perm)) return CSB_ERR_BAD_ARG;
    Workspace w = Workspace::carve(ws, N, E);
    if (ws_bytes < w.bytes) return CSB_ERR_WORKSPACE;
    return forward_impl(d, packed, N, E, x, edge_features, dst_sorted, src_sorted, perm, y, x_latent_out, w.x_f32,
                        w.agg, w.e0, w.x_bf16, 0, nullptr, w.ctrl, (cudaStream_t)stream_);
}

size_t csb_gnn_train_state_bytes(const csb_gnn_dims* d, int N, int E) {
    if (check_dims(d) || N < 0 || E < 0) return 0;
    return TrainState::carve(nullptr, N, E, d->nmessage_passing_steps).bytes;
}

int csb_gnn_forward_train(const csb_gnn_dims* d, const void* packed, int N, int E, const float* x,
                          const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                          const int32_t* perm, float* y, void* state, size_t state_bytes, void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (N < 0 || E < 0) return CSB_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!packed || !x || !y || !state) return CSB_ERR_BAD_ARG;
    if (E > 0 && (!edge_features || !dst_sorted || !src_sorted || !perm)) return CSB_ERR_BAD_ARG;
    TrainState w = TrainState::carve(state, N, E, d->nmessage_passing_steps);
    if (state_bytes < w.bytes) return CSB_ERR_WORKSPACE;
    return forward_impl(d, packed, N, E, x, edge_features, dst_sorted, src_sorted, perm, y, nullptr, w.x_f32, w.agg,
                        w.e0, w.x_stack, (size_t)N * LAT, w.agg_stack, nullptr, (cudaStream_t)stream_);
}

}  // extern "C"
