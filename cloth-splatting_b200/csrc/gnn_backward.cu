// meshnet GNN backward for sm_100a: gradients of EncodeProcessDecode.forward (meshnet/graph_network.py:387-408)
// w.r.t. every parameter, the node inputs and the edge features -- what `loss.backward()` computes in the
// reference's multi-step training loop (train_meshnet_sim.py:491-532) through ~600 autograd nodes.
//
// Two kernels:
//   gnn_bwd_kernel<ROLE>  one CTA per 128-row tile, same resident bf16 weight image as the forward.  It RECOMPUTES
//       the role's MLP from the bf16 operands the training forward kept (x_m, agg_m, e_0), runs LayerNorm backward in
//       registers and then the dgrad chain on the tensor cores.  The dgrad B operand is the *same* shared-memory weight
//       chunk the forward multiplies with, read through an MN-major descriptor (W[out, in] seen as K = out), so no
//       transposed copy of the weights exists.  Bias / LayerNorm / tiny-Linear gradients are column sums over the
//       tile's rows: recursive-halving warp shuffles, accumulated in registers over the CTA's tiles, one atomic per
//       column per warp at the end.  Every (dZ_l, A_l) pair is written once as bf16 [rows,128] for the weight grads.
//   gnn_wgrad_kernel      dW_l[out, in] += dZ_l^T A_l as tcgen05.mma with BOTH operands MN-major (K = the tile's rows),
//       the [128 x 128] fp32 accumulator stays in TMEM over all the tiles of the CTA, one vector red.add at the end.
//       A_l is a plain [rows,128] bf16 array or a row gather of x (edge MLP layer 0: x[dst], x[src], e_0).
//
// Message passing backward (InteractionNetwork, graph_network.py:114-222), step m, given dL/dx_{m+1} in `dx`:
//   NODE: x_{m+1} = x_m + LN(MLP([agg_m, x_m]))  ->  dagg = dIn[:, :128],  dx += dIn[:, 128:]
//   EDGE: agg_m[dst] += LN(MLP([x_m[dst], x_m[src], 2^m e_0]))  ->  dx[dst] += dIn[:, :128] (segmented over the dst-sorted
//         rows), dx[src] += dIn[:, 128:256], de_0 += dIn[:, 256:] (2^m is folded into the packed weight block).
#include "gnn_common.cuh"

namespace csb {
namespace gnn {

// MN-major, swizzle-128B descriptor over [K rows x 64 MN elements] chunks (row pitch 128 B, 8-row atoms of 1024 B):
// LBO = distance between 64-element MN blocks = one chunk, SBO = distance between 8-row K groups = 1024 B.
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(CHUNK_BYTES >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}
constexpr uint32_t IDESC_B_MN = IDESC | (1u << 16);               // dgrad: A K-major, B MN-major
constexpr uint32_t IDESC_AB_MN = IDESC | (1u << 15) | (1u << 16);  // wgrad: both MN-major

__device__ __forceinline__ void umma_i(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// Sum over the 32 lanes of v[c] for every c in [0, 32): lane l returns the total of column l.  31 shuffles.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = upper ? v[i] : v[i + off];
            const float keep = upper ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, off);
        }
    }
    return v[0];
}

// floats of one MLP inside the flat parameter / gradient buffer (layout of csb_gnn_pack)
struct MlpOff {
    size_t W0, b0, W1, b1, W2, b2, lng, lnb;
    __host__ __device__ MlpOff(int k0, int n2) {
        W0 = 0; b0 = (size_t)LAT * k0; W1 = b0 + LAT; b1 = W1 + (size_t)LAT * LAT; W2 = b1 + LAT;
        b2 = W2 + (size_t)n2 * LAT; lng = b2 + n2; lnb = lng + LAT;
    }
};

struct BArgs {
    const uint8_t* unit;             // packed unit image of the role
    int rows, kin, nout, N;
    const int* dst_sorted;           // EDGE
    const int* src_sorted;           // EDGE
    const __nv_bfloat16* x_bf16;     // x_m: EDGE gathers, NODE / DEC input [N,128]
    const __nv_bfloat16* e0;         // EDGE [E,128]
    const __nv_bfloat16* agg_bf16;   // NODE [N,128]
    const float* in;                 // ENC raw input [rows_src, kin]
    const int* perm;                 // ENC: source row of row i (NULL = identity)
    const float* dout;               // upstream gradient rows [*,128]: NODE/ENC by row, EDGE gathered by dst
    const float* dy;                 // DEC [N, nout]
    float* dx;                       // [N,128] fp32: EDGE atomics, NODE += , DEC =
    float* dagg;                     // NODE out [N,128]
    float* de0;                      // EDGE += [E,128]
    float* din;                      // ENC out [rows_src, kin] (may be NULL)
    __nv_bfloat16 *sH1, *sH2, *sdY, *sdZ2, *sdZ1;  // bf16 [rows,128] scratch for the weight gradients
    float* g;                        // this MLP's slice of the flat gradient buffer
};

template <int MODE> struct BCfg {
    static constexpr int A_SLOTS = MODE == EDGE ? 3 : MODE == NODE ? 4 : 2;
    static constexpr int W_CHUNKS = MODE == EDGE ? 10 : MODE == NODE ? 8 : 4;
    static constexpr int IN_BLOCKS = MODE == EDGE ? 3 : MODE == NODE ? 2 : 1;  // 128-wide blocks of the first TC layer's input
    static constexpr int TMEM = MODE == EDGE ? 512 : MODE == NODE ? 256 : 128;
};

__host__ __device__ constexpr int bwd_scratch_floats(int mode) { return mode == ENC ? 4096 : 1024; }
static size_t bwd_smem_bytes(int mode, int kin, int nout) {
    const int a_slots = mode == EDGE ? 3 : mode == NODE ? 4 : 2;
    size_t b = 1024;
    b += (size_t)unit_chunks(mode) * CHUNK_BYTES;
    b += (size_t)a_slots * CHUNK_BYTES;
    b += ((size_t)unit_f32(mode, kin, nout) * 4 + 127) / 128 * 128;
    b += bwd_scratch_floats(mode) * 4 + 2 * TILE_M * 4 + 64;
    return b;
}

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) gnn_bwd_kernel(const __grid_constant__ BArgs a) {
    using C = BCfg<MODE>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sW = smem;
    uint8_t* sA = sW + (size_t)C::W_CHUNKS * CHUNK_BYTES;
    float* sF = reinterpret_cast<float*>(sA + (size_t)C::A_SLOTS * CHUNK_BYTES);
    const int nf32 = unit_f32(MODE, a.kin, a.nout);
    float* sRed = sF + ((nf32 + 31) & ~31);
    int* sDst = reinterpret_cast<int*>(sRed + bwd_scratch_floats(MODE));
    int* sSrc = sDst + TILE_M;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sSrc + TILE_M);
    uint32_t* sTmem = reinterpret_cast<uint32_t*>(bars + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row = (warp & 3) * 32 + lane;
    const int half = warp >> 2;
    const uint32_t bar_w = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);

    if (tid == 0) {
        mbar_init(bar_w, 1);
        mbar_init(bar_mma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(sTmem), C::TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *sTmem;
    if (tid == 0) {
        const uint32_t wbytes = (uint32_t)C::W_CHUNKS * CHUNK_BYTES;
        const uint32_t fbytes = (uint32_t)nf32 * 4;
        mbar_expect_tx(bar_w, wbytes + fbytes);
        for (int c = 0; c < C::W_CHUNKS; ++c)
            bulk_g2s(smem_u32(sW + (size_t)c * CHUNK_BYTES), a.unit + (size_t)c * CHUNK_BYTES, CHUNK_BYTES, bar_w);
        bulk_g2s(smem_u32(sF), a.unit + wbytes, fbytes, bar_w);
    }
    grid_dep_trigger();  // the next kernel's prologue (weight image copy) may overlap this one's tail
    grid_dep_wait();
    bool weights_ready = false;
    uint32_t mma_parity = 0;
    const uint32_t sA_u32 = smem_u32(sA), sW_u32 = smem_u32(sW);
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(half * 64);

    // column sums kept in registers over the CTA's tiles; lane owns column half*64 + blk*32 + lane
    float acc_lnb[2] = {0.f, 0.f}, acc_lng[2] = {0.f, 0.f}, acc_b2[2] = {0.f, 0.f}, acc_b1[2] = {0.f, 0.f},
          acc_b0[2] = {0.f, 0.f};
    constexpr int NW = (MODE == ENC || MODE == DEC) ? 16 : 1;
    float acc_w[NW][2];   // ENC: dW0[:, k]; DEC: dW2[o, :]
    float acc_o[NW];      // DEC: db2[o] (lane 0 of the half-0 warps)
#pragma unroll
    for (int k = 0; k < NW; ++k) { acc_w[k][0] = acc_w[k][1] = 0.f; acc_o[k] = 0.f; }

    // forward MMA over `nchunks` K blocks: A slots [0, nchunks) x weight chunks [wc, wc + nchunks)
    auto issue_fwd = [&](int nchunks, int wc, bool accumulate) {
        for (int s = 0; s < nchunks; ++s) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint64_t ad = umma_desc(sA_u32 + (uint32_t)s * CHUNK_BYTES + k * 32);
                const uint64_t bd = umma_desc(sW_u32 + (uint32_t)(wc + s) * CHUNK_BYTES + k * 32);
                umma_i(tmem, ad, bd, IDESC, (accumulate || s > 0 || k > 0) ? 1u : 0u);
            }
        }
    };
    // dgrad: D[rows, 128 x nblocks] = dZ[rows, 128] W[128 (K = out), in]; W = weight chunks [wc, wc + 2 nblocks)
    auto issue_dgrad = [&](int nblocks, int wc) {
        for (int nb = 0; nb < nblocks; ++nb) {
            for (int s = 0; s < 2; ++s) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint64_t ad = umma_desc(sA_u32 + (uint32_t)s * CHUNK_BYTES + k * 32);
                    const uint64_t bd = umma_desc_mn(sW_u32 + (uint32_t)(wc + 2 * nb) * CHUNK_BYTES + (s * 4 + k) * 2048);
                    umma_i(tmem + (uint32_t)nb * 128, ad, bd, IDESC_B_MN, (s > 0 || k > 0) ? 1u : 0u);
                }
            }
        }
    };
    auto pre_issue_sync = [&]() {
        if (!weights_ready) { mbar_wait(bar_w, 0); weights_ready = true; }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
    };
    auto wait_mma = [&]() {
        mbar_wait(bar_mma, mma_parity);
        mma_parity ^= 1;
        tc_fence_after();
    };
    // thread's 64 values -> bf16 -> A chunk `half`, row `row`
    auto put_a = [&](const float* v) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            uint4 pk = make_uint4(pack_bf16(v[c * 8], v[c * 8 + 1]), pack_bf16(v[c * 8 + 2], v[c * 8 + 3]),
                                  pack_bf16(v[c * 8 + 4], v[c * 8 + 5]), pack_bf16(v[c * 8 + 6], v[c * 8 + 7]));
            *reinterpret_cast<uint4*>(sA + (size_t)half * CHUNK_BYTES + swz_off(row, c)) = pk;
        }
    };

    const int num_tiles = (a.rows + TILE_M - 1) / TILE_M;
    // EDGE: row indices of the CTA's next tile, fetched one tile ahead (threads < 128) so that a tile does not start with
    // two dependent global-memory latencies (indices, then the rows they point to)
    int d_next = -1, s_next = -1;
    bool have_next = false;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int row0 = tile * TILE_M;
        const int grow = row0 + row;
        const bool rvalid = grow < a.rows;
        // A chunks 0,1 (all 128 rows x 128 columns, bf16) -> global [rows,128], coalesced
        auto copy_out = [&](__nv_bfloat16* dst) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                if (row0 + r < a.rows)
                    *reinterpret_cast<uint4*>(dst + (size_t)(row0 + r) * LAT + kb * 64 + c * 8) =
                        *reinterpret_cast<const uint4*>(sA + (size_t)kb * CHUNK_BYTES + swz_off(r, c));
            }
        };
        uint32_t m1[2] = {0u, 0u}, m2[2] = {0u, 0u};  // ReLU masks of the two hidden layers (bit j: column half*64+j)
        // + bias, ReLU, mask, -> A operand
        auto hidden_epilogue = [&](float* v, const float* bias, uint32_t* mask) {
            mask[0] = mask[1] = 0u;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                v[j] = fmaxf(v[j] + bias[j], 0.f);
                if (v[j] > 0.f) mask[j >> 5] |= 1u << (j & 31);
            }
            put_a(v);
        };
        // v *= mask; column sums into acc; -> A operand
        auto dz_epilogue = [&](float* v, const uint32_t* mask, float* acc) {
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = ((mask[j >> 5] >> (j & 31)) & 1u) ? v[j] : 0.f;
#pragma unroll
            for (int blk = 0; blk < 2; ++blk) {
                float t[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = v[blk * 32 + i];
                acc[blk] += warp_colsum32(t, lane);
            }
        };
        // LayerNorm forward + backward on the row (two threads per row): v = pre-LN values + bias on entry,
        // dY on exit; d = upstream gradient of the LN output
        auto ln_backward = [&](float* v, float* d) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) s += v[j];
            sRed[row * 2 + half] = s;
            __syncthreads();
            const float mean = (sRed[row * 2] + sRed[row * 2 + 1]) * (1.0f / LAT);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) { const float dl = v[j] - mean; q = fmaf(dl, dl, q); }
            sRed[256 + row * 2 + half] = q;
            __syncthreads();
            const float var = (sRed[256 + row * 2] + sRed[256 + row * 2 + 1]) * (1.0f / LAT);
            const float rstd = 1.0f / sqrtf(var + 1e-5f);
            const float* lg = sF + 3 * LAT + half * 64;
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = (v[j] - mean) * rstd;  // xhat
#pragma unroll
            for (int blk = 0; blk < 2; ++blk) {
                float t[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = d[blk * 32 + i];
                acc_lnb[blk] += warp_colsum32(t, lane);
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = d[blk * 32 + i] * v[blk * 32 + i];
                acc_lng[blk] += warp_colsum32(t, lane);
            }
            float s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int j = 0; j < 64; ++j) { d[j] *= lg[j]; s1 += d[j]; s2 = fmaf(d[j], v[j], s2); }
            sRed[512 + row * 2 + half] = s1;
            sRed[768 + row * 2 + half] = s2;
            __syncthreads();
            s1 = (sRed[512 + row * 2] + sRed[512 + row * 2 + 1]) * (1.0f / LAT);
            s2 = (sRed[768 + row * 2] + sRed[768 + row * 2 + 1]) * (1.0f / LAT);
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] = rstd * (d[j] - s1 - v[j] * s2);  // dY
#pragma unroll
            for (int blk = 0; blk < 2; ++blk) {
                float t[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = v[blk * 32 + i];
                acc_b2[blk] += warp_colsum32(t, lane);
            }
        };
        auto run_mma = [&](auto&& fn) {
            pre_issue_sync();
            if (tid == 0) {
                tc_fence_after();
                fn();
                umma_commit(bar_mma);
            }
        };
        float v[64];
        auto load_acc = [&](int col) {
            tmem_ld32(taddr + col, v);
            tmem_ld32(taddr + col + 32, v + 32);
        };

        // =========================================================== forward recomputation
        if (MODE == EDGE) {
            if (tid < TILE_M) {
                const int p = row0 + tid;
                int d = d_next, s = s_next;
                if (!have_next) {
                    d = s = -1;
                    if (p < a.rows) {
                        d = a.dst_sorted[p];
                        s = a.src_sorted[p];
                        if ((unsigned)d >= (unsigned)a.N || (unsigned)s >= (unsigned)a.N) d = s = -1;
                    }
                }
                sDst[tid] = d;
                sSrc[tid] = s;
                const int pn = p + (int)gridDim.x * TILE_M;
                d_next = s_next = -1;
                if (pn < a.rows) {  // consumed at the top of the next tile
                    d_next = a.dst_sorted[pn];
                    s_next = a.src_sorted[pn];
                }
            }
            have_next = true;
            __syncthreads();
#pragma unroll 1
            for (int ph = 0; ph < 2; ++ph) {
                uint4 g[12];
#pragma unroll
                for (int it = 0; it < 12; ++it) {
                    const int idx = it * THREADS + tid;
                    const int c = idx & 7, r = (idx >> 3) & 127, kb = 3 * ph + (idx >> 10);
                    const int d = sDst[r];
                    g[it] = make_uint4(0, 0, 0, 0);
                    if (d >= 0) {
                        const __nv_bfloat16* src;
                        if (kb < 2) src = a.x_bf16 + (size_t)d * LAT + kb * 64;
                        else if (kb < 4) src = a.x_bf16 + (size_t)sSrc[r] * LAT + (kb - 2) * 64;
                        else src = a.e0 + (size_t)(row0 + r) * LAT + (kb - 4) * 64;
                        g[it] = *reinterpret_cast<const uint4*>(src + c * 8);
                    }
                }
#pragma unroll
                for (int it = 0; it < 12; ++it) {
                    const int idx = it * THREADS + tid;
                    const int c = idx & 7, r = (idx >> 3) & 127, kl = idx >> 10;
                    *reinterpret_cast<uint4*>(sA + (size_t)kl * CHUNK_BYTES + swz_off(r, c)) = g[it];
                }
                run_mma([&]() { issue_fwd(3, 3 * ph, ph > 0); });
                wait_mma();
            }
        } else if (MODE == NODE) {
            uint4 ga[8], gx[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                const int n = row0 + r;
                ga[it] = gx[it] = make_uint4(0, 0, 0, 0);
                if (n < a.rows) {
                    ga[it] = *reinterpret_cast<const uint4*>(a.agg_bf16 + (size_t)n * LAT + kb * 64 + c * 8);
                    gx[it] = *reinterpret_cast<const uint4*>(a.x_bf16 + (size_t)n * LAT + kb * 64 + c * 8);
                }
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                *reinterpret_cast<uint4*>(sA + (size_t)kb * CHUNK_BYTES + swz_off(r, c)) = ga[it];
                *reinterpret_cast<uint4*>(sA + (size_t)(kb + 2) * CHUNK_BYTES + swz_off(r, c)) = gx[it];
            }
            run_mma([&]() { issue_fwd(4, 0, false); });
            wait_mma();
        } else if (MODE == DEC) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int idx = it * THREADS + tid;
                const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
                const int n = row0 + r;
                uint4 w = make_uint4(0, 0, 0, 0);
                if (n < a.rows) w = *reinterpret_cast<const uint4*>(a.x_bf16 + (size_t)n * LAT + kb * 64 + c * 8);
                *reinterpret_cast<uint4*>(sA + (size_t)kb * CHUNK_BYTES + swz_off(r, c)) = w;
            }
            run_mma([&]() { issue_fwd(2, 0, false); });
            wait_mma();
        }
        float xin[16];
        long long srow = -1;
        if (MODE == ENC) {
            // first Linear on the CUDA cores (fp32), as in the forward
            if (!weights_ready) { mbar_wait(bar_w, 0); weights_ready = true; }
            const float* W0 = sF + 5 * LAT;
            const float* b0 = sF;
            const int kin = a.kin;
            srow = rvalid ? (a.perm ? (long long)a.perm[grow] : (long long)grow) : -1;
#pragma unroll
            for (int k = 0; k < 16; ++k) xin[k] = (srow >= 0 && k < kin) ? a.in[srow * kin + k] : 0.f;
#pragma unroll 1
            for (int c = 0; c < 8; ++c) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int o = half * 64 + c * 8 + j;
                    float s = b0[o];
#pragma unroll
                    for (int k = 0; k < 16; ++k)
                        if (k < kin) s = fmaf(W0[o * kin + k], xin[k], s);
                    v[c * 8 + j] = s;
                }
            }
            m1[0] = m1[1] = 0u;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                v[j] = fmaxf(v[j], 0.f);
                if (v[j] > 0.f) m1[j >> 5] |= 1u << (j & 31);
            }
            put_a(v);
        } else {
            // epilogue of the first tensor-core layer: H1
            load_acc(0);
            hidden_epilogue(v, sF + half * 64, m1);
        }
        // ---- second layer (weights: EDGE 6..7, NODE 4..5, ENC 0..1, DEC 2..3)
        constexpr int WC1 = MODE == EDGE ? 6 : MODE == NODE ? 4 : MODE == ENC ? 0 : 2;
        run_mma([&]() { issue_fwd(2, WC1, false); });
        copy_out(a.sH1);
        wait_mma();
        __syncthreads();
        load_acc(0);
        if (MODE != DEC) {
            hidden_epilogue(v, sF + LAT + half * 64, m2);
            // ---- third layer + LayerNorm
            constexpr int WC2 = WC1 + 2;
            run_mma([&]() { issue_fwd(2, WC2, false); });
            copy_out(a.sH2);
            wait_mma();
            __syncthreads();
            load_acc(0);
            const float* b2 = sF + 2 * LAT + half * 64;
#pragma unroll
            for (int j = 0; j < 64; ++j) v[j] += b2[j];
            float d[64];
            {
                const float* dsrc = nullptr;
                if (MODE == EDGE) { const int dd = sDst[row]; if (dd >= 0) dsrc = a.dout + (size_t)dd * LAT + half * 64; }
                else if (rvalid) dsrc = a.dout + (size_t)grow * LAT + half * 64;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (dsrc) t = reinterpret_cast<const float4*>(dsrc)[c];
                    d[c * 4] = t.x; d[c * 4 + 1] = t.y; d[c * 4 + 2] = t.z; d[c * 4 + 3] = t.w;
                }
            }
            ln_backward(v, d);
            put_a(v);  // dY
            // =========================================================== dgrad chain
            run_mma([&]() { issue_dgrad(1, WC2); });
            copy_out(a.sdY);
            wait_mma();
            __syncthreads();
            load_acc(0);
            dz_epilogue(v, m2, acc_b1);
            put_a(v);  // dZ2
        } else {
            // DEC: Z2 -> H2 (fp32), last Linear and its backward on the CUDA cores
            const float* b1 = sF + LAT + half * 64;
            const float* W2 = sF + 2 * LAT;
            m2[0] = m2[1] = 0u;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
                v[j] = fmaxf(v[j] + b1[j], 0.f);
                if (v[j] > 0.f) m2[j >> 5] |= 1u << (j & 31);
            }
            float dyr[8];
#pragma unroll
            for (int o = 0; o < 8; ++o) dyr[o] = (rvalid && o < a.nout) ? a.dy[(size_t)grow * a.nout + o] : 0.f;
            float dh[64];
#pragma unroll
            for (int j = 0; j < 64; ++j) dh[j] = 0.f;
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                if (o < a.nout) {
                    const float* w = W2 + o * LAT + half * 64;
#pragma unroll
                    for (int j = 0; j < 64; ++j) dh[j] = fmaf(dyr[o], w[j], dh[j]);
#pragma unroll
                    for (int blk = 0; blk < 2; ++blk) {
                        float t[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) t[i] = dyr[o] * v[blk * 32 + i];
                        acc_w[o][blk] += warp_colsum32(t, lane);
                    }
                    if (half == 0) {
                        float s = dyr[o];
#pragma unroll
                        for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, off);
                        acc_o[o] += s;
                    }
                }
            }
            dz_epilogue(dh, m2, acc_b1);
            put_a(dh);  // dZ2
        }
        // ---- dgrad through the second layer
        run_mma([&]() { issue_dgrad(1, WC1); });
        copy_out(a.sdZ2);
        wait_mma();
        __syncthreads();
        load_acc(0);
        dz_epilogue(v, m1, acc_b0);  // dZ1
        if (MODE == ENC) {
            // first Linear backward on the CUDA cores: dW0[c, k] += dZ1[r, c] in[r, k];  din[r, k] = sum_c dZ1[r, c] W0[c, k]
            const int kin = a.kin;
            const float* W0 = sF + 5 * LAT;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                if (k < kin) {
#pragma unroll
                    for (int blk = 0; blk < 2; ++blk) {
                        float t[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) t[i] = v[blk * 32 + i] * xin[k];
                        acc_w[k][blk] += warp_colsum32(t, lane);
                    }
                }
            }
            if (a.din) {
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    if (k < kin) {
                        float s = 0.f;
#pragma unroll
                        for (int j = 0; j < 64; ++j) s = fmaf(v[j], W0[(half * 64 + j) * kin + k], s);
                        sRed[(row * 2 + half) * 16 + k] = s;
                    }
                }
                __syncthreads();
                if (half == 0 && srow >= 0)
                    for (int k = 0; k < kin; ++k)
                        a.din[srow * kin + k] = sRed[(row * 2) * 16 + k] + sRed[(row * 2 + 1) * 16 + k];
            }
            tc_fence_before();
            __syncthreads();
            continue;
        }
        put_a(v);
        run_mma([&]() { issue_dgrad(C::IN_BLOCKS, 0); });
        copy_out(a.sdZ1);
        wait_mma();
        __syncthreads();
        if (MODE == DEC) {
            load_acc(0);
            if (rvalid) {
                float4* o = reinterpret_cast<float4*>(a.dx + (size_t)grow * LAT + half * 64);
#pragma unroll
                for (int c = 0; c < 16; ++c) o[c] = make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
            }
        } else if (MODE == NODE) {
            load_acc(0);
            if (rvalid) {
                float4* o = reinterpret_cast<float4*>(a.dagg + (size_t)grow * LAT + half * 64);
#pragma unroll
                for (int c = 0; c < 16; ++c) o[c] = make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
            }
            load_acc(128);
            if (rvalid) {
                float4* o = reinterpret_cast<float4*>(a.dx + (size_t)grow * LAT + half * 64);
#pragma unroll
                for (int c = 0; c < 16; ++c)  // red.global.add.v4.f32: no load on the tile's critical path (one adder per element
                    atomicAdd(&o[c], make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]));  // and kernel: same sum)
            }
        } else if (MODE == EDGE) {
            float* st = reinterpret_cast<float*>(sA) + warp * (32 * 36);  // per-warp staging: pitch 33 here, 36 below
            const int* wd = sDst + (warp & 3) * 32;
            const int* wsrc = sSrc + (warp & 3) * 32;
            if (tid < TILE_M) {  // the next tile's indices have long arrived: validate, pull its rows towards L2
                if ((unsigned)d_next >= (unsigned)a.N || (unsigned)s_next >= (unsigned)a.N) d_next = s_next = -1;
                if (d_next >= 0) {
                    const char* pd = reinterpret_cast<const char*>(a.x_bf16 + (size_t)d_next * LAT);
                    const char* ps = reinterpret_cast<const char*>(a.x_bf16 + (size_t)s_next * LAT);
                    const char* pe = reinterpret_cast<const char*>(a.e0 + (size_t)(row0 + tid + (size_t)gridDim.x * TILE_M) * LAT);
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pd));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pd + 128));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(ps));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(ps + 128));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pe));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pe + 128));
                }
            }
            // d x[dst]: segmented sum over the warp's dst-sorted rows, one red.add per (segment, column).  Bit i of `ends`:
            // row i is the last of its segment (one ballot per tile instead of a compare + branch per row and column)
            const int my_d = wd[lane];
            const int nx_d = __shfl_down_sync(0xffffffffu, my_d, 1);
            const unsigned ends = __ballot_sync(0xffffffffu, lane == 31 || nx_d != my_d);
            load_acc(0);
#pragma unroll
            for (int rd = 0; rd < 2; ++rd) {
#pragma unroll
                for (int j = 0; j < 32; ++j) st[lane * 33 + j] = v[rd * 32 + j];
                __syncwarp();
                const int col = half * 64 + rd * 32 + lane;
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s += st[i * 33 + lane];
                    if ((ends >> i) & 1u) {
                        const int dd = __shfl_sync(0xffffffffu, my_d, i);
                        if (dd >= 0) atomicAdd(a.dx + (size_t)dd * LAT + col, s);
                        s = 0.f;
                    }
                }
                __syncwarp();
            }
            // d x[src]: rows go out as 16-byte vector red.adds (lane = 4 columns of one of 4 rows per step: 128 bytes per
            // row and warp instruction), 8 steps per 32 columns instead of 32 scalar ones
            load_acc(128);
            float* st4 = st;  // same per-warp region, pitch 36 floats: 16-byte aligned rows
#pragma unroll
            for (int rd = 0; rd < 2; ++rd) {
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<float4*>(st4 + lane * 36 + j) = make_float4(v[rd * 32 + j], v[rd * 32 + j + 1], v[rd * 32 + j + 2], v[rd * 32 + j + 3]);
                __syncwarp();
                const int col = half * 64 + rd * 32 + (lane & 7) * 4;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = i * 4 + (lane >> 3);
                    const int sr = wsrc[r];
                    const float4 t = *reinterpret_cast<const float4*>(st4 + r * 36 + (lane & 7) * 4);
                    if (sr >= 0) atomicAdd(reinterpret_cast<float4*>(a.dx + (size_t)sr * LAT + col), t);
                }
                __syncwarp();
            }
            // d e_0 (2^m already inside the packed weights)
            load_acc(256);
            if (rvalid && sDst[row] >= 0) {
                float4* o = reinterpret_cast<float4*>(a.de0 + (size_t)grow * LAT + half * 64);
#pragma unroll
                for (int c = 0; c < 16; ++c)  // red.global.add.v4.f32: no load on the tile's critical path (one adder per element
                    atomicAdd(&o[c], make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]));  // and kernel: same sum)
            }
        }
        tc_fence_before();
        __syncthreads();
    }

    // ---- flush the register column sums: one atomic per column per warp
    {
        const MlpOff o(MODE == EDGE ? 3 * LAT : MODE == NODE ? 2 * LAT : MODE == ENC ? a.kin : LAT,
                       MODE == DEC ? a.nout : LAT);
#pragma unroll
        for (int blk = 0; blk < 2; ++blk) {
            const int col = half * 64 + blk * 32 + lane;
            atomicAdd(a.g + o.b0 + col, acc_b0[blk]);
            atomicAdd(a.g + o.b1 + col, acc_b1[blk]);
            if (MODE != DEC) {
                atomicAdd(a.g + o.b2 + col, acc_b2[blk]);
                atomicAdd(a.g + o.lng + col, acc_lng[blk]);
                atomicAdd(a.g + o.lnb + col, acc_lnb[blk]);
            }
            if (MODE == ENC) {
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (k < a.kin) atomicAdd(a.g + o.W0 + (size_t)col * a.kin + k, acc_w[k][blk]);
            }
            if (MODE == DEC) {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < a.nout) atomicAdd(a.g + o.W2 + (size_t)k * LAT + col, acc_w[k][blk]);
            }
        }
        if (MODE == DEC && half == 0 && lane == 0) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (k < a.nout) atomicAdd(a.g + o.b2 + k, acc_o[k]);
        }
    }
    if (!weights_ready && tid == 0) mbar_wait(bar_w, 0);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, C::TMEM);
}

// ---- weight gradients ---------------------------------------------------------------------------
struct WJob {
    const __nv_bfloat16* dZ;   // [rows,128]
    const __nv_bfloat16* A;    // [rows,128] or, with `gather`, [N,128] indexed by gather[row]
    const int* gather;
    float* out;                // dW[128, ldo] block: out[m * ldo + n] += scale * sum_r dZ[r, m] A[r, n]
    int ldo;
    float scale;
};
constexpr int MAX_WJOBS = 6;
struct WArgs {
    int rows, N, njobs;
    WJob job[MAX_WJOBS];
};
constexpr int WGRAD_STAGE_BYTES = 4 * CHUNK_BYTES;  // dZ (2 chunks) | A (2 chunks)
constexpr size_t WGRAD_SMEM = 1024 + 2 * WGRAD_STAGE_BYTES + 64;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}

__global__ void __launch_bounds__(THREADS, 1) gnn_wgrad_kernel(const __grid_constant__ WArgs a) {
    const int num_tiles = (a.rows + TILE_M - 1) / TILE_M;
    grid_dep_trigger();
    if ((int)blockIdx.x >= num_tiles) return;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * WGRAD_STAGE_BYTES);
    uint32_t* sTmem = reinterpret_cast<uint32_t*>(bars + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const WJob& job = a.job[blockIdx.y];
    const uint32_t bar[2] = {smem_u32(&bars[0]), smem_u32(&bars[1])};
    if (tid == 0) {
        mbar_init(bar[0], 1);
        mbar_init(bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(sTmem), 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *sTmem;
    const uint32_t s_u32 = smem_u32(smem);

    auto load = [&](int stage, int tile) {
        const int row0 = tile * TILE_M;
        const uint32_t base = s_u32 + (uint32_t)stage * WGRAD_STAGE_BYTES;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * THREADS + tid;
            const int c = idx & 7, r = (idx >> 3) & 127, kb = idx >> 10;
            const int gr = row0 + r;
            const bool ok = gr < a.rows;
            const size_t zr = ok ? (size_t)gr : 0;
            cp_async16(base + (uint32_t)kb * CHUNK_BYTES + swz_off(r, c), job.dZ + zr * LAT + kb * 64 + c * 8, ok);
            size_t ar = zr;
            bool aok = ok;
            if (job.gather) {
                const int gi = ok ? job.gather[gr] : -1;
                aok = (unsigned)gi < (unsigned)a.N;
                ar = aok ? (size_t)gi : 0;
            }
            cp_async16(base + (uint32_t)(2 + kb) * CHUNK_BYTES + swz_off(r, c), job.A + ar * LAT + kb * 64 + c * 8, aok);
        }
    };
    grid_dep_wait();
    int it = 0;
    load(0, blockIdx.x);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int s = it & 1;
        const int next = tile + gridDim.x;
        if (next < num_tiles) {
            if (it >= 1) {  // stage s^1 was read by the MMAs of iteration it-1
                mbar_wait(bar[s ^ 1], (uint32_t)(((it - 1) >> 1) & 1));
            }
            load(s ^ 1, next);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t base = s_u32 + (uint32_t)s * WGRAD_STAGE_BYTES;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const uint64_t ad = umma_desc_mn(base + ks * 2048);
                const uint64_t bd = umma_desc_mn(base + 2 * CHUNK_BYTES + ks * 2048);
                umma_i(tmem, ad, bd, IDESC_AB_MN, (it > 0 || ks > 0) ? 1u : 0u);
            }
            umma_commit(bar[s]);
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    {
        const int last = it - 1;
        mbar_wait(bar[last & 1], (uint32_t)((last >> 1) & 1));
        tc_fence_after();
    }
    // epilogue: TMEM lane = out index m, columns = in index n
    {
        const int m = (warp & 3) * 32 + lane;
        const int half = warp >> 2;
        const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(half * 64);
        float v[64];
        tmem_ld32(taddr, v);
        tmem_ld32(taddr + 32, v + 32);
        float* o = job.out + (size_t)m * job.ldo + half * 64;
#pragma unroll
        for (int j = 0; j < 64; ++j) atomicAdd(o + j, v[j] * job.scale);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 128);
}

// ---- host side -----------------------------------------------------------------------------------
struct BwdWs {
    float *dx, *dagg, *de0;
    __nv_bfloat16 *sH1, *sH2, *sdY, *sdZ2, *sdZ1;
    size_t bytes;
    static BwdWs carve(void* ws, int N, int E) {
        BwdWs w;
        Carver c(ws);
        const size_t R = (size_t)(N > E ? N : E);
        w.dx = c.take<float>((size_t)N * LAT);
        w.dagg = c.take<float>((size_t)N * LAT);
        w.de0 = c.take<float>((size_t)E * LAT);
        w.sH1 = c.take<__nv_bfloat16>(R * LAT);
        w.sH2 = c.take<__nv_bfloat16>(R * LAT);
        w.sdY = c.take<__nv_bfloat16>(R * LAT);
        w.sdZ2 = c.take<__nv_bfloat16>(R * LAT);
        w.sdZ1 = c.take<__nv_bfloat16>(R * LAT);
        w.bytes = c.total();
        return w;
    }
};

static int sm_count() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

template <int MODE>
static int launch_bwd(const BArgs& a, cudaStream_t stream, const char* name) {
    static unsigned long long configured = 0;  // per-device opt-in (common.cuh)
    const size_t smem = bwd_smem_bytes(MODE, a.kin, a.nout);
    CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_bwd_kernel<MODE>, 232448 - 1024, &configured));
    if (smem > 232448 - 1024) return CSB_ERR_UNSUPPORTED;
    const int tiles = (a.rows + TILE_M - 1) / TILE_M;
    if (tiles == 0) return 0;
    const int grid = tiles < sm_count() ? tiles : sm_count();
    CSB_LAUNCH(name, stream, launch_pdl(gnn_bwd_kernel<MODE>, dim3(grid), dim3(THREADS), smem, stream, a));
    return 0;
}

static int launch_wgrad(WArgs& w, cudaStream_t stream) {
    static unsigned long long configured = 0;
    CSB_CUDA_TRY(opt_in_dynamic_smem(gnn_wgrad_kernel, (int)WGRAD_SMEM, &configured));
    const int tiles = (w.rows + TILE_M - 1) / TILE_M;
    if (tiles == 0 || w.njobs == 0) return 0;
    int per_job = sm_count() / w.njobs;
    if (per_job < 1) per_job = 1;
    if (per_job > tiles) per_job = tiles;
    dim3 grid(per_job, w.njobs);
    CSB_LAUNCH("gnn_wgrad", stream, launch_pdl(gnn_wgrad_kernel, grid, dim3(THREADS), WGRAD_SMEM, stream, w));
    return 0;
}

}  // namespace gnn
}  // namespace csb

using namespace csb;
using namespace csb::gnn;

extern "C" {

size_t csb_gnn_backward_workspace_bytes(const csb_gnn_dims* d, int N, int E) {
    if (check_dims(d) || N < 0 || E < 0) return 0;
    return BwdWs::carve(nullptr, N, E).bytes;
}

int csb_gnn_backward(const csb_gnn_dims* d, const void* packed, int N, int E, const float* x,
                     const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                     const int32_t* perm, const void* state, size_t state_bytes, const float* dL_dy,
                     float* grad_flat, float* dL_dx, float* dL_dedge_features, void* ws, size_t ws_bytes,
                     void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (N < 0 || E < 0) return CSB_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!packed || !x || !state || !dL_dy || !grad_flat || !ws) return CSB_ERR_BAD_ARG;
    if (E > 0 && (!edge_features || !dst_sorted || !src_sorted || !perm)) return CSB_ERR_BAD_ARG;
    const int S = d->nmessage_passing_steps;
    TrainState st = TrainState::carve(const_cast<void*>(state), N, E, S);
    if (state_bytes < st.bytes) return CSB_ERR_WORKSPACE;
    BwdWs w = BwdWs::carve(ws, N, E);
    if (ws_bytes < w.bytes) return CSB_ERR_WORKSPACE;
    cudaStream_t stream = (cudaStream_t)stream_;
    PackTable t;
    build_table(d, t, nullptr, nullptr);
    const uint8_t* pk = (const uint8_t*)packed;
    const size_t NL = (size_t)N * LAT;

    if (E > 0) CSB_CUDA_TRY(cudaMemsetAsync(w.de0, 0, (size_t)E * LAT * sizeof(float), stream));
    auto base = [&](int rows) {
        BArgs a{};
        a.rows = rows; a.N = N; a.kin = 0; a.nout = LAT;
        a.sH1 = w.sH1; a.sH2 = w.sH2; a.sdY = w.sdY; a.sdZ2 = w.sdZ2; a.sdZ1 = w.sdZ1;
        return a;
    };
    auto job = [](const __nv_bfloat16* dZ, const __nv_bfloat16* A, const int* gather, float* out, int ldo, float scale) {
        WJob j;
        j.dZ = dZ; j.A = A; j.gather = gather; j.out = out; j.ldo = ldo; j.scale = scale;
        return j;
    };
    // ---- decoder
    {
        const UnitDesc& u = t.u[t.n - 1];
        const MlpOff o(LAT, d->nnode_out);
        BArgs a = base(N);
        a.unit = pk + u.dst; a.nout = d->nnode_out; a.x_bf16 = st.x_stack + (size_t)S * NL; a.dy = dL_dy; a.dx = w.dx;
        a.g = grad_flat + u.src;
        rc = launch_bwd<DEC>(a, stream, "gnn_bwd_decode");
        if (rc) return rc;
        WArgs wa{};
        wa.rows = N; wa.N = N; wa.njobs = 2;
        wa.job[0] = job(w.sdZ2, w.sH1, nullptr, a.g + o.W1, LAT, 1.f);
        wa.job[1] = job(w.sdZ1, a.x_bf16, nullptr, a.g + o.W0, LAT, 1.f);
        rc = launch_wgrad(wa, stream);
        if (rc) return rc;
    }
    // ---- processor, last step first
    for (int m = S - 1; m >= 0; --m) {
        const __nv_bfloat16* xm = st.x_stack + (size_t)m * NL;
        {
            const UnitDesc& u = t.u[3 + 2 * m];
            const MlpOff o(2 * LAT, LAT);
            BArgs a = base(N);
            a.unit = pk + u.dst; a.x_bf16 = xm; a.agg_bf16 = st.agg_stack + (size_t)m * NL;
            a.dout = w.dx; a.dx = w.dx; a.dagg = w.dagg; a.g = grad_flat + u.src;
            rc = launch_bwd<NODE>(a, stream, "gnn_bwd_node");
            if (rc) return rc;
            WArgs wa{};
            wa.rows = N; wa.N = N; wa.njobs = 4;
            wa.job[0] = job(w.sdY, w.sH2, nullptr, a.g + o.W2, LAT, 1.f);
            wa.job[1] = job(w.sdZ2, w.sH1, nullptr, a.g + o.W1, LAT, 1.f);
            wa.job[2] = job(w.sdZ1, a.agg_bf16, nullptr, a.g + o.W0, 2 * LAT, 1.f);
            wa.job[3] = job(w.sdZ1, xm, nullptr, a.g + o.W0 + LAT, 2 * LAT, 1.f);
            rc = launch_wgrad(wa, stream);
            if (rc) return rc;
        }
        if (E > 0) {
            const UnitDesc& u = t.u[2 + 2 * m];
            const MlpOff o(3 * LAT, LAT);
            BArgs a = base(E);
            a.unit = pk + u.dst; a.x_bf16 = xm; a.e0 = st.e0; a.dst_sorted = dst_sorted; a.src_sorted = src_sorted;
            a.dout = w.dagg; a.dx = w.dx; a.de0 = w.de0; a.g = grad_flat + u.src;
            rc = launch_bwd<EDGE>(a, stream, "gnn_bwd_edge");
            if (rc) return rc;
            WArgs wa{};
            wa.rows = E; wa.N = N; wa.njobs = 5;
            wa.job[0] = job(w.sdY, w.sH2, nullptr, a.g + o.W2, LAT, 1.f);
            wa.job[1] = job(w.sdZ2, w.sH1, nullptr, a.g + o.W1, LAT, 1.f);
            wa.job[2] = job(w.sdZ1, xm, dst_sorted, a.g + o.W0, 3 * LAT, 1.f);
            wa.job[3] = job(w.sdZ1, xm, src_sorted, a.g + o.W0 + LAT, 3 * LAT, 1.f);
            wa.job[4] = job(w.sdZ1, st.e0, nullptr, a.g + o.W0 + 2 * LAT, 3 * LAT, u.e_scale);
            rc = launch_wgrad(wa, stream);
            if (rc) return rc;
        }
    }
    // ---- encoders
    {
        const UnitDesc& u = t.u[0];
        const MlpOff o(d->nnode_in, LAT);
        BArgs a = base(N);
        a.unit = pk + u.dst; a.kin = d->nnode_in; a.in = x; a.dout = w.dx; a.din = dL_dx; a.g = grad_flat + u.src;
        rc = launch_bwd<ENC>(a, stream, "gnn_bwd_encode");
        if (rc) return rc;
        WArgs wa{};
        wa.rows = N; wa.N = N; wa.njobs = 2;
        wa.job[0] = job(w.sdY, w.sH2, nullptr, a.g + o.W2, LAT, 1.f);
        wa.job[1] = job(w.sdZ2, w.sH1, nullptr, a.g + o.W1, LAT, 1.f);
        rc = launch_wgrad(wa, stream);
        if (rc) return rc;
    }
    if (E > 0) {
        const UnitDesc& u = t.u[1];
        const MlpOff o(d->nedge_in, LAT);
        BArgs a = base(E);
        a.unit = pk + u.dst; a.kin = d->nedge_in; a.in = edge_features; a.perm = perm; a.dout = w.de0;
        a.din = dL_dedge_features; a.g = grad_flat + u.src;
        rc = launch_bwd<ENC>(a, stream, "gnn_bwd_encode");
        if (rc) return rc;
        WArgs wa{};
        wa.rows = E; wa.N = N; wa.njobs = 2;
        wa.job[0] = job(w.sdY, w.sH2, nullptr, a.g + o.W2, LAT, 1.f);
        wa.job[1] = job(w.sdZ2, w.sH1, nullptr, a.g + o.W1, LAT, 1.f);
        rc = launch_wgrad(wa, stream);
        if (rc) return rc;
    }
    return 0;
}

}  // extern "C"
