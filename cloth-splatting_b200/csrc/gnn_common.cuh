// Shared definitions of the meshnet GNN kernels (gnn.cu forward, gnn_backward.cu): unit images, pack table,
// tcgen05 / TMEM / mbarrier / bulk-copy PTX wrappers.
#pragma once
#include <cuda_bf16.h>
#include "common.cuh"

namespace csb {
namespace gnn {

constexpr int LAT = 128;  // latent width == hidden width (only configuration supported)
constexpr int TILE_M = 128;
constexpr int CHUNK_BYTES = 128 * 128;  // 128 rows x 64 bf16, one swizzle-128B K block
constexpr int THREADS = 256;
constexpr int TMEM_COLS = 128;

enum Mode { ENC = 0, EDGE = 1, NODE = 2, DEC = 3 };

// ---- unit images ------------------------------------------------------------------------------
// bf16 chunks first, then an fp32 block.
//   EDGE: chunks W0(6) W1(2) W2(2); f32 b0 b1 b2 ln_g ln_b
//   NODE: chunks W0(4) W1(2) W2(2); f32 b0 b1 b2 ln_g ln_b
//   ENC : chunks W1(2) W2(2);       f32 b0 b1 b2 ln_g ln_b W0[128 x kin]
//   DEC : chunks W0(2) W1(2);       f32 b0 b1 W2[nout x 128] b2[nout]
__host__ __device__ inline int unit_chunks(int mode) { return mode == EDGE ? 10 : mode == NODE ? 8 : 4; }
__host__ __device__ inline int unit_f32(int mode, int kin, int nout) {
    if (mode == ENC) return 5 * LAT + LAT * kin;
    if (mode == DEC) return 2 * LAT + nout * LAT + ((nout + 3) & ~3);
    return 5 * LAT;
}
__host__ __device__ inline size_t unit_bytes(int mode, int kin, int nout) {
    const size_t b = (size_t)unit_chunks(mode) * CHUNK_BYTES + (size_t)unit_f32(mode, kin, nout) * 4;
    return (b + 127) / 128 * 128;
}
// floats of one MLP (+LN) in the flat parameter buffer: W0[128,k0] b0 W1[128,128] b1 W2[n2,128] b2 (g, beta)
__host__ __device__ inline size_t mlp_floats(int k0, int n2, bool ln) {
    return (size_t)LAT * k0 + LAT + (size_t)LAT * LAT + LAT + (size_t)n2 * LAT + n2 + (ln ? 2 * LAT : 0);
}

struct UnitDesc {
    unsigned long long src;  // float offset in the flat parameter buffer
    unsigned long long dst;  // byte offset in the packed buffer
    int mode, k0, n2;
    float e_scale;           // EDGE: factor of the columns >= 256 of W0 (2^m)
};
constexpr int MAX_UNITS = 2 + 2 * 30 + 1;
struct PackTable {
    int n;
    UnitDesc u[MAX_UNITS];
};

__device__ __forceinline__ uint32_t swz_off(int row, int c16) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((c16 ^ (row & 7)) << 4));
}

// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    while (true) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
        if (clock64() - t0 > 4000000000ll) __trap();  // ~2 s: never hang the device on a protocol bug
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// K-major, swizzle-128B shared-memory matrix descriptor (LBO = 1, SBO = 1024 B, version 1)
// true in exactly one (the lowest active) lane of a converged warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xFFFFFFFF;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 x bf16 -> fp32, M = 128, N = 128, K = 16
constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 columns of fp32: thread l of the warp gets row (lane base + l), columns [col, col+32)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// the same without the wait: several loads in flight, then ONE tmem_ld_wait() before the values are used
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// `n32` consecutive 32-column groups of the thread's TMEM lane, all loads issued before the single wait
template <int N32>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, float* v) {
    uint32_t r[N32 * 32];
#pragma unroll
    for (int i = 0; i < N32; ++i) tmem_ld32_nowait(taddr + i * 32, r + i * 32);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < N32 * 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}

struct Dims {
    int node_in, edge_in, node_out, steps;
};
static int check_dims(const csb_gnn_dims* d) {
    if (!d) return CSB_ERR_BAD_ARG;
    if (d->latent_dim != LAT || d->mlp_hidden_dim != LAT || d->nmlp_layers != 2) return CSB_ERR_UNSUPPORTED;
    if (d->nnode_in < 1 || d->nnode_in > 16 || d->nedge_in < 1 || d->nedge_in > 16) return CSB_ERR_UNSUPPORTED;
    if (d->nnode_out < 1 || d->nnode_out > 8) return CSB_ERR_UNSUPPORTED;
    if (d->nmessage_passing_steps < 0 || d->nmessage_passing_steps > 30) return CSB_ERR_UNSUPPORTED;
    return 0;
}

static void build_table(const csb_gnn_dims* d, PackTable& t, size_t* total_floats, size_t* total_bytes) {
    size_t src = 0, dst = 0;
    int n = 0;
    auto add = [&](int mode, int k0, int n2, bool ln, float scale) {
        UnitDesc& u = t.u[n++];
        u.src = src; u.dst = dst; u.mode = mode; u.k0 = k0; u.n2 = n2; u.e_scale = scale;
        src += mlp_floats(k0, n2, ln);
        dst += unit_bytes(mode, k0, n2);
    };
    add(ENC, d->nnode_in, LAT, true, 1.f);
    add(ENC, d->nedge_in, LAT, true, 1.f);
    float sc = 1.f;
    for (int m = 0; m < d->nmessage_passing_steps; ++m) {
        add(EDGE, 3 * LAT, LAT, true, sc);
        add(NODE, 2 * LAT, LAT, true, 1.f);
        sc *= 2.f;
    }
    add(DEC, LAT, d->nnode_out, false, 1.f);
    t.n = n;
    if (total_floats) *total_floats = src;
    if (total_bytes) *total_bytes = dst;
}

// What a training forward keeps for csb_gnn_backward: the bf16 operands every step consumed (the backward
// recomputes the MLP activations from them tile by tile instead of storing [E,128] x 3 x steps).
struct TrainState {
    float* x_f32;               // [N,128] fp32 master of the node latents (final value after the forward)
    float* agg;                 // [N,128] scatter target (zero after the forward)
    __nv_bfloat16* e0;          // [E,128] encoded edge latents, dst-sorted order
    __nv_bfloat16* x_stack;     // [steps+1][N,128] node latents entering step m (m = steps: decoder input)
    __nv_bfloat16* agg_stack;   // [steps][N,128] bf16 aggregated messages the node MLP of step m consumed
    size_t bytes;
    static TrainState carve(void* ws, int N, int E, int steps) {
        TrainState w;
        Carver c(ws);
        w.x_f32 = c.take<float>((size_t)N * LAT);
        w.agg = c.take<float>((size_t)N * LAT);
        w.e0 = c.take<__nv_bfloat16>((size_t)E * LAT);
        w.x_stack = c.take<__nv_bfloat16>((size_t)(steps + 1) * N * LAT);
        w.agg_stack = c.take<__nv_bfloat16>((size_t)steps * N * LAT);
        w.bytes = c.total();
        return w;
    }
};

}  // namespace gnn
}  // namespace csb
