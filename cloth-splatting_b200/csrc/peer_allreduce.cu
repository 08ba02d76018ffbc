// The one collective of the hot path: the sum over GPUs of the flat parameter-gradient buffer
// after a view-sharded rasterizer step or a graph-sharded GNN rollout.  The reference is single-GPU
// (utils/general_utils.py:136) and loops the views sequentially (scene_reconstruction/train_utils.py:259).
//
// One kernel per rank over NVLink peer memory (no NCCL on the data path): every rank's gradient bucket lives in a
// cudaMalloc'ed "peer buffer" whose IPC handle the other ranks have opened, so a rank can load from and store to
// every bucket of the node.  Two-shot, in place:
//   0. arrive   flag[peer][me] = seq (release.sys); wait until every peer's flag for me shows seq
//   1. reduce   rank r owns slice r: loads its float4s from all W buckets (W loads in flight per thread), adds them
//               in rank order 0..W-1 (the same fixed order on every rank -> bit-identical, run-to-run
//               deterministic sums) and stores the result into slice r of all W buckets (push all-gather)
//   2. depart   system fence, last block of the grid: done[peer][me] = seq; wait for every peer's done flag, so
//               that when the kernel retires on this rank all slices have landed here and nobody still reads it.
// Payloads are small (2.4 MB of Gaussian gradients at C2, 9.3 MB of GNN parameters at C5): the cost is three
// NVLink round trips, not bandwidth -- which is why a latency-optimised single kernel beats a ring.
#include <stddef.h>
#include <string.h>
#include "common.cuh"

namespace csb {
namespace peer {

constexpr int MAX_WORLD = 8;
constexpr int PEER_THREADS = 128;  // small blocks: the slice spreads over many SMs, few remote requests queue per SM
constexpr int HEADER_BYTES = 1024;  // flags in front of the payload
struct Header {
    uint32_t arrive[16];   // [src rank] written by src at the start of call `seq`
    uint32_t done[16];     // [src rank] written by src when its slice has been pushed everywhere
    uint32_t blocks_done;  // local grid counter
    uint32_t timed_out;    // sequence number of a call in which a peer never showed up (0 = none): csb_peer_timed_out
    uint32_t pad[30];
    // [64] self-timing of the calls so far (ns, %globaltimer; block 0 / the last block): waiting for the peers to
    // arrive (rank skew), reducing + pushing, fence + waiting for the peers' done flags; number of calls
    unsigned long long t_arrive, t_reduce, t_depart, calls, t_fence, t_count;
};
static_assert(sizeof(Header) <= HEADER_BYTES, "header");
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

struct Args {
    char* buf[MAX_WORLD];  // base of every rank's peer buffer as mapped in this process (buf[rank] is local)
    int world, rank;
    uint32_t seq;
    unsigned long long n4;  // float4s of payload
    unsigned long long timeout_ns;  // 0 = wait for ever
};
unsigned long long g_timeout_ns = 600ull * 1000000000ull;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Payload accesses are weak, L2-only vector accesses (a warp's 32 x 16 B coalesce into full NVLink packets); the
// ordering comes from the flags: acquire before the loads, block barrier + system fence + release after the stores.
// (.relaxed.sys stores went out as single 16-byte packets: the push phase then ran at ~120 GB/s.)
__device__ __forceinline__ float4 ld_peer(const float4* p) { return __ldcg(p); }
__device__ __forceinline__ void st_peer(float4* p, float4 v) { __stcg(p, v); }
// Spins until the flag has reached seq (flags only grow; compared modulo 2^32).  A peer that does not show up within
// `timeout_ns` of wall-clock time (%globaltimer; default 10 minutes, csb_peer_set_timeout -- rank skew from a first-use
// build, a checkpoint or a re-render on one rank is legitimate and can be long) makes the wait give up: the caller
// records the call in the header and leaves the kernel without touching the payload.  No __trap(): the context, the
// peers' IPC mappings and the other streams survive, and the host raises from PeerBuffer.check_timeout().
__device__ __forceinline__ bool wait_flag(const uint32_t* p, uint32_t seq, unsigned long long timeout_ns) {
    const unsigned long long t0 = now_ns();
    uint32_t spins = 0;
    while ((int)(ld_acquire_sys(p) - seq) < 0) {
        if ((++spins & 1023u) == 0 && timeout_ns && now_ns() - t0 > timeout_ns) return false;
        __nanosleep(20);
    }
    return true;
}

template <int W>
__global__ void __launch_bounds__(PEER_THREADS) peer_all_reduce_kernel(const __grid_constant__ Args a) {
    const int tid = threadIdx.x;
    Header* mine = reinterpret_cast<Header*>(a.buf[a.rank]);
    const unsigned long long t0 = now_ns();

    // 0. arrive: everything this rank wrote into its bucket (earlier kernels of the stream) is ordered before the flag
    if (blockIdx.x == 0 && tid < W) {
        st_release_sys(&reinterpret_cast<Header*>(a.buf[tid])->arrive[a.rank], a.seq);
    }
    bool ok = true;
    if (tid < W) ok = wait_flag(&mine->arrive[tid], a.seq, a.timeout_ns);
    if (__syncthreads_or(!ok)) {  // a peer never arrived: every block of the grid finds out by itself and leaves
        if (tid == 0) mine->timed_out = a.seq;
        return;
    }
    const unsigned long long t1 = now_ns();

    // 1. reduce slice `rank`, push the sums to every bucket
    const unsigned long long slice = (a.n4 + W - 1) / W;
    const unsigned long long lo = slice * a.rank;
    unsigned long long hi = lo + slice;
    if (hi > a.n4) hi = a.n4;
    for (unsigned long long i = lo + (unsigned long long)blockIdx.x * blockDim.x + tid; i < hi;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        float4 v[W];
#pragma unroll
        for (int r = 0; r < W; ++r) v[r] = ld_peer(reinterpret_cast<const float4*>(a.buf[r] + HEADER_BYTES) + i);
        float4 s = v[0];
#pragma unroll
        for (int r = 1; r < W; ++r) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
#pragma unroll
        for (int r = 0; r < W; ++r) st_peer(reinterpret_cast<float4*>(a.buf[r] + HEADER_BYTES) + i, s);
    }

    // 2. depart.  One system fence per block (thread 0, after the block barrier: the fence is cumulative over the
    //    block's pushes it has synchronised with), then the grid counter.  Measured (t_fence): the fence waits ~8 us for
    //    the remote stores to be acknowledged -- the largest part of the kernel that is not rank skew.
    const unsigned long long t2 = now_ns();
    __syncthreads();
    __shared__ bool last;
    __shared__ unsigned long long t_f;
    if (tid == 0) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        t_f = now_ns();
        unsigned prev;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(&mine->blocks_done) : "memory");
        last = prev == gridDim.x - 1;
        if (last) mine->blocks_done = 0;  // every block has counted: ready for the next call
    }
    __syncthreads();
    if (!last) return;
    const unsigned long long t3 = now_ns();
    if (tid < W) {
        st_release_sys(&reinterpret_cast<Header*>(a.buf[tid])->done[a.rank], a.seq);
        if (!wait_flag(&mine->done[tid], a.seq, a.timeout_ns)) mine->timed_out = a.seq;
    }
    __syncwarp();
    if (tid == 0) {  // the last block's view of the call (its own phases; it is also the one that waits at the end)
        mine->t_arrive += t1 - t0;
        mine->t_reduce += t2 - t1;
        mine->t_depart += now_ns() - t2;
        mine->calls += 1;
        mine->t_fence += t_f - t2;   // block barrier + system fence of the last block
        mine->t_count += t3 - t_f;   // grid counter
    }
}

}  // namespace peer
}  // namespace csb

using namespace csb;
using namespace csb::peer;

extern "C" {

size_t csb_peer_header_bytes(void) { return HEADER_BYTES; }

int csb_peer_alloc(size_t bytes, void** out) {
    if (!out || bytes < HEADER_BYTES) return CSB_ERR_BAD_ARG;
    void* p = nullptr;
    CSB_CUDA_TRY(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(p);
        return (int)e;
    }
    *out = p;
    return 0;
}
int csb_peer_free(void* p) {
    if (!p) return 0;
    CSB_CUDA_TRY(cudaFree(p));
    return 0;
}
int csb_peer_export(void* p, void* handle64) {
    if (!p || !handle64) return CSB_ERR_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t h;
    CSB_CUDA_TRY(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, 64);
    return 0;
}
int csb_peer_import(const void* handle64, void** out) {
    if (!handle64 || !out) return CSB_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    CSB_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out = p;
    return 0;
}
int csb_peer_close(void* p) {
    if (!p) return 0;
    CSB_CUDA_TRY(cudaIpcCloseMemHandle(p));
    return 0;
}

int csb_peer_set_timeout(double seconds) {
    if (!(seconds >= 0.0)) return CSB_ERR_BAD_ARG;
    g_timeout_ns = (unsigned long long)(seconds * 1e9);  // 0 = never give up
    return 0;
}
int csb_peer_timed_out(const void* local_buf, unsigned* seq_out) {
    if (!local_buf || !seq_out) return CSB_ERR_BAD_ARG;
    uint32_t v = 0;
    CSB_CUDA_TRY(cudaMemcpy(&v, (const char*)local_buf + offsetof(Header, timed_out), 4, cudaMemcpyDeviceToHost));
    *seq_out = v;
    return 0;
}

int csb_peer_all_reduce(int world, int rank, void* const* bufs, size_t n_floats, unsigned seq, void* stream_) {
    if (world < 1 || world > MAX_WORLD || rank < 0 || rank >= world || !bufs) return CSB_ERR_BAD_ARG;
    if (n_floats % 4 != 0) return CSB_ERR_BAD_ARG;
    if (world == 1 || n_floats == 0) return 0;
    Args a{};
    for (int r = 0; r < world; ++r) {
        if (!bufs[r]) return CSB_ERR_BAD_ARG;
        a.buf[r] = (char*)bufs[r];
    }
    a.world = world;
    a.rank = rank;
    a.seq = seq;
    a.n4 = n_floats / 4;
    a.timeout_ns = g_timeout_ns;
    const unsigned long long slice = (a.n4 + world - 1) / world;
    unsigned long long blocks = (slice + PEER_THREADS - 1) / PEER_THREADS;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 4) blocks = 148 * 4;
    cudaStream_t stream = (cudaStream_t)stream_;
    switch (world) {
#define CSB_PEER_CASE(W)                                                                                        \
    case W:                                                                                                     \
        CSB_LAUNCH("peer_all_reduce", stream, peer_all_reduce_kernel<W><<<(unsigned)blocks, PEER_THREADS, 0, stream>>>(a)); \
        break;
        CSB_PEER_CASE(2)
        CSB_PEER_CASE(3)
        CSB_PEER_CASE(4)
        CSB_PEER_CASE(5)
        CSB_PEER_CASE(6)
        CSB_PEER_CASE(7)
        CSB_PEER_CASE(8)
#undef CSB_PEER_CASE
        default:
            return CSB_ERR_UNSUPPORTED;
    }
    return 0;
}

}  // extern "C"
