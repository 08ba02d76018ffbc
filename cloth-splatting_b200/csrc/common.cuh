// Shared helpers for the sm_100a kernels of the cloth-splatting hot path.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <stdint.h>
#include <stddef.h>
#include "../../include/cloth_splatting_b200.h"

#define CSB_TILE_X 16  // RAST/cuda_rasterizer/config.h:16-17 (results depend on it through getRect)
#define CSB_TILE_Y 16
#define CSB_TILE_PIX (CSB_TILE_X * CSB_TILE_Y)

#define CSB_CUDA_TRY(expr)                         \
    do {                                           \
        cudaError_t _e = (expr);                   \
        if (_e != cudaSuccess) return (int)_e;     \
    } while (0)

#define CSB_LAUNCH_CHECK()                         \
    do {                                           \
        cudaError_t _e = cudaGetLastError();       \
        if (_e != cudaSuccess) return (int)_e;     \
    } while (0)

namespace csb {

// Counts every kernel launch of the library and, while csb_profile_enable(1) is in effect, brackets
// it with CUDA events on the launching stream (api_misc.cu).
struct LaunchScope {
    const char* name;
    cudaStream_t stream;
    int slot;
    LaunchScope(const char* name, cudaStream_t stream);
    ~LaunchScope();
};
#define CSB_LAUNCH(name, stream, ...)              \
    do {                                           \
        {                                          \
            csb::LaunchScope _ls(name, stream);    \
            __VA_ARGS__;                           \
        }                                          \
        CSB_LAUNCH_CHECK();                        \
    } while (0)

// Programmatic dependent launch: the kernel is queued behind its predecessor in the stream and may be scheduled
// before that one has drained; every kernel launched this way starts with grid_dep_wait(), which returns once
// the predecessor grid has completed and its writes are visible (same semantics as stream order, minus the launch
// gap between dependent small kernels).  grid_dep_launch() lets the NEXT kernel of the chain be scheduled early.
#ifdef __CUDACC__
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() {
#ifdef CSB_PDL_EARLY_TRIGGER
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
__device__ __forceinline__ void grid_dep_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: opt in once per (kernel, device), not once per
// process (a process that drives a second GPU would otherwise launch > 48 KB kernels there without the opt-in).
// `done_mask` is a caller-owned static bit mask over device ordinals (devices >= 64 set the attribute every time).
template <typename K>
static inline cudaError_t opt_in_dynamic_smem(K kernel, int bytes, unsigned long long* done_mask) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const unsigned long long bit = dev < 64 ? (1ull << dev) : 0ull;
    if (bit && (__atomic_load_n(done_mask, __ATOMIC_ACQUIRE) & bit)) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && bit) __atomic_fetch_or(done_mask, bit, __ATOMIC_RELEASE);
    return e;
}

// Bump allocator over a caller-provided workspace (256-byte aligned slices).
struct Carver {
    char* base;
    size_t off;
    explicit Carver(void* p) : base((char*)p), off(0) {}
    template <typename T>
    T* take(size_t count) {
        off = align_up(off, 256);
        T* r = (T*)(base ? base + off : nullptr);
        off += count * sizeof(T);
        return r;
    }
    size_t total() const { return align_up(off, 256); }
};

// Per-launch constants of up to CSB_MAX_VIEWS views, passed by value as a kernel parameter.
struct ViewBatch {
    int V, P, W, H, gx, gy, tiles;
    int cull_exact;  // compositing kernels: exact ellipse-vs-quad visit culling (render_lanes.cuh quad_mask); csb_debug_cull_exact
    float focal_x[CSB_MAX_VIEWS], focal_y[CSB_MAX_VIEWS];
    float tan_fovx[CSB_MAX_VIEWS], tan_fovy[CSB_MAX_VIEWS];
    const float* viewmatrix[CSB_MAX_VIEWS];
    const float* projmatrix[CSB_MAX_VIEWS];
    const float* campos[CSB_MAX_VIEWS];
    const float* background[CSB_MAX_VIEWS];
};

// Number of bits needed to tell n tile ids apart the way the reference does
// (rasterizer_impl.cu:35-50: the bit above the MSB of n).
static inline int higher_msb(uint32_t n) {
    int b = 0;
    while (b < 32 && (n >> b)) ++b;
    return b < 1 ? 1 : b;
}

// ---- workspaces ---------------------------------------------------------------
struct GeomState {
    float2* means2D;         // [V*P]
    float4* conic_opacity;   // [V*P]
    float4* rgb_depth;       // [V*P]  r,g,b, view-space z
    float* cov3D;            // [P*6]  (view independent)
    uint8_t* clamped;        // [V*P]  bit c set = channel c clamped at 0
    uint32_t* tiles_touched; // [V*P]
    uint4* rects;            // [V*P]  (min.x, min.y, width, depth bits) of the touched tile rectangle
    uint32_t* depth_keys;    // [V*P]  sort key of the per-view depth order
    uint32_t* order_a;       // [V*P]  ping/pong buffers of the depth sort (values = v*P + idx)
    uint32_t* order_b;       // [V*P]
    uint32_t* keys_b;        // [V*P]
    float* grad_acc;         // [V*P*12] backward accumulators: dmean2D.xy, dconic.xyz, dopacity, dcolor.rgb
    uint32_t* control;       // counters, scan states, histograms (zeroed by one memset per forward)
    size_t control_words;
    size_t bytes;
    static GeomState carve(void* ws, int V, int P);
};

constexpr int TILE_ORDER_HEAD = 513;
struct ImageState {
    float* final_T;      // [V*H*W]
    uint32_t* n_contrib; // [V*H*W]
    uint2* ranges;       // [V*tiles]
    uint32_t* tile_order; // [TILE_ORDER_HEAD + V*tiles]: word 0 = 1 when the ids behind the head list the (view, tile) pairs
                          // longest list first; words 1..256 class histogram, 257..512 class cursors (build_tile_order)
    size_t bytes;
    static ImageState carve(void* ws, int V, int W, int H);
};

struct BinState {
    uint32_t* tiles_a; // [cap] tile id (v*tiles + tile) of each instance, ping
    uint32_t* tiles_b; // [cap] pong
    uint32_t* vals_a;  // [cap] Gaussian idx
    uint32_t* vals_b;  // [cap]
    uint64_t* keys;    // [cap] optional u64 tile|depth keys (flags bit0)
    uint32_t* lookback; // onesweep status words
    size_t lookback_words;
    size_t bytes;
    static BinState carve(void* ws, int V, int P, int W, int H, size_t cap);
};

// control-region layout (u32 words)
enum {
    CTRL_NUM_RENDERED = 0,  // instances to sort/render (0 when the binning workspace overflowed)
    CTRL_OVERFLOW = 1,      // set when the instance count exceeds the capacity
    CTRL_TICKET_SCAN = 2,   // block ticket of the emit kernel
    CTRL_N_ENTRIES = 3,     // V*P, element count of the depth sort
    CTRL_TICKET_DEPTH = 4,  // 4 tickets, one per radix pass of the depth sort
    CTRL_TICKET_TILE = 8,   // 4 tickets of the tile sort
    CTRL_HIST_DEPTH = 16,   // 4 passes x 256 bins
    CTRL_HIST_TILE = CTRL_HIST_DEPTH + 4 * 256,
    CTRL_SCAN_STATE = CTRL_HIST_TILE + 4 * 256,  // u64 per emit block, then the depth sort's look-back words
    CTRL_FIXED_WORDS = CTRL_SCAN_STATE,
};

// ---- radix sort (radix_sort.cu) -------------------------------------------------
// Stable LSD onesweep sort of (u32 key, u32 value) pairs on key bits [0, end_bit).
// n is read from the device (n_dev), n_max bounds the grid.  Buffers ping-pong; the
// *side_out tells which side holds the result (0 = *_a, 1 = *_b); returns 0 or an error code.
#define CSB_SORT_TILE 4096
#define CSB_SORT_TILE_SMALL 1024
// Pairs per thread block of a pass: inputs that would not fill the 148 SMs with 4096-pair tiles (the per-view depth
// sort and the tile sort of the cloth scenes) use 1024-pair tiles -- 4x the blocks, a quarter of the serial work each.
// (Digit widths: sharing the bits evenly between the passes -- 13 tile bits as 7 + 6 instead of 8 + 5, i.e. 128 runs of 32
// pairs per 4096-pair tile instead of 256 runs of 16 -- was measured SLOWER on B200: C3 tile passes 480 -> 512 us.)
static inline size_t sort_tile_size(size_t n_max) {
    return n_max < (size_t)148 * CSB_SORT_TILE ? CSB_SORT_TILE_SMALL : CSB_SORT_TILE;
}
static inline size_t sort_lookback_words(size_t n_max, int passes) {
    // non-decreasing in n_max (workspace sizes are inverted by bisection, csb_rasterize_binning_capacity)
    const size_t switch_at = (size_t)148 * CSB_SORT_TILE;
    const size_t small = (std::min(n_max, switch_at) + CSB_SORT_TILE_SMALL - 1) / CSB_SORT_TILE_SMALL + 1;
    const size_t large = (n_max + CSB_SORT_TILE - 1) / CSB_SORT_TILE + 1;
    return (size_t)passes * std::max(small, large) * 256;
}
// Which buffer side (0 = *_a, 1 = *_b) holds the tile-sorted instance list.
static inline int sorted_side(size_t total_tiles) {
    const int bits = higher_msb((uint32_t)total_tiles);
    return ((bits + 7) / 8) & 1;
}
int fill_view_batch(struct ViewBatch& vb, int num_views, const csb_view* views, int P, int width, int height);
// hist_ready: the caller's producer kernel already accumulated the digit histograms of all passes into `hist`
// (warp_hist_add below), so the separate histogram launch is skipped.
int radix_sort_pairs(uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                     const uint32_t* n_dev, size_t n_max, int end_bit, uint32_t* hist /*4x256, zeroed*/,
                     uint32_t* tickets /*4, zeroed*/, uint32_t* lookback /*zeroed*/, int* side_out,
                     cudaStream_t stream, bool hist_ready = false, const char* launch_name = nullptr);

#ifdef __CUDACC__
// One key per lane -> shared-memory digit histograms of `passes` 8-bit digits (sh[p * 256 + d]).  Neighbouring keys
// mostly share their upper digits (depth exponents, tile rows): the lanes that agree with lane 0 are counted with one
// ballot, only the others fall back to their own atomic.  All 32 lanes must call it; `ok` masks lanes without a key.
__device__ __forceinline__ void warp_hist_add(uint32_t* sh, int passes, uint32_t key, bool ok, int lane) {
    for (int p = 0; p < passes; ++p) {
        const uint32_t dd = ok ? ((key >> (8 * p)) & 255u) : 0x1FFu;
        const uint32_t d0 = __shfl_sync(0xFFFFFFFFu, dd, 0);
        const uint32_t same = __ballot_sync(0xFFFFFFFFu, dd == d0);
        if (lane == 0 && d0 < 256u) atomicAdd(&sh[p * 256 + d0], (uint32_t)__popc(same));
        if (dd != d0 && dd < 256u) atomicAdd(&sh[p * 256 + dd], 1u);
    }
}
#endif

}  // namespace csb
