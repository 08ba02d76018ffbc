// Small C-ABI entry points: version, error strings, workspace introspection for parity tests.
#include <string.h>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "common.cuh"

using namespace csb;

// ---- launch accounting (always) and per-kernel CUDA-event timing (opt-in) ------------------
namespace {
std::atomic<unsigned long long> g_launches{0};
std::atomic<int> g_profile{0};
struct Span {
    const char* name;
    cudaEvent_t a, b;
};
std::mutex g_mu;
std::vector<Span> g_spans;
std::vector<cudaEvent_t> g_free_events;

cudaEvent_t take_event() {
    if (!g_free_events.empty()) {
        cudaEvent_t e = g_free_events.back();
        g_free_events.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
}  // namespace

namespace csb {
int g_render_variant = 0x4F;  // forward: packed kernels chosen by path (15); backward: one pixel pair per lane, delayed reduction (measured best on B200)
int g_cull_exact = 1;
LaunchScope::LaunchScope(const char* name_, cudaStream_t stream_) : name(name_), stream(stream_), slot(-1) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (g_profile.load(std::memory_order_relaxed)) {
        std::lock_guard<std::mutex> lk(g_mu);
        Span sp{name, take_event(), take_event()};
        cudaEventRecord(sp.a, stream);
        slot = (int)g_spans.size();
        g_spans.push_back(sp);
    }
}
LaunchScope::~LaunchScope() {
    if (slot >= 0) {
        std::lock_guard<std::mutex> lk(g_mu);
        if (slot < (int)g_spans.size()) cudaEventRecord(g_spans[slot].b, stream);
    }
}
}  // namespace csb

extern "C" {

int csb_abi_version(void) { return 1; }

const char* csb_error_string(int code) {
    switch (code) {
        case CSB_OK: return "ok";
        case CSB_ERR_BAD_ARG: return "bad argument";
        case CSB_ERR_WORKSPACE: return "workspace too small";
        case CSB_ERR_UNSUPPORTED: return "unsupported configuration";
        case CSB_ERR_STATIC_LIMIT: return "problem exceeds the fixed scratch array of this no-allocation entry point (use the workspace variant)";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
    }
}

unsigned long long csb_launch_count(void) { return g_launches.load(); }

int csb_debug_render_variant(int variant) {
    const int old = csb::g_render_variant;
    if (variant >= 0 && ((variant & 15) <= 4 || (variant & 15) == 15) && (variant >> 4) <= 7) csb::g_render_variant = variant;
    return old;
}

int csb_debug_cull_exact(int on) {
    const int old = csb::g_cull_exact;
    if (on == 0 || on == 1) csb::g_cull_exact = on;
    return old;
}

void csb_profile_enable(int on) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto& sp : g_spans) {
        g_free_events.push_back(sp.a);
        g_free_events.push_back(sp.b);
    }
    g_spans.clear();
    g_profile.store(on ? 1 : 0);
}

int csb_profile_read(const char* name, double* total_ms, unsigned long long* launches) {
    std::lock_guard<std::mutex> lk(g_mu);
    double ms = 0.0;
    unsigned long long n = 0;
    for (auto& sp : g_spans) {
        if (name && strcmp(name, sp.name) != 0) continue;
        cudaError_t e = cudaEventSynchronize(sp.b);
        if (e != cudaSuccess) return (int)e;
        float t = 0.f;
        e = cudaEventElapsedTime(&t, sp.a, sp.b);
        if (e != cudaSuccess) return (int)e;
        ms += t;
        ++n;
    }
    if (total_ms) *total_ms = ms;
    if (launches) *launches = n;
    return 0;
}

int csb_profile_names(char* out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mu);
    std::map<std::string, int> seen;
    std::string all;
    for (auto& sp : g_spans)
        if (!seen[sp.name]++) { all += sp.name; all += ","; }
    if (!out || out_bytes == 0) return (int)all.size() + 1;
    strncpy(out, all.c_str(), out_bytes - 1);
    out[out_bytes - 1] = 0;
    return (int)all.size() + 1;
}

size_t csb_rasterize_binning_capacity(int num_views, int P, int width, int height, size_t bytes) {
    size_t lo = 0, hi = bytes / 24 + 1;  // every instance needs at least 24 bytes
    while (lo < hi) {
        const size_t mid = lo + (hi - lo + 1) / 2;
        if (BinState::carve(nullptr, num_views, P, width, height, mid).bytes <= bytes) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

void* csb_rasterize_state_ptr(const char* name, int num_views, int P, int width, int height, size_t binning_capacity,
                              void* geom_ws, void* image_ws, void* binning_ws) {
    if (!name) return nullptr;
    GeomState g = GeomState::carve(geom_ws, num_views, P);
    ImageState im = ImageState::carve(image_ws, num_views, width, height);
    BinState b = BinState::carve(binning_ws, num_views, P, width, height, binning_capacity);
    const size_t tiles = (size_t)num_views * ((width + CSB_TILE_X - 1) / CSB_TILE_X) *
                         ((height + CSB_TILE_Y - 1) / CSB_TILE_Y);
    const int side = sorted_side(tiles);
    if (!strcmp(name, "means2D")) return g.means2D;
    if (!strcmp(name, "conic_opacity")) return g.conic_opacity;
    if (!strcmp(name, "rgb_depth")) return g.rgb_depth;
    if (!strcmp(name, "cov3D")) return g.cov3D;
    if (!strcmp(name, "clamped")) return g.clamped;
    if (!strcmp(name, "tiles_touched")) return g.tiles_touched;
    if (!strcmp(name, "grad_acc")) return g.grad_acc;
    if (!strcmp(name, "control")) return g.control;
    if (!strcmp(name, "final_T")) return im.final_T;
    if (!strcmp(name, "n_contrib")) return im.n_contrib;
    if (!strcmp(name, "ranges")) return im.ranges;
    if (!strcmp(name, "point_list")) return side ? b.vals_b : b.vals_a;
    if (!strcmp(name, "point_tiles")) return side ? b.tiles_b : b.tiles_a;
    if (!strcmp(name, "keys")) return b.keys;
    return nullptr;
}

}  // extern "C"
