// Small C-ABI entry points: version, error strings, workspace introspection for parity tests.
#include <string.h>
#include "common.cuh"

using namespace csb;

extern "C" {

int csb_abi_version(void) { return 1; }

const char* csb_error_string(int code) {
    switch (code) {
        case CSB_OK: return "ok";
        case CSB_ERR_BAD_ARG: return "bad argument";
        case CSB_ERR_WORKSPACE: return "workspace too small";
        case CSB_ERR_UNSUPPORTED: return "unsupported configuration";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
    }
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
