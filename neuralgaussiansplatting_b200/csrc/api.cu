// extern "C" entry points of libb200splat.so (see include/b200splat.h).
// Host orchestration only: carves the caller's workspaces and enqueues the kernels on the
// caller's stream. Replaces CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (reference rasterizer_impl.cu:141-153,198-434).

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "kernels.cuh"
#if defined(B200S_NVTX) && B200S_NVTX
#include <nvtx3/nvToolsExt.h>
#endif

using namespace b200s;

namespace b200s {
int idxmap_forward_impl(const B200sIdxmapArgs* a, cudaStream_t stream);
int idxmap_backward_impl(int P, int W, int H, const int32_t* out_idx, const float* grad_feature,
                         float* dL_dfeature_vector, cudaStream_t stream);
size_t idxmap_workspace_size_impl(int P, int W, int H);
int idxmap_forward_static_impl(const B200sIdxmapArgs* a, cudaStream_t stream);
}

// ------------------------------------------------------------------ measurement hooks
namespace b200s {
struct Bracket { int stage; cudaEvent_t e0, e1; };
static std::atomic<bool> g_prof_on{false};
static std::atomic<uint64_t> g_launches{0};
static std::mutex g_prof_mu;
static std::vector<Bracket> g_brackets;
static std::vector<cudaEvent_t> g_open[B200S_NUM_STAGES];
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
void prof_begin(int stage, cudaStream_t stream) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, stream);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_open[stage].push_back(e);
}
void prof_end(int stage, cudaStream_t stream) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (g_open[stage].empty()) return;
    cudaEvent_t e0 = g_open[stage].back();
    g_open[stage].pop_back();
    cudaEvent_t e1;
    if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return; }
    cudaEventRecord(e1, stream);
    g_brackets.push_back({stage, e0, e1});
}
#if defined(B200S_NVTX) && B200S_NVTX
static const char* const kStageNames[B200S_NUM_STAGES] = {
    "b200s:preprocess", "b200s:scan", "b200s:expand", "b200s:sort", "b200s:ranges", "b200s:blend_fwd",
    "b200s:blend_bwd", "b200s:preprocess_bwd", "b200s:idxmap", "b200s:idxmap_bwd", "b200s:misc"};
void nvtx_push(int stage) { nvtxRangePushA(stage >= 0 && stage < B200S_NUM_STAGES ? kStageNames[stage] : "b200s:?"); }
void nvtx_pop() { nvtxRangePop(); }
#endif
}  // namespace b200s

extern "C" int b200s_profile_begin(void) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& b : g_brackets) { cudaEventDestroy(b.e0); cudaEventDestroy(b.e1); }
    g_brackets.clear();
    g_prof_on.store(true);
    return 0;
}
extern "C" int b200s_profile_end(float* ms, int32_t* calls) {
    g_prof_on.store(false);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (int i = 0; i < B200S_NUM_STAGES; ++i) { if (ms) ms[i] = 0.f; if (calls) calls[i] = 0; }
    int rc = 0;
    for (auto& b : g_brackets) {
        float t = 0.f;
        cudaError_t e = cudaEventSynchronize(b.e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, b.e0, b.e1);
        if (e != cudaSuccess) rc = (int)e;
        else { if (ms) ms[b.stage] += t; if (calls) calls[b.stage] += 1; }
        cudaEventDestroy(b.e0); cudaEventDestroy(b.e1);
    }
    g_brackets.clear();
    for (auto& v : g_open) { for (auto e : v) cudaEventDestroy(e); v.clear(); }
    return rc;
}
extern "C" uint64_t b200s_launch_count(void) { return g_launches.load(); }

namespace {

// Warps per tile for the blend kernels; overridable for tuning (B200S_BLEND_WARPS=1|2|4|8).
int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return atoi(v);
}
// Default: 2 warps per tile (fewest instructions per pair: a lane sums 4 pixels before the warp
// reduction) unless the image has too few tiles to fill the 148 SMs with that many warps.
int auto_warps(int tiles) {
    int nw = 2;
    while (nw < 8 && (int64_t)tiles * nw < 148 * 48) nw *= 2;
    return nw;
}
int blend_fwd_warps(int tiles) { static int v = env_int("B200S_BLEND_FWD_WARPS", 0); return v > 0 ? v : auto_warps(tiles); }
int blend_bwd_warps(int tiles) { static int v = env_int("B200S_BLEND_BWD_WARPS", 0); return v > 0 ? v : auto_warps(tiles); }

struct Dims { int gx, gy, tiles; float focal_x, focal_y; };
Dims dims_of(int W, int H, float tan_fovx, float tan_fovy) {
    Dims d;
    d.gx = (W + TILE - 1) / TILE;
    d.gy = (H + TILE - 1) / TILE;
    d.tiles = d.gx * d.gy;
    // rasterizer_impl.cu:222-223 (host float arithmetic)
    d.focal_y = H / (2.0f * tan_fovy);
    d.focal_x = W / (2.0f * tan_fovx);
    return d;
}

int check_forward_args(const B200sForwardArgs* a) {
    if (!a) return B200S_E_BADARG;
    if (a->P < 0 || a->W <= 0 || a->H <= 0) return B200S_E_BADARG;
    if (a->P == 0) return 0;
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->campos || !a->background)
        return B200S_E_BADARG;
    if (!a->out_color || !a->radii) return B200S_E_BADARG;
    if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) return B200S_E_BADARG;
    const bool has_sr = a->scales != nullptr && a->rotations != nullptr;
    if (has_sr == (a->cov3D_precomp != nullptr)) return B200S_E_BADARG;
    if (a->shs && a->M <= 0) return B200S_E_BADARG;
    if (a->W > 65535 * TILE || a->H > 65535 * TILE) return B200S_E_UNSUPPORTED;
    if (a->blend_mode != 0 && a->blend_mode != 1) return B200S_E_BADARG;
    if (a->P >= (1 << 30)) return B200S_E_UNSUPPORTED;   // packed look-back word of scan_compact
    if (!a->geom_ws || !a->img_ws) return B200S_E_BADARG;
    if (a->geom_bytes < b200s_geometry_workspace_size(a->P)) return B200S_E_WORKSPACE;
    if (a->img_bytes < b200s_image_workspace_size(a->W, a->H)) return B200S_E_WORKSPACE;
    if (a->rotations && (reinterpret_cast<uintptr_t>(a->rotations) & 15)) return B200S_E_BADARG;
    if ((reinterpret_cast<uintptr_t>(a->geom_ws) & 255) || (reinterpret_cast<uintptr_t>(a->img_ws) & 255))
        return B200S_E_BADARG;
    return 0;
}

}  // namespace

extern "C" {

int b200s_abi_version(void) { return B200S_ABI_VERSION; }

const char* b200s_build_info(void) {
    return "b200splat sm_100a nvcc " __DATE__ " " __TIME__;
}

size_t b200s_geometry_workspace_size(int32_t P) {
    return carve_geom(nullptr, P > 0 ? P : 1).bytes;
}
size_t b200s_image_workspace_size(int32_t W, int32_t H) {
    return carve_img(nullptr, W, H).bytes;
}
size_t b200s_binning_workspace_size(int64_t capacity, int32_t W, int32_t H) {
    (void)W; (void)H;
    return carve_bin(nullptr, capacity).bytes;
}
size_t b200s_backward_workspace_size(int32_t P, int64_t num_rendered) {
    return carve_bwd(nullptr, P, num_rendered).bytes;
}
int64_t b200s_binning_capacity_for_bytes(size_t bytes) {
    // largest capacity whose workspace fits in `bytes` (sizes are monotone in capacity, and equal
    // sizes imply identical internal offsets, so this inverts b200s_binning_workspace_size)
    int64_t lo = 0, hi = (1ll << 30) - 1;
    if (carve_bin(nullptr, 1).bytes > bytes) return -1;
    while (lo < hi) {
        const int64_t mid = lo + (hi - lo + 1) / 2;
        if (carve_bin(nullptr, mid).bytes <= bytes) lo = mid; else hi = mid - 1;
    }
    return lo;
}
size_t b200s_sizeof(int32_t which) {
    switch (which) {
        case 0: return sizeof(B200sHeader);
        case 1: return sizeof(B200sForwardArgs);
        case 2: return sizeof(B200sBackwardArgs);
        case 3: return sizeof(B200sIdxmapArgs);
        case 4: return sizeof(B200sPeerReduceArgs);
        case 5: return sizeof(B200sAdamArgs);
        case 6: return sizeof(B200sLossArgs);
        case 7: return sizeof(B200sShGradArgs);
        default: return 0;
    }
}

int b200s_forward_preprocess(const B200sForwardArgs* a, b200s_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_forward_args(a);
    if (rc) return rc;
    const Dims d = dims_of(a->W, a->H, a->tan_fovx, a->tan_fovy);
    GeomWS g = carve_geom(a->geom_ws, a->P > 0 ? a->P : 1);
    // header + counters + scan status are small; clear them every call
    B200S_CUDA_OK(cudaMemsetAsync(g.header, 0, sizeof(B200sHeader), stream));
    B200S_CUDA_OK(cudaMemsetAsync(g.counters, 0, 64 * sizeof(uint32_t), stream));
    if (a->P == 0) return 0;
    B200S_CUDA_OK(cudaMemsetAsync((char*)a->geom_ws + g.zero_off, 0, g.zero_bytes, stream));
    PreprocessParams p;
    p.P = a->P; p.D = a->D; p.M = a->M; p.W = a->W; p.H = a->H;
    p.grid_x = d.gx; p.grid_y = d.gy;
    p.tan_fovx = a->tan_fovx; p.tan_fovy = a->tan_fovy; p.focal_x = d.focal_x; p.focal_y = d.focal_y;
    p.scale_modifier = a->scale_modifier; p.prefiltered = a->prefiltered;
    p.means3D = a->means3D; p.shs = a->shs; p.colors_precomp = a->colors_precomp; p.opacities = a->opacities;
    p.scales = a->scales; p.rotations = a->rotations; p.cov3D_precomp = a->cov3D_precomp;
    p.viewmatrix = a->viewmatrix; p.projmatrix = a->projmatrix; p.campos = a->campos;
    p.radii = a->radii; p.rec = g.rec; p.tiles_touched = g.tiles_touched; p.depth_bits = g.depth_bits;
    p.clamped = g.clamped;
    p.header = g.header; p.counters = g.counters;
    {
        ProfScope ps(B200S_STAGE_PREPROCESS, stream);
        rc = launch_preprocess_fwd(p, stream);
    }
    if (rc) return rc;
    B200S_AFTER_LAUNCH(stream, a->debug);
    return launch_scan_and_depth_sort(a->P, g, a->debug != 0, stream);
}

int b200s_forward_render(const B200sForwardArgs* a, b200s_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_forward_args(a);
    if (rc) return rc;
    const Dims d = dims_of(a->W, a->H, a->tan_fovx, a->tan_fovy);
    const size_t plane = (size_t)a->W * a->H;
    if (a->P == 0) {
        // RasterizeGaussiansCUDA returns the zero-initialised image when P == 0
        // (rasterize_points.cu:68,80): keep that.
        B200S_CUDA_OK(cudaMemsetAsync(a->out_color, 0, 3 * plane * sizeof(float), stream));
        return 0;
    }
    if (!a->bin_ws || a->bin_capacity < 0) return B200S_E_BADARG;
    if (reinterpret_cast<uintptr_t>(a->bin_ws) & 255) return B200S_E_BADARG;
    if (a->bin_bytes < b200s_binning_workspace_size(a->bin_capacity, a->W, a->H)) return B200S_E_WORKSPACE;
    if (a->bin_capacity >= (1ll << 30)) return B200S_E_UNSUPPORTED;
    GeomWS g = carve_geom(a->geom_ws, a->P);
    ImgWS im = carve_img(a->img_ws, a->W, a->H);
    BinWS bw = carve_bin(a->bin_ws, a->bin_capacity);
    int final_buf = 0;
    // stage 2 may be re-run with a larger workspace after an overflow: reset its device state
    B200S_CUDA_OK(cudaMemsetAsync(&g.header->overflow, 0, sizeof(uint32_t), stream));
    B200S_CUDA_OK(cudaMemsetAsync(g.counters + CTR_SORT_TICKET0, 0, SORT_MAX_PASSES * sizeof(uint32_t), stream));
    B200S_CUDA_OK(cudaMemsetAsync(g.counters + CTR_EXPAND_TICKET, 0, sizeof(uint32_t), stream));
    B200S_CUDA_OK(cudaMemsetAsync(g.exp_status, 0, 4 * ((size_t)(a->P + EXP_TILE - 1) / EXP_TILE + 1), stream));
    rc = launch_binning(a->P, g, bw, a->bin_capacity, a->bin_capacity, d.gx, d.tiles, im.ranges,
                        &final_buf, a->debug != 0, stream);
    if (rc) return rc;
    BlendFwdParams b;
    b.W = a->W; b.H = a->H; b.grid_x = d.gx; b.grid_y = d.gy;
    b.ranges = im.ranges; b.point_list = bw.vals[final_buf]; b.rec = g.rec;
    b.background = a->background;
    b.final_T = im.final_T; b.n_contrib = im.n_contrib; b.tile_last = im.tile_last; b.out_color = a->out_color;
    b.fast = a->blend_mode == 1;
    b.P = a->P; b.list_cap = a->bin_capacity;
    b.tile_order = nullptr;
    // The forward's work per tile is not predicted by its list length (pixels terminate early): measured gain 7 us
    // for a 9 us ordering kernel at C3, none at C4. Opt-in (B200S_TILE_ORDER=2); the backward's order is exact.
    if (tile_order_enabled() && tile_order_fwd_enabled()) {
        ProfScope ps(B200S_STAGE_RANGES, stream);
        rc = launch_tile_order(im.ranges, nullptr, d.tiles, im.order_fwd, stream);
        if (rc) return rc;
        b.tile_order = im.order_fwd;
    }
    {
        ProfScope ps(B200S_STAGE_BLEND_FWD, stream);
        rc = launch_blend_fwd(b, blend_fwd_warps(d.tiles), stream);
    }
    if (rc) return rc;
    B200S_AFTER_LAUNCH(stream, a->debug);
    return 0;
}

int b200s_forward(const B200sForwardArgs* a, b200s_stream_t stream) {
    int rc = b200s_forward_preprocess(a, stream);
    if (rc) return rc;
    return b200s_forward_render(a, stream);
}

int b200s_read_header(const void* geom_ws, B200sHeader* header_host, b200s_stream_t stream) {
    if (!geom_ws || !header_host) return B200S_E_BADARG;
    B200S_CUDA_OK(cudaMemcpyAsync(header_host, geom_ws, sizeof(B200sHeader), cudaMemcpyDeviceToHost,
                                  (cudaStream_t)stream));
    return 0;
}

int b200s_forward_batch(const B200sForwardArgs* views, int32_t count, B200sHeader* headers_host,
                        b200s_stream_t stream) {
    if (count < 0 || (count > 0 && (!views || !headers_host))) return B200S_E_BADARG;
    for (int32_t i = 0; i < count; ++i) {
        int rc = b200s_forward(&views[i], stream);
        if (rc) return rc;
        if (views[i].P > 0) {
            rc = b200s_read_header(views[i].geom_ws, &headers_host[i], stream);
            if (rc) return rc;
        } else {
            memset(&headers_host[i], 0, sizeof(B200sHeader));
        }
    }
    return 0;
}

int b200s_backward(const B200sBackwardArgs* a, b200s_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!a) return B200S_E_BADARG;
    if (a->P < 0 || a->W <= 0 || a->H <= 0) return B200S_E_BADARG;
    if (a->P == 0) return 0;
    if (!a->means3D || !a->viewmatrix || !a->projmatrix || !a->campos || !a->background || !a->radii ||
        !a->dL_dout_color || !a->geom_ws || !a->img_ws || !a->bwd_ws)
        return B200S_E_BADARG;
    if (!a->dL_dmeans3D || !a->dL_dmeans2D || !a->dL_dopacity) return B200S_E_BADARG;
    if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) return B200S_E_BADARG;
    const bool has_sr = a->scales != nullptr && a->rotations != nullptr;
    if (has_sr == (a->cov3D_precomp != nullptr)) return B200S_E_BADARG;
    if (a->num_rendered < 0 || a->num_rendered > a->bin_capacity) return B200S_E_BADARG;
    if (a->num_rendered > 0 && !a->bin_ws) return B200S_E_BADARG;
    if (a->blend_mode != 0 && a->blend_mode != 1) return B200S_E_BADARG;   // same domain as the forward's check
    if (a->bwd_bytes < b200s_backward_workspace_size(a->P, a->deterministic ? a->num_rendered : 0))
        return B200S_E_WORKSPACE;
    if (a->dL_drotations && (reinterpret_cast<uintptr_t>(a->dL_drotations) & 15)) return B200S_E_BADARG;
    if (a->dL_dconic && (reinterpret_cast<uintptr_t>(a->dL_dconic) & 15)) return B200S_E_BADARG;
    if (reinterpret_cast<uintptr_t>(a->bwd_ws) & 255) return B200S_E_BADARG;

    const Dims d = dims_of(a->W, a->H, a->tan_fovx, a->tan_fovy);
    GeomWS g = carve_geom(const_cast<void*>(a->geom_ws), a->P);
    ImgWS im = carve_img(const_cast<void*>(a->img_ws), a->W, a->H);
    BinWS bw = carve_bin(const_cast<void*>(a->bin_ws), a->bin_capacity);
    BwdWS ws = carve_bwd(a->bwd_ws, a->P, a->deterministic ? a->num_rendered : 0);
    const int passes = sort_num_passes(d.tiles);
    const int final_buf = passes & 1;

    // zero the raw-sum scratch; the same launch orders the tiles for the backward blend (longest list first)
    const bool ordered = tile_order_enabled() && a->num_rendered > 0;
    {
        ProfScope misc(B200S_STAGE_MISC, stream);
        void* zp = a->deterministic ? (void*)ws.partial : (void*)ws.accum;
        const size_t zb = sizeof(float) * ACC_STRIDE *
                          (a->deterministic ? (size_t)(a->num_rendered > 0 ? a->num_rendered : 1) : (size_t)a->P);
        // (Clearing only the rows of visible Gaussians — 48-byte rows at 43 % density — was measured SLOWER than the
        // dense fill: 0.056 vs 0.039 ms of `misc` at C3, partial-sector stores. B200S_SPARSE_ZERO=1 keeps it testable.)
        static const bool sparse_zero = [] { const char* v = getenv("B200S_SPARSE_ZERO"); return v && *v && atoi(v) != 0; }();
        int rc0 = launch_bwd_prologue(zp, zb, (a->deterministic || !sparse_zero) ? nullptr : a->radii, a->P, im.ranges, im.tile_last,
                                      d.tiles, ordered ? im.order_bwd : nullptr, stream);
        if (rc0) return rc0;
    }
    if (a->num_rendered > 0) {
        BlendBwdParams b;
        b.W = a->W; b.H = a->H; b.grid_x = d.gx; b.grid_y = d.gy;
        b.ranges = im.ranges; b.point_list = bw.vals[final_buf]; b.rec = g.rec;
        b.point_offsets = g.point_offsets;
        b.background = a->background; b.final_T = im.final_T; b.n_contrib = im.n_contrib; b.tile_last = im.tile_last;
        b.dL_dpix = a->dL_dout_color; b.accum = ws.accum; b.partial = ws.partial; b.deterministic = a->deterministic;
        b.fast = a->blend_mode == 1;
        b.P = a->P; b.list_cap = a->num_rendered;
        b.tile_order = ordered ? im.order_bwd : nullptr;
        int rc;
        {
            ProfScope ps(B200S_STAGE_BLEND_BWD, stream);
            rc = launch_blend_bwd(b, blend_bwd_warps(d.tiles), stream);
        }
        if (rc) return rc;
        B200S_AFTER_LAUNCH(stream, a->debug);
    }
    PreprocessBwdParams p;
    p.P = a->P; p.D = a->D; p.M = a->M; p.W = a->W; p.H = a->H;
    p.tan_fovx = a->tan_fovx; p.tan_fovy = a->tan_fovy; p.focal_x = d.focal_x; p.focal_y = d.focal_y;
    p.scale_modifier = a->scale_modifier;
    p.means3D = a->means3D; p.shs = a->shs; p.colors_precomp = a->colors_precomp; p.scales = a->scales;
    p.rotations = a->rotations; p.cov3D_precomp = a->cov3D_precomp; p.viewmatrix = a->viewmatrix;
    p.projmatrix = a->projmatrix; p.campos = a->campos; p.radii = a->radii; p.rec = g.rec; p.clamped = g.clamped;
    p.point_offsets = g.point_offsets;
    p.accum = a->deterministic ? nullptr : ws.accum;
    p.partial = a->deterministic ? ws.partial : nullptr;
    p.dL_dmeans3D = a->dL_dmeans3D; p.dL_dmeans2D = a->dL_dmeans2D; p.dL_dopacity = a->dL_dopacity;
    p.dL_dsh = a->dL_dsh; p.dL_dsh_factors = a->shs ? a->dL_dsh_factors : nullptr; p.dL_dscales = a->dL_dscales; p.dL_drotations = a->dL_drotations;
    p.dL_dcolors = a->dL_dcolors; p.dL_dcov3D = a->dL_dcov3D; p.dL_dconic = a->dL_dconic;
    int rc;
    {
        ProfScope ps(B200S_STAGE_PREPROCESS_BWD, stream);
        rc = launch_preprocess_bwd(p, stream);
    }
    if (rc) return rc;
    B200S_AFTER_LAUNCH(stream, a->debug);
    return 0;
}

int b200s_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, b200s_stream_t stream) {
    (void)projmatrix;   // the reference's in_frustum only tests view-space depth (auxiliary.h:154)
    if (P < 0) return B200S_E_BADARG;
    if (P == 0) return 0;
    if (!means3D || !viewmatrix || !present) return B200S_E_BADARG;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

size_t b200s_idxmap_workspace_size(int32_t P, int32_t W, int32_t H) {
    return idxmap_workspace_size_impl(P, W, H);
}
int b200s_idxmap_forward(const B200sIdxmapArgs* args, b200s_stream_t stream) {
    ProfScope ps(B200S_STAGE_IDXMAP, (cudaStream_t)stream);
    return idxmap_forward_impl(args, (cudaStream_t)stream);
}
int b200s_idxmap_forward_static(const B200sIdxmapArgs* args, b200s_stream_t stream) {
    // no stage bracket here: the profiling hooks create events, and this entry point promises to create nothing
    return idxmap_forward_static_impl(args, (cudaStream_t)stream);
}
int b200s_idxmap_backward(int32_t P, int32_t W, int32_t H, const int32_t* out_idx, const float* grad_feature,
                          float* dL_dfeature_vector, b200s_stream_t stream) {
    ProfScope ps(B200S_STAGE_IDXMAP_BWD, (cudaStream_t)stream);
    return idxmap_backward_impl(P, W, H, out_idx, grad_feature, dL_dfeature_vector, (cudaStream_t)stream);
}

}  // extern "C"

extern "C" size_t b200s_peer_allreduce_flag_bytes(int32_t ctas, int32_t world) {
    if (ctas <= 0 || world <= 0) return 0;
    return (size_t)2 * (size_t)ctas * (size_t)world * sizeof(uint32_t);   // two barrier phases
}

extern "C" int64_t b200s_peer_allreduce_rows_per_rank(int32_t rows, int32_t world) {
    if (rows <= 0 || world <= 0) return 0;
    const int64_t per = ((int64_t)rows + world - 1) / world;
    return (per + 1023) / 1024 * 1024;
}

extern "C" size_t b200s_peer_allreduce_vis_bytes(int32_t rows, int32_t world) {
    return (size_t)b200s_peer_allreduce_rows_per_rank(rows, world) * (size_t)(world > 0 ? world : 0);
}

extern "C" int b200s_peer_allreduce_sum_f32(const B200sPeerReduceArgs* a, b200s_stream_t stream_) {
    if (!a || a->world < 1 || a->world > B200S_MAX_PEERS || a->rank < 0 || a->rank >= a->world) return B200S_E_BADARG;
    if (a->ctas < 1 || a->ctas > 1024 || (a->n_floats & 3u) || a->epoch == 0) return B200S_E_BADARG;
    if ((a->use_multimem & 3) && !a->multicast_buf) return B200S_E_BADARG;
    for (int r = 0; r < a->world; ++r)
        if (!a->peer_bufs[r] || !a->peer_flags[r] || ((uintptr_t)a->peer_bufs[r] & 15u)) return B200S_E_BADARG;
    if (a->radii) {
        if (a->sparse_rows <= 0 || (a->sparse_offset_floats & 3u) || (a->vis_offset_bytes & 3u)) return B200S_E_BADARG;
        if (a->sparse_offset_floats + (uint64_t)a->sparse_rows * 48u > a->n_floats) return B200S_E_BADARG;
    }
    if ((a->gather_floats & 3u) || (a->gather_offset_floats & 3u)) return B200S_E_BADARG;
    if (a->gather_floats && a->gather_offset_floats < a->n_floats) return B200S_E_BADARG;
    if ((a->n_floats == 0 && a->gather_floats == 0) || a->world == 1) return 0;
    return launch_peer_allreduce(*a, (cudaStream_t)stream_);
}

extern "C" int b200s_sh_grad_reconstruct(const B200sShGradArgs* a, b200s_stream_t stream_) {
    if (!a || a->P < 0 || a->views < 1 || a->views > B200S_MAX_PEERS || a->D < 0 || a->D > 3) return B200S_E_BADARG;
    if (a->P == 0) return 0;
    if (!a->means3D || !a->factors || !a->dL_dsh) return B200S_E_BADARG;
    if (a->M < (a->D + 1) * (a->D + 1) || a->M > 16) return B200S_E_BADARG;
    if (a->stride_floats < 3ull * (uint64_t)a->P + 4ull) return B200S_E_BADARG;
    if ((uintptr_t)a->dL_dsh & 15u) return B200S_E_BADARG;
    if (a->fast != 0 && a->fast != 1) return B200S_E_BADARG;
    return launch_sh_grad_reconstruct(*a, (cudaStream_t)stream_);
}

extern "C" int b200s_adam_step(const B200sAdamArgs* a, b200s_stream_t stream_) {
    if (!a || a->count < 0 || a->count > B200S_ADAM_MAX_TENSORS) return B200S_E_BADARG;
    if (!(a->bias_correction2_sqrt > 0.0f)) return B200S_E_BADARG;
    for (int i = 0; i < a->count; ++i) {
        const B200sAdamTensor& t = a->t[i];
        if (t.n < 0 || t.mode < B200S_ADAM_IDENTITY || t.mode > B200S_ADAM_NORMALIZE) return B200S_E_BADARG;
        if (t.n == 0) continue;
        if (!t.param || !t.exp_avg || !t.exp_avg_sq || !t.grad) return B200S_E_BADARG;
        if (((uintptr_t)t.param | (uintptr_t)t.exp_avg | (uintptr_t)t.exp_avg_sq | (uintptr_t)t.grad |
             (uintptr_t)t.activated) & 15u) return B200S_E_BADARG;
        if (t.mode == B200S_ADAM_SH && t.n % 48) return B200S_E_BADARG;
        if (t.mode == B200S_ADAM_NORMALIZE && t.n % 4) return B200S_E_BADARG;
    }
    return launch_adam_step(*a, (cudaStream_t)stream_);
}

extern "C" size_t b200s_loss_workspace_size(int32_t C_, int32_t H, int32_t W) {
    if (C_ <= 0 || H <= 0 || W <= 0) return 0;
    return loss_workspace_size(C_, H, W);
}

extern "C" int b200s_l1_ssim_loss(const B200sLossArgs* a, b200s_stream_t stream_) {
    if (!a || a->C <= 0 || a->H <= 0 || a->W <= 0) return B200S_E_BADARG;
    if (!a->image || !a->gt || !a->loss_out || !a->ws) return B200S_E_BADARG;
    if (a->ws_bytes < loss_workspace_size(a->C, a->H, a->W)) return B200S_E_WORKSPACE;
    return launch_l1_ssim(*a, (cudaStream_t)stream_);
}

extern "C" int b200s_count_pairs(const B200sForwardArgs* a, unsigned long long* counts_dev, b200s_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!a || !counts_dev) return B200S_E_BADARG;
    B200S_CUDA_OK(cudaMemsetAsync(counts_dev, 0, 3 * sizeof(unsigned long long), stream));
    if (a->P <= 0) return 0;
    if (!a->geom_ws || !a->img_ws || !a->bin_ws) return B200S_E_BADARG;
    const Dims d = dims_of(a->W, a->H, a->tan_fovx, a->tan_fovy);
    GeomWS g = carve_geom(a->geom_ws, a->P);
    ImgWS im = carve_img(a->img_ws, a->W, a->H);
    BinWS bw = carve_bin(a->bin_ws, a->bin_capacity);
    BlendFwdParams b;
    b.W = a->W; b.H = a->H; b.grid_x = d.gx; b.grid_y = d.gy;
    b.ranges = im.ranges; b.point_list = bw.vals[sort_num_passes(d.tiles) & 1]; b.rec = g.rec;
    b.background = a->background; b.final_T = im.final_T; b.n_contrib = im.n_contrib; b.tile_last = im.tile_last;
    b.out_color = a->out_color; b.fast = 0;
    b.P = a->P; b.list_cap = a->bin_capacity;
    b.tile_order = nullptr;
    return launch_count_pairs(b, counts_dev, stream);
}

// ---------------------------------------------------------------------------- state export
namespace {
__global__ void export_rec_kernel(int P, const SplatRec* rec, const int* radii, const uint8_t* clamped, int which,
                                  float* dst_f, uint8_t* dst_b) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const bool vis = radii[i] > 0;
    SplatRec r;
    r.q0 = r.q1 = r.q2 = make_float4(0, 0, 0, 0);
    if (vis) r = rec[i];
    switch (which) {
        case B200S_STATE_DEPTHS: dst_f[i] = r.q2.y; break;
        case B200S_STATE_MEANS2D: dst_f[2 * i] = r.q0.x; dst_f[2 * i + 1] = r.q0.y; break;
        case B200S_STATE_CONIC_OPACITY:
            dst_f[4 * i] = r.q0.z; dst_f[4 * i + 1] = r.q0.w; dst_f[4 * i + 2] = r.q1.x; dst_f[4 * i + 3] = r.q1.y; break;
        case B200S_STATE_RGB: dst_f[3 * i] = r.q1.z; dst_f[3 * i + 1] = r.q1.w; dst_f[3 * i + 2] = r.q2.x; break;
        case B200S_STATE_CLAMPED: {
            const uint8_t c = vis ? clamped[i] : 0;
            dst_b[3 * i] = c & 1; dst_b[3 * i + 1] = (c >> 1) & 1; dst_b[3 * i + 2] = (c >> 2) & 1; break;
        }
    }
}
// The sorted 64-bit keys are not materialised by the two-level sort (the blend only needs the
// sorted Gaussian ids): rebuild them as tile << 32 | depth bits of the sorted ids.
__global__ void export_keys_kernel(int tiles, const uint2* ranges, const uint32_t* point_list, const SplatRec* rec,
                                   unsigned long long* dst) {
    const int tile = blockIdx.x;
    if (tile >= tiles) return;
    const uint2 r = ranges[tile];
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
        dst[i] = ((unsigned long long)tile << 32) | (unsigned long long)__float_as_uint(rec[point_list[i]].q2.y);
}
}  // namespace

extern "C" int64_t b200s_export_state(const B200sForwardArgs* a, int32_t which, int64_t num_rendered, void* dst,
                                      int64_t dst_bytes, b200s_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!a || !dst) return B200S_E_BADARG;
    const int64_t P = a->P;
    const Dims d = dims_of(a->W, a->H, a->tan_fovx, a->tan_fovy);
    GeomWS g = carve_geom(a->geom_ws, P > 0 ? P : 1);
    ImgWS im = carve_img(a->img_ws, a->W, a->H);
    const int64_t px = (int64_t)a->W * a->H;
    const void* src = nullptr;
    int64_t bytes = 0;
    switch (which) {
        case B200S_STATE_DEPTHS: bytes = 4 * P; break;
        case B200S_STATE_MEANS2D: bytes = 8 * P; break;
        case B200S_STATE_CONIC_OPACITY: bytes = 16 * P; break;
        case B200S_STATE_RGB: bytes = 12 * P; break;
        case B200S_STATE_CLAMPED: bytes = 3 * P; break;
        case B200S_STATE_TILES_TOUCHED: src = g.tiles_touched; bytes = 4 * P; break;
        case B200S_STATE_POINT_OFFSETS: src = g.point_offsets; bytes = 4 * P; break;
        case B200S_STATE_RANGES: src = im.ranges; bytes = 8 * (int64_t)d.tiles; break;
        case B200S_STATE_N_CONTRIB: src = im.n_contrib; bytes = 4 * px; break;
        case B200S_STATE_FINAL_T: src = im.final_T; bytes = 4 * px; break;
        case B200S_STATE_POINT_LIST:
        case B200S_STATE_POINT_LIST_KEYS: {
            if (!a->bin_ws || num_rendered > a->bin_capacity) return B200S_E_BADARG;
            BinWS bw = carve_bin(a->bin_ws, a->bin_capacity);
            const int fb = sort_num_passes(d.tiles) & 1;
            if (which == B200S_STATE_POINT_LIST) { src = bw.vals[fb]; bytes = 4 * num_rendered; }
            else {
                bytes = 8 * num_rendered;
                if (bytes > dst_bytes) return B200S_E_WORKSPACE;
                if (bytes == 0) return 0;
                count_launch();
                export_keys_kernel<<<d.tiles, 256, 0, stream>>>(d.tiles, im.ranges, bw.vals[fb], g.rec,
                                                                (unsigned long long*)dst);
                if (cudaGetLastError() != cudaSuccess) return B200S_E_BADARG;
                return bytes;
            }
            break;
        }
        default: return B200S_E_BADARG;
    }
    if (bytes > dst_bytes) return B200S_E_WORKSPACE;
    if (bytes == 0) return 0;
    if (src) {
        if (cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, stream) != cudaSuccess) return B200S_E_BADARG;
    } else {
        count_launch();
        export_rec_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>((int)P, g.rec, a->radii, g.clamped, which,
                                                                         (float*)dst, (uint8_t*)dst);
        if (cudaGetLastError() != cudaSuccess) return B200S_E_BADARG;
    }
    return bytes;
}
