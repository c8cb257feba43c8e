// Kernel parameter blocks and host launchers shared between the translation units.
#pragma once
#include "common.cuh"

namespace b200s {

struct PreprocessParams {
    int P, D, M, W, H;
    int grid_x, grid_y;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
    int prefiltered;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;
    const float* scales;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    int* radii;
    SplatRec* rec;
    uint32_t* tiles_touched;
    uint32_t* depth_bits;
    uint8_t* clamped;
    B200sHeader* header;
    uint32_t* counters;
};

struct BlendFwdParams {
    int W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const SplatRec* rec;
    const float* background;
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* tile_last;
    float* out_color;
    const uint32_t* tile_order;   // blockIdx.x -> tile id (longest list first) or NULL = identity
    int fast;          // blend_mode 1: ex2.approx falloff
    int P;             // checked build only: bounds of point_list entries / of the list itself
    int64_t list_cap;
};

struct BlendBwdParams {
    int W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const SplatRec* rec;
    const uint32_t* point_offsets;   // deterministic mode: slot of an instance
    const float* background;
    const float* final_T;
    const uint32_t* n_contrib;
    const uint32_t* tile_last;
    const float* dL_dpix;
    float* accum;      // [P][ACC_STRIDE] atomic mode
    float* partial;    // [R][ACC_STRIDE] deterministic mode
    int deterministic;
    const uint32_t* tile_order;   // blockIdx.x -> tile id (longest list first) or NULL = identity
    int fast;          // blend_mode 1: must match the forward that produced the state
    int P;             // checked build only
    int64_t list_cap;
};

struct PreprocessBwdParams {
    int P, D, M, W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    const int* radii;
    const SplatRec* rec;
    const uint8_t* clamped;
    const uint32_t* point_offsets;
    const float* accum;       // [P][ACC_STRIDE] (atomic mode) or NULL
    const float* partial;     // [R][ACC_STRIDE] (deterministic mode) or NULL
    float* dL_dmeans3D;
    float* dL_dmeans2D;
    float* dL_dopacity;
    float* dL_dsh;
    float* dL_dsh_factors;   // [3P + 4] or NULL (factored SH gradient)
    float* dL_dscales;
    float* dL_drotations;
    float* dL_dcolors;
    float* dL_dcov3D;
    float* dL_dconic;
};

// measurement hooks (api.cu)
void prof_begin(int stage, cudaStream_t stream);
void prof_end(int stage, cudaStream_t stream);
void count_launch(int n = 1);
// Built with -DB200S_NVTX=1 (python -m neuralgaussiansplatting_b200.build --tag nvtx -DB200S_NVTX=1) every stage
// bracket also opens an NVTX range named after the stage, so an nsys / ncu --nvtx timeline shows K1..K16 by name.
// Off by default: the product library carries no NVTX calls.
#if defined(B200S_NVTX) && B200S_NVTX
void nvtx_push(int stage);
void nvtx_pop();
#else
static inline void nvtx_push(int) {}
static inline void nvtx_pop() {}
#endif
struct ProfScope {
    int stage; cudaStream_t stream;
    ProfScope(int st, cudaStream_t s) : stage(st), stream(s) { nvtx_push(stage); prof_begin(stage, stream); }
    ~ProfScope() { prof_end(stage, stream); nvtx_pop(); }
    ProfScope(const ProfScope&) = delete;
    ProfScope& operator=(const ProfScope&) = delete;
};
int launch_count_pairs(const BlendFwdParams& p, unsigned long long* counts, cudaStream_t stream);

int launch_preprocess_fwd(const PreprocessParams& p, cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream);
int launch_scan_and_depth_sort(int P, const GeomWS& g, bool debug, cudaStream_t stream);
int sort_num_passes(int tiles);
int launch_binning(int P, const GeomWS& g, const BinWS& bw, int64_t capacity, int64_t n_hint, int grid_x,
                   int tiles, uint2* ranges, int* final_buf, bool debug, cudaStream_t stream);
int launch_tile_order(const uint2* ranges, const uint32_t* tile_last, int tiles, uint32_t* order, cudaStream_t stream);
bool tile_order_enabled();
bool tile_order_fwd_enabled();
int launch_bwd_prologue(void* zero, size_t zero_bytes, const int* radii, int P, const uint2* ranges,
                        const uint32_t* tile_last, int tiles, uint32_t* order, cudaStream_t stream);
int launch_blend_fwd(const BlendFwdParams& p, int nw, cudaStream_t stream);
int launch_blend_bwd(const BlendBwdParams& p, int nw, cudaStream_t stream);
int launch_preprocess_bwd(const PreprocessBwdParams& p, cudaStream_t stream);
int launch_peer_allreduce(const B200sPeerReduceArgs& a, cudaStream_t stream);
int launch_sh_grad_reconstruct(const B200sShGradArgs& a, cudaStream_t stream);
int launch_adam_step(const B200sAdamArgs& a, cudaStream_t stream);
size_t loss_workspace_size(int C, int H, int W);
int launch_l1_ssim(const B200sLossArgs& a, cudaStream_t stream);

}  // namespace b200s
