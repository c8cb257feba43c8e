/*
 * b200splat — C ABI of the B200-native (sm_100a) differentiable Gaussian-splatting
 * rasterizer. This is the drop-in boundary: every entry point below is what the
 * reference's torch binding for this path would bind instead of
 * CudaRasterizer::Rasterizer (reference file:line cited per function).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; no torch / C++ types cross the ABI.
 *  - Every pointer is a DEVICE pointer borrowed for the duration of the call
 *    (stream-ordered) unless the name ends in _host. The library never allocates
 *    or frees device memory; the caller sizes workspaces with the *_workspace_size
 *    queries and passes them in (the reference instead calls back into three
 *    std::function<char*(size_t)> allocators, cuda_rasterizer/rasterizer.h:38-40).
 *  - All kernels are launched on the explicit `stream` (the reference launches on
 *    the legacy default stream).
 *  - Return value: 0 = ok; >0 = CUDA error code (cudaError_t) captured with
 *    cudaGetLastError / cudaStreamSynchronize; <0 = B200S_E_* argument error.
 *    Nothing throws across the ABI. With args->debug != 0 each stage is followed
 *    by a stream synchronise + error check (reference CHECK_CUDA,
 *    cuda_rasterizer/auxiliary.h:166-174).
 *  - "Absent" optional inputs are NULL (the reference binding passes the data_ptr
 *    of an empty tensor, which is nullptr: rasterizer_impl.cu:321,389,411).
 *  - Matrices are 16 floats consumed column-major exactly as the reference does
 *    (auxiliary.h:58-76): element [i + 4*j] multiplies input component j for
 *    output component i.
 */
#ifndef B200SPLAT_H_INCLUDED
#define B200SPLAT_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200S_ABI_VERSION 3

/* tile size is fixed by the reference (cuda_rasterizer/config.h:16-17) */
#define B200S_BLOCK_X 16
#define B200S_BLOCK_Y 16
#define B200S_NUM_CHANNELS 3   /* config.h:15 */
#define B200S_NUM_FEATURES 64  /* rasterizer2 cuda_rasterizer/config.h:3 */

enum {
    B200S_OK = 0,
    B200S_E_BADARG = -1,       /* NULL / inconsistent arguments */
    B200S_E_WORKSPACE = -2,    /* a workspace is smaller than *_workspace_size says */
    B200S_E_OVERFLOW = -3,     /* num_rendered exceeds the binning capacity */
    B200S_E_UNSUPPORTED = -4,  /* e.g. SH without 3 channels (rasterizer_impl.cu:242-245) */
    B200S_E_INTERNAL = -5      /* device-side watchdog tripped (look-back spin bound) */
};

typedef void* b200s_stream_t; /* cudaStream_t */

/* Header the forward writes at the start of the geometry workspace (device) and
 * that b200s_read_header copies back. */
typedef struct B200sHeader {
    uint32_t num_rendered;   /* sum of tiles_touched == reference's return value */
    uint32_t overflow;       /* 1 if num_rendered > binning capacity (nothing was binned) */
    uint32_t error;          /* device-side watchdog flags, 0 when healthy */
    uint32_t num_visible;    /* Gaussians with radii > 0 */
    uint32_t reserved[12];
} B200sHeader;

/* ---- forward: replaces CudaRasterizer::Rasterizer::forward
 *      (diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:198-336) and
 *      the argument marshalling of RasterizeGaussiansCUDA (rasterize_points.cu:35-115).
 *      rasterizer2's stripped copy (rasterizer2/.../rasterizer_impl.cu:300-443) is the
 *      same call with colors_precomp == cov3D_precomp == NULL. */
typedef struct B200sForwardArgs {
    int32_t P;               /* number of Gaussians */
    int32_t D;               /* active SH degree 0..3 */
    int32_t M;               /* SH coefficients per Gaussian (0 when shs == NULL) */
    int32_t W, H;            /* image size in pixels */
    float tan_fovx, tan_fovy;
    float scale_modifier;
    int32_t prefiltered;     /* auxiliary.h:156-160: culled-but-prefiltered sets header.error */
    int32_t debug;
    const float* background;      /* [3] */
    const float* means3D;         /* [P,3] */
    const float* shs;             /* [P,M,3] or NULL */
    const float* colors_precomp;  /* [P,3] or NULL (exactly one of shs/colors_precomp) */
    const float* opacities;       /* [P] */
    const float* scales;          /* [P,3] or NULL */
    const float* rotations;       /* [P,4] or NULL */
    const float* cov3D_precomp;   /* [P,6] or NULL (exactly one of scales+rot / cov3D) */
    const float* viewmatrix;      /* [16] */
    const float* projmatrix;      /* [16] */
    const float* campos;          /* [3] */
    float* out_color;             /* [3,H,W] fully written */
    int32_t* radii;               /* [P] fully written */
    void* geom_ws;  size_t geom_bytes;      /* >= b200s_geometry_workspace_size(P) */
    void* img_ws;   size_t img_bytes;       /* >= b200s_image_workspace_size(W,H) */
    void* bin_ws;   size_t bin_bytes;       /* >= b200s_binning_workspace_size(capacity,W,H) */
    int64_t bin_capacity;                   /* instances the binning workspace can hold */
    int32_t blend_mode;      /* 0 = strict (expf as the reference, bit-exact n_contrib — the parity gate);
                              * 1 = fast: falloff as ex2.approx(power*log2 e) on a pre-scaled conic; image within 1e-5,
                              *     n_contrib may differ on isolated threshold pixels. b200s_backward must be called
                              *     with the same blend_mode as the forward whose state it consumes. */
    int32_t reserved;
} B200sForwardArgs;

size_t b200s_geometry_workspace_size(int32_t P);
size_t b200s_image_workspace_size(int32_t W, int32_t H);
size_t b200s_binning_workspace_size(int64_t capacity, int32_t W, int32_t H);
/* Inverse of the above: the capacity a binning workspace of `bytes` bytes was sized for
 * (lets the backward re-derive the layout from the opaque buffer alone, as the reference
 * does with BinningState::fromChunk(ptr, R), rasterizer_impl.cu:372). */
int64_t b200s_binning_capacity_for_bytes(size_t bytes);

/* Stage 1: fused preprocess (cull, EWA covariance/conic, SH, tile rect) + prefix sum.
 * Leaves header.num_rendered on the device. Never blocks. */
int b200s_forward_preprocess(const B200sForwardArgs* args, b200s_stream_t stream);
/* Stage 2: duplicateWithKeys, radix sort, identifyTileRanges, forward blend. Reads
 * num_rendered from the device; if it exceeds bin_capacity it sets header.overflow and
 * renders nothing (caller re-runs stage 2 with a larger workspace). Never blocks. */
int b200s_forward_render(const B200sForwardArgs* args, b200s_stream_t stream);
/* Both stages back to back. */
int b200s_forward(const B200sForwardArgs* args, b200s_stream_t stream);
/* Asynchronous copy of the header into (pinned) host memory. */
int b200s_read_header(const void* geom_ws, B200sHeader* header_host, b200s_stream_t stream);
/* View batch (the loop of render.py:24-35 / the view loop of train.py:75-93 without gradients):
 * `count` forwards enqueued back to back on one stream with NO host synchronisation in between. The views may
 * (and normally do) share ONE geometry / image / binning workspace — they run in stream order — and differ in
 * camera, out_color and radii pointers. After view i its header is copied into headers_host[i] (pinned memory);
 * the caller synchronises ONCE after the call and then checks every header: `error`, and `overflow` /
 * num_rendered > bin_capacity (that view rendered nothing: re-run it alone with a larger binning workspace). */
int b200s_forward_batch(const B200sForwardArgs* views, int32_t count, B200sHeader* headers_host,
                        b200s_stream_t stream);

/* ---- backward: replaces CudaRasterizer::Rasterizer::backward
 *      (rasterizer_impl.cu:340-434) and RasterizeGaussiansBackwardCUDA
 *      (rasterize_points.cu:117-196). The library writes EVERY element of every
 *      non-NULL output (zeros for culled Gaussians), so the caller may allocate them
 *      uninitialised — the reference's nine torch::zeros (rasterize_points.cu:151-159)
 *      are not needed. Optional outputs may be NULL. */
typedef struct B200sBackwardArgs {
    int32_t P, D, M, W, H;
    float tan_fovx, tan_fovy;
    float scale_modifier;
    int32_t debug;
    int32_t deterministic;   /* 1 = fixed-order reduction, bit-reproducible gradients */
    const float* background;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;       /* [P] (needed only by the fast blend mode; may be NULL) */
    const float* scales;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    const int32_t* radii;         /* [P] from the forward */
    const float* dL_dout_color;   /* [3,H,W] */
    const void* geom_ws;  size_t geom_bytes;   /* the forward's workspaces, unmodified */
    const void* img_ws;   size_t img_bytes;
    const void* bin_ws;   size_t bin_bytes;
    int64_t bin_capacity;
    void* bwd_ws;  size_t bwd_bytes;           /* >= b200s_backward_workspace_size(P, R) */
    int64_t num_rendered;                      /* R, as returned by the forward */
    float* dL_dmeans3D;      /* [P,3] */
    float* dL_dmeans2D;      /* [P,3] (z component is written as 0) */
    float* dL_dopacity;      /* [P]   */
    float* dL_dsh;           /* [P,M,3] or NULL */
    float* dL_dscales;       /* [P,3] or NULL */
    float* dL_drotations;    /* [P,4] or NULL */
    float* dL_dcolors;       /* [P,3] or NULL: gradient w.r.t. per-Gaussian RGB */
    float* dL_dcov3D;        /* [P,6] or NULL */
    float* dL_dconic;        /* [P,4] or NULL: (a, b, 0, c) as the reference lays it out */
    int32_t blend_mode;
    int32_t reserved;
    /* Factored SH gradient (optional, shs path only). dL/dsh[k][c] of one view is the outer product
     * basis_k(dir) * dL/dRGB[c] (backward.cu:20-139), so a view-parallel job need not move 48 floats per
     * Gaussian: when non-NULL the backward writes the 3 post-clamp dL/dRGB floats per Gaussian to
     * dL_dsh_factors[3*i .. 3*i+2] (zeros for culled Gaussians) and this view's camera centre to
     * dL_dsh_factors[3*P .. 3*P+2]; b200s_sh_grad_reconstruct rebuilds (and sums over views) dL/dsh from such
     * factor blocks bit-identically. dL_dsh may then be NULL (the 192-B rows are not written). */
    float* dL_dsh_factors;   /* [3*P + 4] or NULL */
} B200sBackwardArgs;

size_t b200s_backward_workspace_size(int32_t P, int64_t num_rendered);
int b200s_backward(const B200sBackwardArgs* args, b200s_stream_t stream);

/* ---- markVisible: replaces Rasterizer::markVisible (rasterizer_impl.cu:141-153) */
int b200s_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                       const float* projmatrix, uint8_t* present, b200s_stream_t stream);

/* ---- rasterizer2 "idxmap" extra channels: replaces Rasterizer::getidxmap /
 *      getidxmap_backward (rasterizer2/cuda_rasterizer/rasterizer_impl.cu:851-1008)
 *      and GetIdxMapCUDA / GetIdxMapBackwardCUDA (rasterizer2/rasterize_points.cu:183-295). */
typedef struct B200sIdxmapArgs {
    int32_t P, W, H;
    int32_t debug;
    const float* means3D;         /* [P,3] */
    const float* viewmatrix;      /* [16] */
    const float* projmatrix;      /* [16] */
    const float* campos;          /* [3] */
    const float* feature_vector;  /* [P,64] */
    int32_t* out_idx;             /* [H,W]    fully written; -1 = no point */
    float* out_color;             /* [H,W,3]  fully written (zeros where no point) */
    float* out_depth;             /* [H,W]    fully written */
    float* out_feature;           /* [H,W,64] fully written */
    int32_t* radii;               /* [P] 1 if the point's centre is on screen else 0 */
    void* ws; size_t ws_bytes;    /* >= b200s_idxmap_workspace_size(P,W,H) */
    uint32_t* num_rendered_dev;   /* optional [1]: sum of clipped footprint areas */
} B200sIdxmapArgs;

size_t b200s_idxmap_workspace_size(int32_t P, int32_t W, int32_t H);
int b200s_idxmap_forward(const B200sIdxmapArgs* args, b200s_stream_t stream);
/* dL_dfeature_vector [P,64] is fully written (zeros + scatter-add of channels 25..63). */
int b200s_idxmap_backward(int32_t P, int32_t W, int32_t H, const int32_t* out_idx,
                          const float* grad_feature, float* dL_dfeature_vector,
                          b200s_stream_t stream);

/* ---- idxmap forward without a caller workspace. No counterpart in the reference (its binding
 *      allocates through torch on every call); added for hosts of this library that cannot
 *      hand over a workspace: a render loop embedded in an application whose device memory is
 *      already owned in full by another allocator (a pre-sized arena, a framework pool capped
 *      at a memory fraction), or a loop that must not run into cudaMalloc's device-wide
 *      synchronisation. Same inputs, outputs, results and status codes as the workspace variant
 *      above; args->ws and args->ws_bytes are ignored. The 8-byte-per-pixel z-buffer is a fixed
 *      array inside the library image instead, so:
 *        - W*H may be at most B200S_IDXMAP_STATIC_MAX_PIXELS (2^21, enough for 1920x1080);
 *          a larger view returns B200S_E_WORKSPACE before anything is enqueued;
 *        - every GPU that loads the library holds those 16 MiB until the process exits;
 *        - the array is one per GPU and shared by all calls: it is re-initialised at the start
 *          of each call, and two calls that overlap on one GPU (two streams, two threads)
 *          corrupt each other. Serialise them;
 *        - the device the stream belongs to must be the current device. The first call on each
 *          device looks up the array's device address (cudaGetSymbolAddress, checked); make
 *          that first call before starting any stream capture;
 *        - the call itself performs no cudaMalloc / cudaFree and creates no events (it is also
 *          left out of the stage timings of the profiling hooks for that reason). */
#define B200S_IDXMAP_STATIC_MAX_PIXELS 2097152
int b200s_idxmap_forward_static(const B200sIdxmapArgs* args, b200s_stream_t stream);

/* ---- introspection for parity tests: copies one private state array of the last
 *      forward into a device buffer the caller owns. `which` is a B200S_STATE_* id;
 *      returns bytes written or a negative error. */
enum {
    B200S_STATE_DEPTHS = 0,        /* f32[P]  (0 for culled) */
    B200S_STATE_MEANS2D = 1,       /* f32[2P] */
    B200S_STATE_CONIC_OPACITY = 2, /* f32[4P] */
    B200S_STATE_RGB = 3,           /* f32[3P] */
    B200S_STATE_TILES_TOUCHED = 4, /* u32[P]  */
    B200S_STATE_POINT_OFFSETS = 5, /* u32[P]  */
    B200S_STATE_POINT_LIST = 6,    /* u32[R]  sorted values */
    B200S_STATE_POINT_LIST_KEYS = 7, /* u64[R] sorted keys */
    B200S_STATE_RANGES = 8,        /* u32[2*tiles] */
    B200S_STATE_N_CONTRIB = 9,     /* u32[W*H] */
    B200S_STATE_FINAL_T = 10,      /* f32[W*H] */
    B200S_STATE_CLAMPED = 11       /* u8[3P] */
};
int64_t b200s_export_state(const B200sForwardArgs* args, int32_t which, int64_t num_rendered,
                           void* dst, int64_t dst_bytes, b200s_stream_t stream);

/* ---- measurement hooks (bench.py / tests only; never needed for results).
 * Stage timing: between b200s_profile_begin() and b200s_profile_end() every stage of every
 * forward / backward call is bracketed by cudaEvents on the caller's stream (no
 * synchronisation until _end, which waits for the last event and sums per stage).
 * b200s_launch_count() is the cumulative number of kernels this library has launched. */
enum {
    B200S_STAGE_PREPROCESS = 0,    /* fused preprocess forward (K1) */
    B200S_STAGE_SCAN = 1,          /* prefix sum of tiles_touched (K2) */
    B200S_STAGE_EXPAND = 2,        /* duplicateWithKeys (K3) */
    B200S_STAGE_SORT = 3,          /* histogram + onesweep passes (K4) */
    B200S_STAGE_RANGES = 4,        /* identifyTileRanges (K5) */
    B200S_STAGE_BLEND_FWD = 5,     /* forward compositing (K6) */
    B200S_STAGE_BLEND_BWD = 6,     /* backward compositing (K7) */
    B200S_STAGE_PREPROCESS_BWD = 7,/* fused preprocess backward (K8+K9) */
    B200S_STAGE_IDXMAP = 8,        /* idxmap forward (K11..K15) */
    B200S_STAGE_IDXMAP_BWD = 9,    /* idxmap backward (K16) */
    B200S_STAGE_MISC = 10,         /* memsets and small bookkeeping kernels */
    B200S_NUM_STAGES = 11
};
int b200s_profile_begin(void);
/* ms[B200S_NUM_STAGES] (summed elapsed time) and calls[B200S_NUM_STAGES] (brackets recorded). */
int b200s_profile_end(float* ms, int32_t* calls);
uint64_t b200s_launch_count(void);

/* Roofline accounting: replays the per-pixel traversal of the last forward and counts
 * counts_dev[0] = (pixel, Gaussian) pairs the forward traverses (reference `contributor` at
 * exit, forward.cu:325-351), [1] = pairs that contribute, [2] = pairs the backward traverses
 * (sum of n_contrib). counts_dev: device u64[3]. */
int b200s_count_pairs(const B200sForwardArgs* args, unsigned long long* counts_dev, b200s_stream_t stream);

/* ---- gradient all-reduce over NVLink peer memory (the one collective of the
 *      path; the reference has no multi-GPU code, utils/general_utils.py:133 — the baseline
 *      this replaces is ncclAllReduce(sum, fp32) over the 59 gradient floats per Gaussian).
 *      One launch: cross-rank start barrier, two-shot fixed-order sum (rank r reduces slice r
 *      from every peer and stores the total to every peer), end barrier. Every rank of the
 *      group must call it with the same n_floats / ctas / epoch, in the same order.
 *      peer_bufs[r]  : rank r's gradient buffer as mapped into THIS process (symmetric
 *                      allocation; peer_bufs[rank] is the local one), 16-byte aligned.
 *      peer_flags[r] : rank r's flag array (u32, >= b200s_peer_allreduce_flag_bytes, zeroed
 *                      once at allocation), same mapping rules.
 *      multicast_buf : NVSwitch multicast mapping of the buffer or NULL; used only when
 *                      use_multimem != 0. Bit 0: the summed prefix goes through multimem.ld_reduce /
 *                      multimem.st (the switch sums, in ITS order: replicas identical, not the
 *                      rank-order sum). Bit 1: the all-gather slots are PUSHED with multimem.st (a pure
 *                      copy, exact) instead of pulled with peer loads.
 *      epoch         : strictly increasing per call (1, 2, ...) — flags are never reset.
 *      In-place: after the call (stream-ordered) every rank's buffer holds the sum. */
#define B200S_MAX_PEERS 16
typedef struct B200sPeerReduceArgs {
    void* peer_bufs[B200S_MAX_PEERS];
    void* peer_flags[B200S_MAX_PEERS];
    void* multicast_buf;
    uint64_t n_floats;       /* multiple of 4 */
    int32_t rank, world;
    uint32_t epoch;
    int32_t ctas;            /* CTAs of 512 threads the collective may occupy (1..1024) */
    int32_t use_multimem;
    /* Visibility-sparse tail (optional; radii == NULL reduces everything densely). The LAST tensor of the
     * buffer, floats [sparse_offset_floats, sparse_offset_floats + 48*sparse_rows) with at most alignment padding
     * (left untouched) up to n_floats, is sparse_rows rows of 48 floats (dL/dSH, degree 3).
     * radii: THIS rank's radii[sparse_rows] (int32, the forward's output for the view the gradients belong
     * to): rows with radii <= 0 are all-zero on this rank (b200s_backward zero-fills them) and are neither read
     * by the owners nor, if no rank saw them, written. vis_offset_bytes: offset from peer_bufs[r] of a
     * b200s_peer_allreduce_vis_bytes() scratch array inside the same peer mapping (zeroed once at allocation).
     * Used for world in {2, 4, 8} without multimem; the result is bit-identical to the dense reduction. */
    int32_t sparse_rows;
    const void* radii;
    uint64_t sparse_offset_floats;   /* multiple of 4 */
    uint64_t vis_offset_bytes;       /* multiple of 4 */
    /* All-gather region (optional, gather_floats == 0: none). Besides the summed prefix [0, n_floats) every
     * buffer holds `world` slots of gather_floats floats at gather_offset_floats; rank r has filled slot r of
     * ITS buffer before the call, and after it every rank's slot r holds rank r's data (each rank pulls the
     * world-1 foreign slots over NVLink inside the same launch, between the same two barriers). Used for the
     * factored SH gradient: 3 floats per Gaussian travel instead of 48 (see b200s_sh_grad_reconstruct). */
    uint64_t gather_offset_floats;   /* multiple of 4, >= n_floats */
    uint64_t gather_floats;          /* per-rank slot length, multiple of 4 */
} B200sPeerReduceArgs;
size_t b200s_peer_allreduce_flag_bytes(int32_t ctas, int32_t world);
/* rows of the sparse tensor each rank owns (a multiple of the kernel's 1024-row work unit) and the size of
 * the visibility scratch array (world * rows_per_rank bytes). */
int64_t b200s_peer_allreduce_rows_per_rank(int32_t rows, int32_t world);
size_t b200s_peer_allreduce_vis_bytes(int32_t rows, int32_t world);
int b200s_peer_allreduce_sum_f32(const B200sPeerReduceArgs* args, b200s_stream_t stream);

/* ---- factored SH gradient: dL/dsh summed over `views` factor blocks (the view-parallel ranks of one step after
 *      the gather above, or several local views). Replaces moving / all-reducing the materialised dL/dsh rows
 *      (48 of the 59 gradient floats per Gaussian; computeColorFromSH backward, backward.cu:20-139):
 *          dL_dsh[i][k][c] = sum over v in view order of  fl( basis_k(normalize(means3D[i] - campos_v)) * g_v[i][c] )
 *      where block v = factors + v*stride_floats holds g_v [P,3] followed by campos_v at float 3*P, exactly as
 *      b200s_backward wrote them (dL_dsh_factors). Terms with g_v[i] == (0,0,0) are skipped, products and sums are
 *      single fp32 roundings in view order, so the result equals the fixed-order sum of the per-view dL/dsh rows
 *      b200s_backward would have written (same bits up to the sign of zero). Every element of dL_dsh is written. */
typedef struct B200sShGradArgs {
    int32_t P, D, M, views;
    const float* means3D;        /* [P,3] */
    const float* factors;        /* [views][stride_floats] */
    uint64_t stride_floats;      /* >= 3*P + 4 */
    float* dL_dsh;               /* [P,M,3], 16-byte aligned */
    int32_t fast;                /* 0: the exact form above (separate product and sum roundings, IEEE sqrt / divide).
                                  * 1: fused multiply-adds and an rsqrt-based direction: about half the instructions;
                                  *    the same sum to ~1e-7 relative, identical on every rank that holds the same
                                  *    factors (for collectives whose dense part is not a rank-order sum anyway). */
    int32_t reserved_;
} B200sShGradArgs;
int b200s_sh_grad_reconstruct(const B200sShGradArgs* args, b200s_stream_t stream);

/* ---- fused activation-backward + Adam + activation-forward for the replicated Gaussian
 *      parameters. Replaces, for the Gaussian tensors, autograd's chain
 *      through get_opacity / get_scaling / get_rotation / get_features (scene/gaussian_model.py:113-141),
 *      torch.optim.Adam(l, lr=0.0, eps=1e-15).step() over the param groups of training_setup
 *      (scene/gaussian_model.py:252-279, train.py:127-128) and the next iteration's activations
 *      + torch.cat. One launch over up to B200S_ADAM_MAX_TENSORS tensors, all float32, 16-byte
 *      aligned, updated in place. `grad` is dL/d(ACTIVATED value) exactly as b200s_backward emits it
 *      (after the all-reduce); `activated` receives activation(param_new) for the next forward (NULL
 *      or == param: not written). Adam arithmetic follows torch's single-tensor path operation by
 *      operation: m.lerp_(g, 1-b1); v.mul_(b2).addcmul_(g, g, 1-b2); denom = sqrt(v)/bc2_sqrt + eps;
 *      p.addcdiv_(m, denom, -step_size), with step_size = lr / (1 - b1^t) computed by the caller in
 *      double (as torch does) and passed negated. */
#define B200S_ADAM_MAX_TENSORS 8
enum {
    B200S_ADAM_IDENTITY = 0,   /* xyz */
    B200S_ADAM_SH = 1,         /* [P,16,3]: coefficient 0 steps with neg_step (f_dc), 1..15 with neg_step_rest (f_rest) */
    B200S_ADAM_EXP = 2,        /* scaling */
    B200S_ADAM_SIGMOID = 3,    /* opacity */
    B200S_ADAM_NORMALIZE = 4   /* rotation, rows of 4 (torch.nn.functional.normalize, eps 1e-12) */
};
typedef struct B200sAdamTensor {
    float* param;              /* raw (pre-activation) parameter */
    float* exp_avg;
    float* exp_avg_sq;
    const float* grad;         /* dL/d activated */
    float* activated;          /* out: activation(param) after the update, or NULL */
    int64_t n;                 /* floats (SH: multiple of 48, NORMALIZE: multiple of 4) */
    float neg_step;            /* -(lr / bias_correction1) */
    float neg_step_rest;       /* SH only */
    int32_t mode;
    int32_t reserved;
} B200sAdamTensor;
typedef struct B200sAdamArgs {
    B200sAdamTensor t[B200S_ADAM_MAX_TENSORS];
    int32_t count;
    float beta1, beta2;
    float one_minus_beta1, one_minus_beta2;   /* float(1 - beta) computed in double, as torch passes them */
    float eps;
    float bias_correction2_sqrt;               /* sqrt(1 - beta2^step) */
    int32_t reserved;
} B200sAdamArgs;
int b200s_adam_step(const B200sAdamArgs* args, b200s_stream_t stream);

/* ---- fused photometric loss: value and gradient of
 *      (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt))
 *      as train.py:91-92 forms it from utils/loss_utils.py:17-18 (l1_loss) and :27-63 (gaussian,
 *      create_window, ssim, _ssim: 11x11 window, sigma 1.5, zero padding 5, C1 = 0.01^2, C2 = 0.03^2,
 *      mean over all C*H*W). image, gt: [C,H,W] float32. loss_out: float[3] = {loss, l1, ssim}.
 *      dL_dimage: [C,H,W] = grad_scale * dloss/dimage, or NULL for the value only. ws: scratch of
 *      b200s_loss_workspace_size bytes (three derivative maps + per-block partial sums, which are added in
 *      index order: the value is reproducible run to run). */
typedef struct B200sLossArgs {
    int32_t C, H, W;
    float lambda_dssim;
    float grad_scale;            /* upstream dL/dloss (1 for loss.backward()) */
    int32_t reserved;
    const float* image;
    const float* gt;
    float* loss_out;
    float* dL_dimage;
    void* ws; size_t ws_bytes;
} B200sLossArgs;
size_t b200s_loss_workspace_size(int32_t C, int32_t H, int32_t W);
int b200s_l1_ssim_loss(const B200sLossArgs* args, b200s_stream_t stream);

/* sizeof() of the ABI structs as compiled: 0 B200sHeader, 1 B200sForwardArgs,
 * 2 B200sBackwardArgs, 3 B200sIdxmapArgs, 4 B200sPeerReduceArgs,
 * 5 B200sAdamArgs, 6 B200sLossArgs, 7 B200sShGradArgs. Bindings assert their own layout against it. */
size_t b200s_sizeof(int32_t which);

/* Library / build identification. */
int b200s_abi_version(void);
const char* b200s_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* B200SPLAT_H_INCLUDED */
