// K11..K16 — rasterizer2's "idxmap" extra channels: nearest point per pixel plus
// depth / view-direction encoding / learned feature gather, and the feature-gradient scatter.
// Replaces RASTER::preprocess, duplicateIdxDepthImg, the 63-bit SortPairs, identifyPixelRanges,
// GETMAP and CopyFeature (reference rasterizer2/cuda_rasterizer/raster.cu:9-97,
// rasterizer_impl.cu:26-42,578-615,680-847,851-1008).
//
// The reference expands every point into one (x<<48|y<<32|depth, idx) pair per covered pixel,
// radix-sorts ALL pairs on 63 bits and keeps only the first entry of every pixel. The first
// entry of a pixel's sorted run is the pair with the smallest (depth bits, insertion order)
// and insertion order is ascending point index, so the winner is exactly
//        min over covering points of  (depth_bits << 32 | idx)
// which a 64-bit atomicMin z-buffer computes with no sort at all (8 B/pixel instead of
// ~200 B/instance of sort traffic) and with identical tie-breaking.

#include <mutex>
#include "kernels.cuh"

namespace b200s {

constexpr unsigned long long ZBUF_EMPTY = ~0ull;
constexpr int NFEAT = B200S_NUM_FEATURES;   // 64
constexpr int NENC = 24;                    // 3 dims x 4 frequencies x (sin, cos)
constexpr int FEAT_FIRST_LEARNED = 25;      // channels 25..63 are copied from feature_vector

struct IdxWS {
    unsigned long long* zbuf;   // [W*H]
    size_t bytes;
};
static IdxWS carve_idx(void* base, int W, int H) {
    IdxWS s;
    s.zbuf = (unsigned long long*)base;
    s.bytes = align_up(8 * (size_t)W * H);
    return s;
}

// K11 + K12 fused: footprint of every point, splatted straight into the z-buffer.
// (raster.cu:27-61 and rasterizer_impl.cu:578-615.) The footprint geometry follows the
// reference expression by expression; all of it feeds integer outputs.
__global__ void __launch_bounds__(256)
idx_splat_kernel(int P, int W, int H, const float* __restrict__ means3D, const float* __restrict__ view,
                 const float* __restrict__ proj, int* __restrict__ radii, unsigned long long* zbuf,
                 uint32_t* num_rendered) {
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    uint32_t touched = 0;
    if (idx < P) {
        const float px = means3D[3 * idx], py = means3D[3 * idx + 1], pz = means3D[3 * idx + 2];
        int rad = 0;
        const float depth = affine_row_ref(view, 2, px, py, pz);
        if (!(depth <= 0.2f)) {
            const float hx = affine_row_ref(proj, 0, px, py, pz);
            const float hy = affine_row_ref(proj, 1, px, py, pz);
            const float hw = affine_row_ref(proj, 3, px, py, pz);
            const float p_w = __frcp_rn(fadd(hw, 0.0000001f));
            const float fx = ndc2pix_ref(fmul(hx, p_w), W);
            const float fy = ndc2pix_ref(fmul(hy, p_w), H);
            const int cx = (int)fx, cy = (int)fy;                     // (int)ndc2Pix(...) raster.cu:47
            const float rf = __fdiv_rn(3.0f, depth);                  // S / p_view.z, S = 3 (raster.cu:49,82)
            const int x0 = (int)fmaxf(0.0f, fsub(fx, rf));
            const int y0 = (int)fmaxf(0.0f, fsub(fy, rf));
            const int x1 = (int)fminf((float)W, fadd(fadd(fx, rf), 1.0f));
            const int y1 = (int)fminf((float)H, fadd(fadd(fy, rf), 1.0f));
            if (!(cx < 0 || cx >= W || cy < 0 || cy >= H)) {
                rad = 1;
                touched = (uint32_t)((x1 - x0) * (y1 - y0));
                const unsigned long long key =
                    ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned long long)(uint32_t)idx;
                for (int x = x0; x < x1; ++x)
                    for (int y = y0; y < y1; ++y) atomicMin(&zbuf[(size_t)y * W + x], key);
            }
        }
        radii[idx] = rad;
    }
    if (num_rendered != nullptr) {
        uint32_t s = touched;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL_MASK, s, o);
        if ((threadIdx.x & 31) == 0 && s) atomicAdd(num_rendered, s);
    }
}

// K15 GETMAP (rasterizer_impl.cu:731-817): one warp resolves 32 consecutive pixels; each lane
// computes its pixel's depth + 24 encoding values into shared memory, then the warp writes
// every pixel's 64-float feature row as two fully coalesced 128-B stores and gathers the
// learned channels 25..63 with coalesced loads.
__global__ void __launch_bounds__(256)
idx_resolve_kernel(int P, int W, int H, const unsigned long long* __restrict__ zbuf,
                   const float* __restrict__ means3D, const float* __restrict__ campos,
                   const float* __restrict__ feature_vector, int* __restrict__ out_idx,
                   float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_feature) {
    __shared__ float s_enc[8][32][NENC + 1 + 1];   // depth + 24 enc (+1 pad)
    __shared__ int s_g[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t npix = (int64_t)W * H;
    const int64_t pix0 = ((int64_t)blockIdx.x * 8 + warp) * 32;
    const int64_t pix = pix0 + lane;
    int g = -1;
    float depth = 0.f, dxn = 0.f, dyn = 0.f, dzn = 0.f;
    if (pix < npix) {
        const unsigned long long z = zbuf[pix];
        if (z != ZBUF_EMPTY) {
            g = (int)(uint32_t)(z & 0xffffffffull);
            depth = __uint_as_float((uint32_t)(z >> 32));
            // dir = normalize(mean - campos) (rasterizer_impl.cu:770-772)
            float vx = means3D[3 * (size_t)g] - campos[0];
            float vy = means3D[3 * (size_t)g + 1] - campos[1];
            float vz = means3D[3 * (size_t)g + 2] - campos[2];
            const float len = sqrtf(vx * vx + vy * vy + vz * vz);
            dxn = vx / len; dyn = vy / len; dzn = vz / len;
        }
        out_idx[pix] = g;
        out_depth[pix] = depth;
        out_color[3 * pix] = dxn; out_color[3 * pix + 1] = dyn; out_color[3 * pix + 2] = dzn;
    }
    s_g[warp][lane] = g;
    float* e = s_enc[warp][lane];
    e[0] = depth;
    if (g >= 0) {
        // positional_encoding_3d (rasterizer_impl.cu:26-42): index dim*8 + f*2 + {sin, cos}
#pragma unroll
        for (int dim = 0; dim < 3; ++dim) {
            const float v = dim == 0 ? dxn : (dim == 1 ? dyn : dzn);
#pragma unroll
            for (int f = 0; f < 4; ++f) {
                const float frequency = powf(2.0f, (float)f);
                const float arg = frequency * 3.14159265358979323846f * v;
                e[1 + dim * 8 + f * 2] = sinf(arg);
                e[1 + dim * 8 + f * 2 + 1] = cosf(arg);
            }
        }
    }
    __syncwarp();
    // four pixels per trip: their eight gathers are issued together before the first store needs its data
    // (the plain one-pixel loop sat on `long_scoreboard` at the store, profiles/r03d_ncu_stalls_idxmap_C3.txt)
    for (int q0 = 0; q0 < 32; q0 += 4) {
        if (pix0 + q0 >= npix) break;
        float v0[4], v1[4];            // channels lane and lane+32 of pixels q0..q0+3
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gq = s_g[warp][q0 + j];
            v0[j] = 0.f; v1[j] = 0.f;
            if (gq >= 0) {             // pixels past the end of the image carry g = -1
                v0[j] = lane < FEAT_FIRST_LEARNED ? s_enc[warp][q0 + j][lane] : __ldg(&feature_vector[(size_t)gq * NFEAT + lane]);
                v1[j] = __ldg(&feature_vector[(size_t)gq * NFEAT + 32 + lane]);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t pq = pix0 + q0 + j;
            if (pq < npix) {
                out_feature[(size_t)pq * NFEAT + lane] = v0[j];
                out_feature[(size_t)pq * NFEAT + 32 + lane] = v1[j];
            }
        }
    }
}

// K16 CopyFeature (rasterizer_impl.cu:820-847) with a correct (atomic) accumulation: the
// reference's plain `+=` races when several pixels hit the same point.
__global__ void __launch_bounds__(256)
idx_feature_backward_kernel(int P, int64_t npix, const int* __restrict__ out_idx,
                            const float* __restrict__ grad_feature, float* __restrict__ dL_dfeature) {
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int64_t pix0 = warp_global * 32;
    if (pix0 >= npix) return;
    const int64_t pix = pix0 + lane;
    const int g_mine = pix < npix ? out_idx[pix] : -1;
    for (int q = 0; q < 32; ++q) {
        const int g = __shfl_sync(FULL_MASK, g_mine, q);
        if (g < 0 || g >= P) continue;
        const int64_t pq = pix0 + q;
        if (lane >= FEAT_FIRST_LEARNED)
            atomicAdd(&dL_dfeature[(size_t)g * NFEAT + lane], grad_feature[(size_t)pq * NFEAT + lane]);
        atomicAdd(&dL_dfeature[(size_t)g * NFEAT + 32 + lane], grad_feature[(size_t)pq * NFEAT + 32 + lane]);
    }
}

size_t idxmap_workspace_size_impl(int P, int W, int H) {
    (void)P;
    return carve_idx(nullptr, W, H).bytes;
}

int idxmap_forward_impl(const B200sIdxmapArgs* a, cudaStream_t stream) {
    if (!a) return B200S_E_BADARG;
    if (a->P < 0 || a->W <= 0 || a->H <= 0) return B200S_E_BADARG;
    if (!a->out_idx || !a->out_color || !a->out_depth || !a->out_feature) return B200S_E_BADARG;
    const size_t npix = (size_t)a->W * a->H;
    if (a->num_rendered_dev) B200S_CUDA_OK(cudaMemsetAsync(a->num_rendered_dev, 0, 4, stream));
    if (a->P == 0) {
        // GetIdxMapCUDA returns its zero-initialised maps untouched when P == 0
        // (rasterize_points.cu:206-210,231): idx is 0 there, not -1.
        B200S_CUDA_OK(cudaMemsetAsync(a->out_idx, 0, 4 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_color, 0, 12 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_depth, 0, 4 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_feature, 0, 4 * (size_t)NFEAT * npix, stream));
        return 0;
    }
    if (!a->means3D || !a->viewmatrix || !a->projmatrix || !a->campos || !a->feature_vector || !a->radii || !a->ws)
        return B200S_E_BADARG;
    if (a->W > 32767 || a->H > 65535) return B200S_E_UNSUPPORTED;   // ushort2 pixels / 63-bit keys
    if (a->ws_bytes < idxmap_workspace_size_impl(a->P, a->W, a->H)) return B200S_E_WORKSPACE;
    IdxWS ws = carve_idx(a->ws, a->W, a->H);
    B200S_CUDA_OK(cudaMemsetAsync(ws.zbuf, 0xff, 8 * npix, stream));
    count_launch();
    idx_splat_kernel<<<(a->P + 255) / 256, 256, 0, stream>>>(a->P, a->W, a->H, a->means3D, a->viewmatrix,
                                                           a->projmatrix, a->radii, ws.zbuf, a->num_rendered_dev);
    B200S_AFTER_LAUNCH(stream, a->debug);
    const unsigned blocks = (unsigned)((npix + 255) / 256);
    count_launch();
    idx_resolve_kernel<<<blocks, 256, 0, stream>>>(a->P, a->W, a->H, ws.zbuf, a->means3D, a->campos,
                                                   a->feature_vector, a->out_idx, a->out_color, a->out_depth,
                                                   a->out_feature);
    B200S_AFTER_LAUNCH(stream, a->debug);
    return 0;
}

int idxmap_backward_impl(int P, int W, int H, const int32_t* out_idx, const float* grad_feature,
                         float* dL_dfeature_vector, cudaStream_t stream) {
    if (P < 0 || W <= 0 || H <= 0) return B200S_E_BADARG;
    if (P == 0) return 0;
    if (!out_idx || !grad_feature || !dL_dfeature_vector) return B200S_E_BADARG;
    const int64_t npix = (int64_t)W * H;
    B200S_CUDA_OK(cudaMemsetAsync(dL_dfeature_vector, 0, 4 * (size_t)NFEAT * P, stream));
    const unsigned blocks = (unsigned)((npix + 255) / 256);
    count_launch();
    idx_feature_backward_kernel<<<blocks, 256, 0, stream>>>(P, npix, out_idx, grad_feature, dL_dfeature_vector);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// idxmap forward for callers that have no workspace to give (b200s_idxmap_forward_static).
//
// The regular entry point needs an 8-byte-per-pixel z-buffer from the caller. Some hosts cannot
// produce one: a render loop living inside an application whose device memory is already handed
// out in full to another allocator (a pre-sized arena, a framework pool capped at a memory
// fraction), or a loop that must never hit cudaMalloc's device-wide synchronisation. For those the
// z-buffer below is part of the library image: every GPU that loads libb200splat.so carries these
// 16 MiB until the process exits, whether or not the entry point is ever called.
//
// The array is shared by every call on a device and keeps what the previous call left in it; each
// call therefore refills the part it uses with the "empty" pattern before splatting. Two calls that
// overlap on one GPU (different streams, different host threads) would splat into the same pixels
// and corrupt each other: callers must serialise them.
constexpr int64_t IDX_STATIC_MAX_PIXELS = B200S_IDXMAP_STATIC_MAX_PIXELS;
static __device__ __align__(256) unsigned long long g_idx_static_zbuf[IDX_STATIC_MAX_PIXELS];

// In host code `g_idx_static_zbuf` is only a handle for the runtime, not an address a kernel may
// dereference; the device address is asked for once per device (with that device current) and
// remembered. cudaGetSymbolAddress may initialise the module on that device, which is why the first
// call per device is documented as the set-up call.
constexpr int IDX_STATIC_MAX_DEVICES = 64;
static std::mutex g_idx_static_mu;
static unsigned long long* g_idx_static_ptr[IDX_STATIC_MAX_DEVICES] = {};

static int idx_static_zbuf_of_current_device(unsigned long long** out) {
    int dev = -1;
    B200S_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= IDX_STATIC_MAX_DEVICES) return B200S_E_UNSUPPORTED;
    std::lock_guard<std::mutex> lk(g_idx_static_mu);
    if (g_idx_static_ptr[dev] == nullptr) {
        void* p = nullptr;
        B200S_CUDA_OK(cudaGetSymbolAddress(&p, g_idx_static_zbuf));
        if (p == nullptr) return B200S_E_INTERNAL;
        g_idx_static_ptr[dev] = (unsigned long long*)p;
    }
    *out = g_idx_static_ptr[dev];
    return 0;
}

int idxmap_forward_static_impl(const B200sIdxmapArgs* a, cudaStream_t stream) {
    if (!a) return B200S_E_BADARG;
    if (a->P < 0 || a->W <= 0 || a->H <= 0) return B200S_E_BADARG;
    if (!a->out_idx || !a->out_color || !a->out_depth || !a->out_feature) return B200S_E_BADARG;
    const size_t npix = (size_t)a->W * a->H;
    // the size limit is checked before anything is enqueued, so a refused call leaves every output untouched
    if ((int64_t)npix > IDX_STATIC_MAX_PIXELS) return B200S_E_WORKSPACE;
    if (a->W > 32767 || a->H > 65535) return B200S_E_UNSUPPORTED;
    if (a->num_rendered_dev) B200S_CUDA_OK(cudaMemsetAsync(a->num_rendered_dev, 0, 4, stream));
    if (a->P == 0) {   // same zero-filled maps as the workspace variant; the z-buffer is not needed
        B200S_CUDA_OK(cudaMemsetAsync(a->out_idx, 0, 4 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_color, 0, 12 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_depth, 0, 4 * npix, stream));
        B200S_CUDA_OK(cudaMemsetAsync(a->out_feature, 0, 4 * (size_t)NFEAT * npix, stream));
        return 0;
    }
    if (!a->means3D || !a->viewmatrix || !a->projmatrix || !a->campos || !a->feature_vector || !a->radii)
        return B200S_E_BADARG;
    unsigned long long* zbuf = nullptr;
    const int rc = idx_static_zbuf_of_current_device(&zbuf);
    if (rc) return rc;
    // the previous call's winners are still in the array: every pixel of this view starts empty again
    B200S_CUDA_OK(cudaMemsetAsync(zbuf, 0xff, 8 * npix, stream));
    count_launch();
    idx_splat_kernel<<<(a->P + 255) / 256, 256, 0, stream>>>(a->P, a->W, a->H, a->means3D, a->viewmatrix,
                                                           a->projmatrix, a->radii, zbuf, a->num_rendered_dev);
    B200S_AFTER_LAUNCH(stream, a->debug);
    const unsigned blocks = (unsigned)((npix + 255) / 256);
    count_launch();
    idx_resolve_kernel<<<blocks, 256, 0, stream>>>(a->P, a->W, a->H, zbuf, a->means3D, a->campos, a->feature_vector,
                                                   a->out_idx, a->out_color, a->out_depth, a->out_feature);
    B200S_AFTER_LAUNCH(stream, a->debug);
    return 0;
}

}  // namespace b200s
