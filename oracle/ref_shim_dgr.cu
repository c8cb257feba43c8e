// TEST INFRASTRUCTURE — not part of the product path.
//
// extern "C" shim around the UNMODIFIED reference rasterizer
// (submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu, compiled where
// they lie under /root/reference by oracle/build_ref.sh into oracle/_ref/).
// It replaces the reference's torch binding (rasterize_points.cu:35-196) with
// a raw-pointer interface so tests/ and `bench.py --impl reference` can drive
// the reference kernels on the same device buffers as the product and read back
// the private state blobs (GeometryState / BinningState / ImageState,
// rasterizer_impl.cu:155-194) for bit-exact comparison.
//
// Only tests/, __graft_entry__.smoke() and bench.py's reference/cpu_baseline
// legs may load the resulting library.

#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
#include "rasterizer.h"
#include "rasterizer_impl.h"

namespace {

struct Blob {
    char* ptr = nullptr;
    size_t cap = 0;
    size_t used = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
            cap = want;
        }
        used = n;
        return ptr;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = used = 0; }
};

struct RefCtx {
    Blob geom, binning, img;
    int P = 0, W = 0, H = 0, R = 0;
};

}  // namespace

extern "C" {

void* ref_dgr_create() { return new RefCtx(); }

void ref_dgr_destroy(void* h) {
    RefCtx* c = (RefCtx*)h;
    if (!c) return;
    c->geom.release(); c->binning.release(); c->img.release();
    delete c;
}

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-115) minus torch.
// out_color [3,H,W] and radii [P] must be zero-filled by the caller, as the
// reference binding does with torch::full (rasterize_points.cu:68-69).
int ref_dgr_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
                    const float* means3D, const float* shs, const float* colors_precomp,
                    const float* opacities, const float* scales, float scale_modifier,
                    const float* rotations, const float* cov3D_precomp,
                    const float* viewmatrix, const float* projmatrix, const float* campos,
                    float tan_fovx, float tan_fovy, int prefiltered,
                    float* out_color, int* radii, int debug)
{
    RefCtx* c = (RefCtx*)h;
    c->P = P; c->W = W; c->H = H; c->R = 0;
    if (P == 0) return 0;
    std::function<char*(size_t)> g = [c](size_t n) { return c->geom.get(n); };
    std::function<char*(size_t)> b = [c](size_t n) { return c->binning.get(n); };
    std::function<char*(size_t)> i = [c](size_t n) { return c->img.get(n); };
    int R = -1;
    try {
        R = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, bg, W, H, means3D, shs,
            colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp,
            viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, prefiltered != 0,
            out_color, radii, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_dgr_forward] %s\n", e.what());
        return -1;
    }
    c->R = R;
    return R;
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:117-196). All nine
// gradient buffers must be zero-filled by the caller (torch::zeros, :151-159).
int ref_dgr_backward(void* h, int P, int D, int M, int R, const float* bg, int W, int H,
                     const float* means3D, const float* shs, const float* colors_precomp,
                     const float* scales, float scale_modifier, const float* rotations,
                     const float* cov3D_precomp, const float* viewmatrix,
                     const float* projmatrix, const float* campos,
                     float tan_fovx, float tan_fovy, const int* radii,
                     const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                     float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                     int debug)
{
    RefCtx* c = (RefCtx*)h;
    if (P == 0) return 0;
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs,
            colors_precomp, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
            projmatrix, campos, tan_fovx, tan_fovy, radii, c->geom.ptr, c->binning.ptr,
            c->img.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
            dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_dgr_backward] %s\n", e.what());
        return -1;
    }
    return 0;
}

int ref_dgr_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix,
                         unsigned char* present)
{
    if (P == 0) return 0;
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, (bool*)present);
    return 0;
}

// Copies one private state array of the last forward to a host buffer.
// which: 0 depths f32[P]  1 clamped u8[3P]  2 means2D f32[2P]  3 cov3D f32[6P]
//        4 conic_opacity f32[4P]  5 rgb f32[3P]  6 tiles_touched u32[P]
//        7 point_offsets u32[P]  8 point_list u32[R]  9 point_list_keys u64[R]
//        10 ranges u32[2*tiles... stored as uint2 per PIXEL-count entry; first T used]
//        11 n_contrib u32[W*H]  12 accum_alpha f32[W*H]
//        13 point_list_keys_unsorted u64[R]  14 point_list_unsorted u32[R]
// Returns bytes copied, or -1.
long long ref_dgr_fetch(void* h, int which, void* host_dst, long long max_bytes)
{
    RefCtx* c = (RefCtx*)h;
    if (!c || !c->geom.ptr) return -1;
    char* gp = c->geom.ptr; char* bp = c->binning.ptr; char* ip = c->img.ptr;
    size_t P = c->P, R = c->R, NP = (size_t)c->W * c->H;
    CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(gp, P);
    CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(ip, NP);
    CudaRasterizer::BinningState bs;
    if (bp) bs = CudaRasterizer::BinningState::fromChunk(bp, R);
    const void* src = nullptr; size_t bytes = 0;
    switch (which) {
        case 0: src = gs.depths; bytes = 4 * P; break;
        case 1: src = gs.clamped; bytes = 3 * P; break;
        case 2: src = gs.means2D; bytes = 8 * P; break;
        case 3: src = gs.cov3D; bytes = 24 * P; break;
        case 4: src = gs.conic_opacity; bytes = 16 * P; break;
        case 5: src = gs.rgb; bytes = 12 * P; break;
        case 6: src = gs.tiles_touched; bytes = 4 * P; break;
        case 7: src = gs.point_offsets; bytes = 4 * P; break;
        case 8: src = bp ? bs.point_list : nullptr; bytes = 4 * R; break;
        case 9: src = bp ? bs.point_list_keys : nullptr; bytes = 8 * R; break;
        case 10: src = is.ranges; bytes = 8 * (size_t)((c->W + 15) / 16) * ((c->H + 15) / 16); break;
        case 11: src = is.n_contrib; bytes = 4 * NP; break;
        case 12: src = is.accum_alpha; bytes = 4 * NP; break;
        case 13: src = bp ? bs.point_list_keys_unsorted : nullptr; bytes = 8 * R; break;
        case 14: src = bp ? bs.point_list_unsorted : nullptr; bytes = 4 * R; break;
        default: return -1;
    }
    if (bytes == 0) return 0;
    if (!src || (long long)bytes > max_bytes) return -1;
    if (cudaMemcpy(host_dst, src, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (long long)bytes;
}

}  // extern "C"
