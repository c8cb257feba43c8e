// TEST INFRASTRUCTURE — not part of the product path.
//
// extern "C" shim around the UNMODIFIED reference `rasterizer2`
// (submodules/rasterizer2/cuda_rasterizer/*.cu, compiled where they lie under
// /root/reference by oracle/build_ref.sh into oracle/_ref/). Replaces the torch
// binding (rasterizer2/rasterize_points.cu:26-295) with raw pointers: the
// stripped alpha path (forward/backward, no precomp inputs) and the idxmap
// pipeline (getidxmap / getidxmap_backward).

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
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
            cap = want;
        }
        return ptr;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = 0; }
};

struct RefCtx {
    Blob geom, binning, img;      // alpha path
    Blob idx, binimg, pixel;      // idxmap path
    int P = 0, W = 0, H = 0, R = 0, R2 = 0;
};

}  // namespace

extern "C" {

void* ref_r2_create() { return new RefCtx(); }

void ref_r2_destroy(void* h) {
    RefCtx* c = (RefCtx*)h;
    if (!c) return;
    c->geom.release(); c->binning.release(); c->img.release();
    c->idx.release(); c->binimg.release(); c->pixel.release();
    delete c;
}

// rasterizer2/rasterize_points.cu:26-100 (17-arg forward, no precomp inputs).
int ref_r2_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
                   const float* means3D, const float* shs, const float* opacities,
                   const float* scales, float scale_modifier, const float* rotations,
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
            opacities, scales, scale_modifier, rotations, viewmatrix, projmatrix, campos,
            tan_fovx, tan_fovy, prefiltered != 0, out_color, radii, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_r2_forward] %s\n", e.what());
        return -1;
    }
    c->R = R;
    return R;
}

// rasterizer2/rasterize_points.cu:102-180.
int ref_r2_backward(void* h, int P, int D, int M, int R, const float* bg, int W, int H,
                    const float* means3D, const float* shs, const float* scales,
                    float scale_modifier, const float* rotations, const float* viewmatrix,
                    const float* projmatrix, const float* campos, float tan_fovx,
                    float tan_fovy, const int* radii, const float* dL_dpix,
                    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                    float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                    float* dL_dscale, float* dL_drot, int debug)
{
    RefCtx* c = (RefCtx*)h;
    if (P == 0) return 0;
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, scales,
            scale_modifier, rotations, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
            radii, c->geom.ptr, c->binning.ptr, c->img.ptr, dL_dpix, dL_dmean2D, dL_dconic,
            dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
            debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_r2_backward] %s\n", e.what());
        return -1;
    }
    return 0;
}

// rasterizer2/rasterize_points.cu:183-260. The four output maps and radii must be
// zero-filled by the caller (torch::full(..., 0), :206-210).
int ref_r2_getidxmap(void* h, int P, int M, int W, int H, const float* means3D,
                     const float* normal, const float* scales, const float* rotations,
                     const float* viewmatrix, const float* projmatrix, float tan_fovx,
                     float tan_fovy, const float* feature_vector, const float* sh,
                     const float* campos, int* out_idx, float* out_color, float* out_depth,
                     float* out_feature, int* radii, int debug)
{
    RefCtx* c = (RefCtx*)h;
    c->P = P; c->W = W; c->H = H; c->R2 = 0;
    if (P == 0) return 0;
    std::function<char*(size_t)> a = [c](size_t n) { return c->idx.get(n); };
    std::function<char*(size_t)> b = [c](size_t n) { return c->binimg.get(n); };
    std::function<char*(size_t)> p = [c](size_t n) { return c->pixel.get(n); };
    int R = -1;
    try {
        R = CudaRasterizer::Rasterizer::getidxmap(a, b, p, P, M, W, H, means3D, normal,
            scales, rotations, viewmatrix, projmatrix, tan_fovx, tan_fovy, feature_vector,
            sh, campos, out_idx, out_color, out_depth, out_feature, radii, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "[ref_r2_getidxmap] %s\n", e.what());
        return -1;
    }
    c->R2 = R;
    return R;
}

// rasterizer2/rasterize_points.cu:262-295. dL_dfeature_vector [P,64] zero-filled
// by the caller (torch::zeros, :280).
int ref_r2_getidxmap_backward(int P, int W, int H, const float* means3D,
                              const float* campos, int* out_idx, float* grad_feature,
                              float* dL_dfeature_vector, int debug)
{
    if (P == 0) return 0;
    CudaRasterizer::Rasterizer::getidxmap_backward(P, W, H, means3D, campos, out_idx,
                                                   grad_feature, dL_dfeature_vector,
                                                   debug != 0);
    return 0;
}

// which: 0 depths f32[P]  1 pixels u16[2P]  2 img_xy f32[2P]  3 radii_float f32[P]
//        4 touched u32[P]  5 point_offsets u32[P]  6 point_list u32[R2]
//        7 point_list_keys u64[R2]  8 ranges u32[2*W*H]
long long ref_r2_fetch_idx(void* h, int which, void* host_dst, long long max_bytes)
{
    RefCtx* c = (RefCtx*)h;
    if (!c || !c->idx.ptr) return -1;
    char* ap = c->idx.ptr; char* bp = c->binimg.ptr; char* pp = c->pixel.ptr;
    size_t P = c->P, R = c->R2, NP = (size_t)c->W * c->H;
    CudaRasterizer::Idx2Img s = CudaRasterizer::Idx2Img::fromChunk(ap, P);
    CudaRasterizer::PixelState ps = CudaRasterizer::PixelState::fromChunk(pp, NP);
    CudaRasterizer::BinningImg bs;
    if (bp) bs = CudaRasterizer::BinningImg::fromChunk(bp, R);
    const void* src = nullptr; size_t bytes = 0;
    switch (which) {
        case 0: src = s.depths; bytes = 4 * P; break;
        case 1: src = s.pixels; bytes = 4 * P; break;
        case 2: src = s.img_xy; bytes = 8 * P; break;
        case 3: src = s.radii_float; bytes = 4 * P; break;
        case 4: src = s.touched; bytes = 4 * P; break;
        case 5: src = s.point_offsets; bytes = 4 * P; break;
        case 6: src = bp ? bs.point_list : nullptr; bytes = 4 * R; break;
        case 7: src = bp ? bs.point_list_keys : nullptr; bytes = 8 * R; break;
        case 8: src = ps.ranges; bytes = 8 * NP; break;
        default: return -1;
    }
    if (bytes == 0) return 0;
    if (!src || (long long)bytes > max_bytes) return -1;
    if (cudaMemcpy(host_dst, src, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (long long)bytes;
}

}  // extern "C"
