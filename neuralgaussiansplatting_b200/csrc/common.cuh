// Shared definitions for the b200splat kernels (sm_100a only).
#pragma once

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/b200splat.h"

#ifndef __CUDA_ARCH__
#define B200S_HOST 1
#endif

namespace b200s {

constexpr int TILE = 16;                 // B200S_BLOCK_X == B200S_BLOCK_Y
constexpr int TILE_PIXELS = TILE * TILE;
constexpr uint32_t FULL_MASK = 0xffffffffu;

// ---------------------------------------------------------------------------------
// Per-Gaussian record written by the forward preprocess for VISIBLE Gaussians only
// (radii > 0) and gathered by binning and both blend kernels. 48 B, 16-B aligned, so a
// gather touches exactly two 32-B sectors (the reference gathers from three separate
// arrays: means2D, conic_opacity, rgb — rasterizer_impl.cu:161-164).
//   q0 = { x, y, conic.a, conic.b }          pixel-space mean, inverse 2-D covariance
//   q1 = { conic.c, opacity, r, g }
//   q2 = { b, depth, rect_min (x | y<<16), rect_wh (w | h<<16) }   tile rectangle
// ---------------------------------------------------------------------------------
struct __align__(16) SplatRec {
    float4 q0, q1, q2;
};

// ---------------------------------------------------------------------------------
// Workspace layouts. Every sub-array starts 256-B aligned.
// ---------------------------------------------------------------------------------
constexpr size_t WS_ALIGN = 256;
__host__ __device__ inline size_t align_up(size_t x, size_t a = WS_ALIGN) {
    return (x + a - 1) / a * a;
}

// Threads per block of the look-back scans. All blocks of these small kernels are resident at once, so the look-back
// of block i walks ~i/32 windows of predecessors before it meets an inclusive word: fewer, larger blocks shorten the
// chain (B200S_SCAN_THREADS / B200S_SS_THREADS: tuning knobs, A/B'd with build variants). Measured on B200
// (profiles/r02h_scan_variants.log): sorted_scan 256 -> 1024 threads: expand stage 0.111 -> 0.098 ms at C3, 0.206 ->
// 0.198 at C4; scan_compact 256 -> 512 / 1024: 0.037 -> 0.040 / 0.043 ms (slower: it is not chain-bound), so it stays.
#ifndef B200S_SCAN_THREADS
#define B200S_SCAN_THREADS 256
#endif
#ifndef B200S_SS_THREADS
#define B200S_SS_THREADS 1024
#endif
constexpr int SCAN_THREADS = B200S_SCAN_THREADS;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// Radix sort geometry (see binning.cu).
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int SORT_MAX_PASSES = 4;    // onesweep passes over the tile bits (<= 32 bits of tile id)
constexpr int DEPTH_PASSES = 4;       // onesweep passes over the 32 depth bits

// Expansion geometry: one block turns EXP_TILE depth-sorted Gaussians into their instances.
constexpr int EXP_THREADS = 256;
constexpr int SS_THREADS = B200S_SS_THREADS;   // sorted_scan_kernel (the scan in front of the expansion)
constexpr int EXP_ITEMS = 4;
constexpr int EXP_TILE = SS_THREADS * EXP_ITEMS;
constexpr int EXP_OUT_TILE = EXP_THREADS * 8;   // instances emitted per expansion block

struct GeomWS {
    B200sHeader* header;      // 64 B
    uint32_t* counters;       // 64 x u32 device tickets (scan ticket, sort tickets, ...)
    SplatRec* rec;            // [P]
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P] inclusive prefix sum of tiles_touched (Gaussian-index order)
    uint8_t* clamped;         // [P] bit c set when SH colour channel c was clamped at 0
    uint32_t* depth_bits;     // [P] view-space depth of visible Gaussians (float bits)
    uint32_t* vis_keys[2];    // [P] depth bits of the compacted visible Gaussians (ping-pong)
    uint32_t* vis_ids[2];     // [P] their Gaussian ids; after the depth sort: front-to-back order
    unsigned long long* scan_status;   // [ceil(P / SCAN_TILE) + 1] decoupled look-back words
    uint32_t* depth_hist;     // [DEPTH_PASSES][SORT_RADIX]
    uint32_t* depth_lookback; // [DEPTH_PASSES][ceil(P / SORT_TILE)][SORT_RADIX]
    uint32_t* exp_status;     // [ceil(P / EXP_TILE) + 1] look-back words of the expansion
    size_t zero_off, zero_bytes;   // scan_status..exp_status are contiguous: cleared every call
    int64_t depth_tiles;
    size_t bytes;
};

inline GeomWS carve_geom(void* base, int64_t P) {
    GeomWS g;
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t n) { char* p = b + off; off = align_up(off + n); return p; };
    g.header = (B200sHeader*)take(sizeof(B200sHeader));
    g.counters = (uint32_t*)take(64 * sizeof(uint32_t));
    g.rec = (SplatRec*)take(sizeof(SplatRec) * (size_t)P);
    g.tiles_touched = (uint32_t*)take(4 * (size_t)P);
    g.point_offsets = (uint32_t*)take(4 * (size_t)P);
    g.clamped = (uint8_t*)take((size_t)P);
    g.depth_bits = (uint32_t*)take(4 * (size_t)P);
    for (int i = 0; i < 2; ++i) {
        g.vis_keys[i] = (uint32_t*)take(4 * (size_t)P);
        g.vis_ids[i] = (uint32_t*)take(4 * (size_t)P);
    }
    g.depth_tiles = (P + SORT_TILE - 1) / SORT_TILE;
    g.zero_off = off;
    g.scan_status = (unsigned long long*)take(8 * ((size_t)(P + SCAN_TILE - 1) / SCAN_TILE + 1));
    g.depth_hist = (uint32_t*)take(4 * (size_t)DEPTH_PASSES * SORT_RADIX);
    g.depth_lookback = (uint32_t*)take(4 * (size_t)DEPTH_PASSES * g.depth_tiles * SORT_RADIX);
    g.exp_status = (uint32_t*)take(4 * ((size_t)(P + EXP_TILE - 1) / EXP_TILE + 1));
    g.zero_bytes = off - g.zero_off;
    g.bytes = off;
    return g;
}

struct ImgWS {
    float* final_T;        // [W*H]
    uint32_t* n_contrib;   // [W*H]
    uint2* ranges;         // [tiles]
    uint32_t* tile_last;   // [tiles] max n_contrib over the tile's pixels
    uint32_t* order_fwd;   // [tiles] tile ids, longest list first (launch order of the forward blend)
    uint32_t* order_bwd;   // [tiles] same for the backward (list length capped by tile_last)
    size_t bytes;
};

inline ImgWS carve_img(void* base, int W, int H) {
    ImgWS s;
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t n) { char* p = b + off; off = align_up(off + n); return p; };
    size_t px = (size_t)W * H;
    size_t tiles = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    s.final_T = (float*)take(4 * px);
    s.n_contrib = (uint32_t*)take(4 * px);
    s.ranges = (uint2*)take(8 * tiles);
    s.tile_last = (uint32_t*)take(4 * tiles);
    s.order_fwd = (uint32_t*)take(4 * tiles);
    s.order_bwd = (uint32_t*)take(4 * tiles);
    s.bytes = off;
    return s;
}

struct BinWS {
    uint32_t* keys[2];     // [capacity] tile id of every instance (ping-pong)
    uint32_t* vals[2];     // [capacity] Gaussian id of every instance
    uint32_t* hist;        // [SORT_MAX_PASSES][SORT_RADIX] digit histograms -> exclusive bases
    uint32_t* lookback;    // [SORT_MAX_PASSES][sort_tiles][SORT_RADIX]
    uint32_t* block_first; // [capacity / EXP_OUT_TILE + 2] first Gaussian of every expansion block
    size_t zero_bytes;     // hist+lookback are contiguous: bytes to clear before a sort
    int64_t sort_tiles;
    size_t bytes;
};

inline BinWS carve_bin(void* base, int64_t capacity) {
    BinWS s;
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t n) { char* p = b + off; off = align_up(off + n); return p; };
    if (capacity < 1) capacity = 1;
    s.sort_tiles = (capacity + SORT_TILE - 1) / SORT_TILE;
    s.keys[0] = (uint32_t*)take(4 * (size_t)capacity);
    s.keys[1] = (uint32_t*)take(4 * (size_t)capacity);
    s.vals[0] = (uint32_t*)take(4 * (size_t)capacity);
    s.vals[1] = (uint32_t*)take(4 * (size_t)capacity);
    size_t z0 = off;
    s.hist = (uint32_t*)take(4 * (size_t)SORT_MAX_PASSES * SORT_RADIX);
    s.lookback = (uint32_t*)take(4 * (size_t)SORT_MAX_PASSES * s.sort_tiles * SORT_RADIX);
    s.zero_bytes = off - z0;
    s.block_first = (uint32_t*)take(4 * ((size_t)capacity / EXP_OUT_TILE + 2));
    s.bytes = off;
    return s;
}

// Per-Gaussian accumulators of the backward blend: 12 floats / Gaussian
//   [0..2] dL/dRGB  [3..4] dL/dmean2D (pixel units, before the 0.5*W / 0.5*H factor is
//   applied... see blend_bwd.cu)  [5..7] dL/dconic (a, b, c)  [8] dL/dopacity  [9..11] pad
constexpr int ACC_STRIDE = 12;

struct BwdWS {
    float* accum;          // [P][ACC_STRIDE]   (atomic mode)
    float* partial;        // [R][ACC_STRIDE]   (deterministic mode), slot = unsorted instance
    size_t bytes;
};

inline BwdWS carve_bwd(void* base, int64_t P, int64_t R) {
    BwdWS s;
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t n) { char* p = b + off; off = align_up(off + n); return p; };
    s.accum = (float*)take(4 * (size_t)ACC_STRIDE * (size_t)(P > 0 ? P : 1));
    s.partial = (float*)take(4 * (size_t)ACC_STRIDE * (size_t)(R > 0 ? R : 1));
    s.bytes = off;
    return s;
}

// Device-side counters (GeomWS::counters) -------------------------------------------
enum { CTR_SCAN_TICKET = 0, CTR_SORT_TICKET0 = 1 /* ..+SORT_MAX_PASSES */, CTR_DEPTH_TICKET0 = 8 /* ..+DEPTH_PASSES */,
       CTR_EXPAND_TICKET = 15, CTR_VISIBLE = 16 };

// Device error bits (B200sHeader::error)
enum { ERR_LOOKBACK_TIMEOUT = 1u, ERR_PREFILTERED = 2u };

#define B200S_CUDA_OK(expr)                                   \
    do {                                                      \
        cudaError_t _e = (expr);                              \
        if (_e != cudaSuccess) return (int)_e;                \
    } while (0)

// After a kernel launch: catch launch errors; in debug mode also synchronise
// (reference CHECK_CUDA, auxiliary.h:166-174).
#define B200S_AFTER_LAUNCH(stream, debug)                     \
    do {                                                      \
        cudaError_t _e = cudaGetLastError();                  \
        if (_e != cudaSuccess) return (int)_e;                \
        if (debug) {                                          \
            _e = cudaStreamSynchronize(stream);               \
            if (_e != cudaSuccess) return (int)_e;            \
        }                                                     \
    } while (0)

// cudaFuncSetAttribute applies to the CURRENT device only, so "done once" has to be tracked per device (a
// process may drive several GPUs, and several host threads may race here): one atomic flag per (call site,
// device). Returns a cudaError_t as int.
#define B200S_SET_FUNC_ATTR_ONCE(kernel, attr, value)                                            \
    do {                                                                                         \
        static std::atomic<unsigned char> _done[64];                                             \
        int _dev = 0;                                                                            \
        B200S_CUDA_OK(cudaGetDevice(&_dev));                                                     \
        if (_dev < 0 || _dev >= 64 || !_done[_dev].load(std::memory_order_acquire)) {            \
            B200S_CUDA_OK(cudaFuncSetAttribute(kernel, attr, value));                            \
            if (_dev >= 0 && _dev < 64) _done[_dev].store(1, std::memory_order_release);         \
        }                                                                                        \
    } while (0)

#ifdef __CUDACC__
// Checked build (python -m neuralgaussiansplatting_b200.build --tag checked -DB200S_CHECKED=1): every index that is
// derived from device data (list entries, tile ids, scatter destinations, compaction slots) is range-checked and the
// kernel traps with a message on the first violation. Stands in for compute-sanitizer, which this pool refuses to
// run (profiles/r02e_sanitizer_unavailable.txt). The product build compiles the checks out.
#if defined(B200S_CHECKED) && B200S_CHECKED
#include <cstdio>
#define B200S_CHECK(cond)                                                                                   \
    do {                                                                                                    \
        if (!(cond)) {                                                                                      \
            printf("b200s check failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__,        \
                   (int)blockIdx.x, (int)threadIdx.x);                                                      \
            __trap();                                                                                       \
        }                                                                                                   \
    } while (0)
#else
#define B200S_CHECK(cond) ((void)0)
#endif
// ---------------------------------------------------------------------------------
// Bit-exact geometry. The integer outputs (radii, tiles_touched, depth bits, ranges,
// n_contrib) are functions of fp32 expression trees whose rounding is fixed by how
// nvcc contracted the reference's source into FMAs. The operation order below mirrors
// the SASS of the reference kernels built for sm_100 (decoded instruction by
// instruction; see DESIGN.md "Bit-exactness"). Explicit _rn intrinsics are never
// re-associated or contracted by the compiler.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// a0*b0 + a1*b1 + a2*b2 as the reference's GLM mat3 products compile:
// fma(a2, b2, fma(a0, b0, a1*b1))
__device__ __forceinline__ float dot3_ref(float a0, float b0, float a1, float b1, float a2, float b2) {
    return ffma(a2, b2, ffma(a0, b0, fmul(a1, b1)));
}

// m[r]*x + m[4+r]*y + m[8+r]*z + m[12+r]  (auxiliary.h:58-76) compiles to
// fma(z, m8, fma(x, m0, y*m4)) + m12
__device__ __forceinline__ float affine_row_ref(const float* __restrict__ m, int r, float x, float y, float z) {
    return fadd(ffma(z, m[8 + r], ffma(x, m[r], fmul(y, m[4 + r]))), m[12 + r]);
}

// ndc2Pix (auxiliary.h:41-44) silently evaluates in double: ((v + 1.0) * S - 1.0) * 0.5,
// compiled as DADD, DFMA(.., S, -1), DMUL 0.5, F2F.F32.F64.
__device__ __forceinline__ float ndc2pix_ref(float v, int S) {
    double t = __dadd_rn((double)v, 1.0);
    t = __fma_rn(t, (double)S, -1.0);
    t = __dmul_rn(t, 0.5);
    return __double2float_rn(t);
}

struct TileRect { int x0, y0, x1, y1; };

// getRect (auxiliary.h:46-56): float arithmetic, division by 16 as multiplication by
// 0.0625 (exact), truncation, signed max with 0 then unsigned min with the grid size.
__device__ __forceinline__ TileRect get_rect_ref(float px, float py, int radius, int gx, int gy) {
    const float r = (float)radius;
    TileRect t;
    int a = __float2int_rz(fmul(fsub(px, r), 0.0625f));
    int b = __float2int_rz(fmul(fsub(py, r), 0.0625f));
    int c = __float2int_rz(fmul(fadd(fadd(fadd(px, r), 16.0f), -1.0f), 0.0625f));
    int d = __float2int_rz(fmul(fadd(fadd(fadd(py, r), 16.0f), -1.0f), 0.0625f));
    t.x0 = (int)min((unsigned)max(a, 0), (unsigned)gx);
    t.y0 = (int)min((unsigned)max(b, 0), (unsigned)gy);
    t.x1 = (int)min((unsigned)max(c, 0), (unsigned)gx);
    t.y1 = (int)min((unsigned)max(d, 0), (unsigned)gy);
    return t;
}

// computeCov3D (forward.cu:118-152): Sigma = (S R)^T (S R) for the UN-normalised
// quaternion (r,x,y,z) = (q.x,q.y,q.z,q.w). Multiplications by the zero off-diagonals of
// S contribute exact zeros, so M[c][k] = s_k * R[c][k] with a single rounding.
__device__ __forceinline__ void cov3d_ref(float sx, float sy, float sz, float mod,
                                          float r, float x, float y, float z, float* cov) {
    const float S0 = fmul(sx, mod), S1 = fmul(sy, mod), S2 = fmul(sz, mod);
    const float xz = fmul(x, z), rx = fmul(r, x), rz = fmul(r, z);
    const float yy = fmul(y, y), zz = fmul(z, z);
    const float xz_p_ry = ffma(r, y, xz);
    const float xz_m_ry = ffma(-r, y, xz);
    const float yz_m_rx = ffma(y, z, -rx);
    const float yz_p_rx = ffma(y, z, rx);
    const float xy_m_rz = ffma(x, y, -rz);
    const float xy_p_rz = ffma(x, y, rz);
    const float xx_p_yy = ffma(x, x, yy);
    const float yy_p_zz = fadd(yy, zz);
    const float xx_p_zz = ffma(x, x, zz);
    // GLM column-major: Rcr = R[c][r]
    const float R00 = fsub(1.0f, fadd(yy_p_zz, yy_p_zz));
    const float R01 = fadd(xy_m_rz, xy_m_rz);
    const float R02 = fadd(xz_p_ry, xz_p_ry);
    const float R10 = fadd(xy_p_rz, xy_p_rz);
    const float R11 = fsub(1.0f, fadd(xx_p_zz, xx_p_zz));
    const float R12 = fadd(yz_m_rx, yz_m_rx);
    const float R20 = fadd(xz_m_ry, xz_m_ry);
    const float R21 = fadd(yz_p_rx, yz_p_rx);
    const float R22 = fsub(1.0f, fadd(xx_p_yy, xx_p_yy));
    const float M00 = fmul(S0, R00), M01 = fmul(S1, R01), M02 = fmul(S2, R02);
    const float M10 = fmul(S0, R10), M11 = fmul(S1, R11), M12 = fmul(S2, R12);
    const float M20 = fmul(S0, R20), M21 = fmul(S1, R21), M22 = fmul(S2, R22);
    cov[0] = dot3_ref(M00, M00, M01, M01, M02, M02);
    cov[1] = dot3_ref(M00, M10, M01, M11, M02, M12);
    cov[2] = dot3_ref(M00, M20, M01, M21, M02, M22);
    cov[3] = dot3_ref(M10, M10, M11, M11, M12, M12);
    cov[4] = dot3_ref(M10, M20, M11, M21, M12, M22);
    cov[5] = dot3_ref(M20, M20, M21, M21, M22, M22);
}

// computeCov2D (forward.cu:74-113). Returns (cov.x, cov.y, cov.z) with the +0.3 low-pass.
// Also returns the T = W*J entries for the backward.
struct Cov2D { float a, b, c; float T00, T01, T02, T10, T11, T12; float tx, ty, tz; float txtz, tytz; };

__device__ __forceinline__ Cov2D cov2d_ref(float px, float py, float pz, float focal_x, float focal_y,
                                           float tan_fovx, float tan_fovy, const float* cov3D,
                                           const float* __restrict__ view) {
    Cov2D o;
    const float tx = affine_row_ref(view, 0, px, py, pz);
    const float ty = affine_row_ref(view, 1, px, py, pz);
    const float tz = affine_row_ref(view, 2, px, py, pz);
    const float limx = fmul(tan_fovx, 1.3f);
    const float limy = fmul(tan_fovy, 1.3f);
    const float txtz = __fdiv_rn(tx, tz);
    const float tytz = __fdiv_rn(ty, tz);
    const float cx = fminf(fmaxf(txtz, -limx), limx);
    const float cy = fminf(fmaxf(tytz, -limy), limy);
    o.txtz = txtz; o.tytz = tytz;
    o.tx = fmul(cx, tz); o.ty = fmul(cy, tz); o.tz = tz;
    const float tz2 = fmul(tz, tz);
    const float J00 = __fdiv_rn(focal_x, tz);
    const float J02 = __fdiv_rn(fmul(-o.tx, focal_x), tz2);
    const float J11 = __fdiv_rn(focal_y, tz);
    const float J12 = __fdiv_rn(fmul(-o.ty, focal_y), tz2);
    // T = W * J with W = view rotation (columns (v0,v4,v8),(v1,v5,v9),(v2,v6,v10)); the
    // products with J's zeros are exact zeros.
    o.T00 = ffma(view[2], J02, fmul(view[0], J00));
    o.T01 = ffma(view[6], J02, fmul(view[4], J00));
    o.T02 = ffma(view[10], J02, fmul(view[8], J00));
    o.T10 = ffma(view[2], J12, fmul(view[1], J11));
    o.T11 = ffma(view[6], J12, fmul(view[5], J11));
    o.T12 = ffma(view[10], J12, fmul(view[9], J11));
    const float c0 = cov3D[0], c1 = cov3D[1], c2 = cov3D[2], c3 = cov3D[3], c4 = cov3D[4], c5 = cov3D[5];
    // X = T^T * Vrk^T
    const float X00 = dot3_ref(o.T00, c0, o.T01, c1, o.T02, c2);
    const float X01 = dot3_ref(o.T10, c0, o.T11, c1, o.T12, c2);
    const float X10 = dot3_ref(o.T00, c1, o.T01, c3, o.T02, c4);
    const float X11 = dot3_ref(o.T10, c1, o.T11, c3, o.T12, c4);
    const float X20 = dot3_ref(o.T00, c2, o.T01, c4, o.T02, c5);
    const float X21 = dot3_ref(o.T10, c2, o.T11, c4, o.T12, c5);
    // cov = X * T
    o.a = fadd(dot3_ref(o.T00, X00, o.T01, X10, o.T02, X20), 0.3f);
    o.b = dot3_ref(o.T00, X01, o.T01, X11, o.T02, X21);
    o.c = fadd(dot3_ref(o.T10, X01, o.T11, X11, o.T12, X21), 0.3f);
    return o;
}

// Blend falloff exactly as renderCUDA compiles (forward.cu:332-335):
//   power = fma(fma(dx, dx*ca, dy*(dy*cc)), -0.5, -(dy*(dx*cb)))
__device__ __forceinline__ float power_ref(float dx, float dy, float ca, float cb, float cc) {
    const float s = ffma(dx, fmul(dx, ca), fmul(dy, fmul(dy, cc)));
    return ffma(s, -0.5f, -fmul(dy, fmul(dx, cb)));
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Lanes of the warp holding the same 9-bit digit (8-bit radix digit or the padding sentinel 256).
// match.any by default; the nine-ballot form (B200S_USE_MATCH=0) measured ~5 % slower on B200
// (r01: C3 sort stage 0.500 ms vs 0.475 ms).
#ifndef B200S_USE_MATCH
#define B200S_USE_MATCH 1
#endif
__device__ __forceinline__ uint32_t digit_peers(uint32_t d) {
#if B200S_USE_MATCH
    return __match_any_sync(FULL_MASK, d);
#else
    uint32_t peers = FULL_MASK;
#pragma unroll
    for (int b = 0; b < 9; ++b) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(FULL_MASK, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
#endif
}

// Ballot form for the 8-bit digits + sentinel bit (cost independent of the digit distribution).
__device__ __forceinline__ uint32_t digit_peers_ballot(uint32_t d) {
    uint32_t peers = FULL_MASK;
#pragma unroll
    for (int b = 0; b < 9; ++b) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(FULL_MASK, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
#endif  // __CUDACC__

}  // namespace b200s
