// Pieces shared by the forward and backward blend kernels (blend_fwd.cu, blend_bwd.cu):
// cp.async staging of SplatRec gathers, the conservative "can this Gaussian reach alpha >= 1/255
// inside this pixel rectangle" test, and the pixel-block geometry of a warp.
//
// Hierarchical culling. The reference walks every entry of a tile's list with all 256 pixels
// (forward.cu:325-351). Any (pixel, Gaussian) pair with power > 0 or alpha < 1/255 is
// `continue`d there without touching the pixel's state, so a pair that PROVABLY fails the
// alpha test can be skipped without changing one output bit. We prove it at two levels:
//   level 1 (staging)  : the whole 16x16 tile          -> entry never reaches shared lists
//   level 2 (per warp) : each 8x4 pixel block of a warp -> block skipped by a uniform branch
// Both use the exact minimum of the conic form over the rectangle plus a rounding margin.
#pragma once
#include "kernels.cuh"

namespace b200s {

constexpr float BLEND_ALPHA_MIN = 1.0f / 255.0f;
#ifndef B200S_BLEND_SPT
#define B200S_BLEND_SPT 2
#endif
constexpr int BLEND_SPT = B200S_BLEND_SPT;   // list entries staged per thread per round

// Minimum resident CTAs per SM requested from ptxas (register cap) per warps-per-tile variant
// (B200S_{FWD,BWD}_MINB{1,2,4,8}; 1 = no cap).
#ifndef B200S_FWD_MINB1
#define B200S_FWD_MINB1 16
#endif
#ifndef B200S_FWD_MINB2
#define B200S_FWD_MINB2 16
#endif
#ifndef B200S_FWD_MINB4
#define B200S_FWD_MINB4 8
#endif
#ifndef B200S_FWD_MINB8
#define B200S_FWD_MINB8 1
#endif
#ifndef B200S_BWD_MINB1
#define B200S_BWD_MINB1 16
#endif
#ifndef B200S_BWD_MINB2
#define B200S_BWD_MINB2 12
#endif
#ifndef B200S_BWD_MINB4
#define B200S_BWD_MINB4 8
#endif
#ifndef B200S_BWD_MINB8
#define B200S_BWD_MINB8 1
#endif
constexpr int fwd_min_blocks(int nw) {
    return nw == 1 ? B200S_FWD_MINB1 : nw == 2 ? B200S_FWD_MINB2 : nw == 4 ? B200S_FWD_MINB4 : B200S_FWD_MINB8;
}
constexpr int bwd_min_blocks(int nw) {
    return nw == 1 ? B200S_BWD_MINB1 : nw == 2 ? B200S_BWD_MINB2 : nw == 4 ? B200S_BWD_MINB4 : B200S_BWD_MINB8;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

constexpr float LOG2E_F = 1.4426950408889634f;
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Per-Gaussian constants of the rectangle test.
struct ReachCtx {
    float x, y, a, b, c;
    float nb_c, nb_a;   // -b/c, -b/a : unconstrained minimisers along a vertical / horizontal edge
    float lim;          // reach iff min over the rectangle of (a dx^2 + 2b dx dy + c dy^2) <= lim
    bool claim;         // false: conic not positive definite -> never cull
};

// thr = power threshold below which opacity*exp(power) < 1/255 is certain (with margin).
// (X0..Y1) = the TILE rectangle; its worst-case |form| bounds the fp32 rounding margin for
// every sub-rectangle of the tile.
__device__ __forceinline__ ReachCtx make_reach_ctx(float x, float y, float a, float b, float c, float thr,
                                                   float X0, float Y0, float X1, float Y1) {
    ReachCtx r;
    r.x = x; r.y = y; r.a = a; r.b = b; r.c = c;
    r.claim = (a > 0.0f && c > 0.0f && a * c > b * b);
    r.nb_c = __fdividef(-b, c);
    r.nb_a = __fdividef(-b, a);
    const float mx = fmaxf(fabsf(X0 - x), fabsf(X1 - x)), my = fmaxf(fabsf(Y0 - y), fabsf(Y1 - y));
    const float S = a * mx * mx + 2.0f * fabsf(b) * mx * my + c * my * my;
    r.lim = -2.0f * thr + 4e-6f * S + 1e-4f;
    return r;
}

// Can the Gaussian reach alpha >= 1/255 (with power <= 0) at ANY point of [X0,X1]x[Y0,Y1]?
// NaNs fall through to "yes".
__device__ __forceinline__ bool rect_reach(const ReachCtx& r, float X0, float Y0, float X1, float Y1) {
    const float ex = fminf(fmaxf(r.x, X0), X1) - r.x;
    const float ey = fminf(fmaxf(r.y, Y0), Y1) - r.y;
    const float dx0 = X0 - r.x, dx1 = X1 - r.x, dy0 = Y0 - r.y, dy1 = Y1 - r.y;
    // nearest vertical edge (dx = ex), minimise over dy; nearest horizontal edge likewise
    const float dy = fminf(fmaxf(r.nb_c * ex, dy0), dy1);
    const float qv = r.a * ex * ex + 2.0f * r.b * ex * dy + r.c * dy * dy;
    const float dx = fminf(fmaxf(r.nb_a * ey, dx0), dx1);
    const float qh = r.a * dx * dx + 2.0f * r.b * dx * ey + r.c * ey * ey;
    float qmin = 3.0e38f;
    if (ex != 0.0f) qmin = qv;
    if (ey != 0.0f) qmin = fminf(qmin, qh);
    const bool inside = (ex == 0.0f && ey == 0.0f);
    return inside || !r.claim || !(qmin > r.lim);
}

// power threshold of a Gaussian: power < thr  =>  opacity*exp(power) < 1/255
__device__ __forceinline__ float power_threshold(float op) { return -logf(255.0f * op) - 1e-4f; }

// Pixel blocks of a warp. A tile is 2 (x) by 4 (y) blocks of 8x4 pixels; with NW warps per tile a
// warp owns PX = 8/NW blocks and lane l owns pixel (l & 7, l >> 3) of each of them.
template <int NW>
struct WarpBlocks {
    static constexpr int PX = 8 / NW;
    static constexpr int NX = PX >= 2 ? 2 : 1;
    static constexpr int NY = PX >= 2 ? PX / 2 : 1;
    __device__ static __forceinline__ int col_block(int warp, int ix) { return (warp * PX + ix) & 1; }
    __device__ static __forceinline__ int row_block(int warp, int iy) { return (warp * PX + 2 * iy) >> 1; }
};

}  // namespace b200s
