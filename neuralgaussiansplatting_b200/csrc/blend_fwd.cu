// K6 — forward alpha compositing, one CTA of NW warps per 16x16 tile.
// Replaces renderCUDA<3> forward (reference forward.cu:261-374).
//
// Design (vs. the reference's 256 threads x 1 pixel):
//  * each lane owns PX = 8/NW pixels (8x4 sub-blocks), so a staged Gaussian is read from
//    shared memory once per PX pixels and the dx-only / dy-only factors of the quadratic
//    form are shared between pixels of the same column / row; PX independent T-chains per
//    lane provide the ILP the reference gets from occupancy.
//  * while staging a batch every thread tests ITS Gaussian against the tile rectangle
//    (exact minimum of the conic form over the tile, conservative margin) and only survivors
//    are written to shared memory. Entries that cannot reach alpha >= 1/255 anywhere in the
//    tile are `continue`d by every pixel in the reference, so dropping them changes no
//    output bit; survivors keep their list position so n_contrib stays bit-exact.
//  * a per-Gaussian power threshold skips expf for pixels that provably fail the 1/255 test.
//  * colour comes from the staged record (the reference gathers features[] from global
//    memory per contributing pixel, forward.cu:355).
//  * the arithmetic of every decision (power, alpha, T test) mirrors the reference SASS.

#include "kernels.cuh"

namespace b200s {


constexpr float ALPHA_MIN = 1.0f / 255.0f;

// Conservative test: can the Gaussian reach alpha >= 1/255 (with power <= 0) at ANY point of
// the pixel rectangle [X0,X1]x[Y0,Y1]? Also returns the per-Gaussian power threshold below
// which alpha < 1/255 is certain. NaNs fall through to "keep".
__device__ __forceinline__ bool splat_reaches_tile(float x, float y, float a, float b, float c, float op,
                                                   float X0, float Y0, float X1, float Y1, float& thr) {
    thr = -logf(255.0f * op) - 1e-4f;           // power < thr  =>  op*exp(power) < 1/255
    if (op < ALPHA_MIN) return false;           // exp(power) <= 1 for power <= 0
    const float ex = fminf(fmaxf(x, X0), X1) - x;
    const float ey = fminf(fmaxf(y, Y0), Y1) - y;
    if (ex == 0.0f && ey == 0.0f) return true;  // centre inside the tile
    if (!(a > 0.0f && c > 0.0f && a * c > b * b)) return true;   // not positive definite: no claim
    const float dx0 = X0 - x, dx1 = X1 - x, dy0 = Y0 - y, dy1 = Y1 - y;
    float qmin = 3.0e38f;
    if (ex != 0.0f) {   // nearest vertical edge, minimise over dy
        const float dy = fminf(fmaxf(__fdividef(-b * ex, c), dy0), dy1);
        qmin = fminf(qmin, a * ex * ex + 2.0f * b * ex * dy + c * dy * dy);
    }
    if (ey != 0.0f) {   // nearest horizontal edge, minimise over dx
        const float dx = fminf(fmaxf(__fdividef(-b * ey, a), dx0), dx1);
        qmin = fminf(qmin, a * dx * dx + 2.0f * b * dx * ey + c * ey * ey);
    }
    const float mx = fmaxf(fabsf(dx0), fabsf(dx1)), my = fmaxf(fabsf(dy0), fabsf(dy1));
    const float S = a * mx * mx + 2.0f * fabsf(b) * mx * my + c * my * my;
    return !(qmin > -2.0f * thr + 4e-6f * S + 1e-4f);
}

template <int NW>
__global__ void __launch_bounds__(32 * NW)
blend_fwd_kernel(const __grid_constant__ BlendFwdParams p) {
    constexpr int PX = 8 / NW;                 // pixels per lane
    constexpr int NX = PX >= 2 ? 2 : 1;        // distinct pixel columns per lane
    constexpr int NY = PX >= 2 ? PX / 2 : 1;   // distinct pixel rows per lane
    constexpr int BATCH = 32 * NW;

    __shared__ float4 s_a[2][BATCH];           // x, y, conic a, conic b
    __shared__ float4 s_b[2][BATCH];           // conic c, opacity, power threshold, list pos + 1
    __shared__ float4 s_c[2][BATCH];           // r, g, b
    __shared__ int s_cnt[2][NW];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tile_x = tile % p.grid_x, tile_y = tile / p.grid_x;
    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);

    // pixel ownership
    float pxf[NX], pyf[NY];
    int pix_x[NX], pix_y[NY];
    const int b0 = warp * PX;
#pragma unroll
    for (int i = 0; i < NX; ++i) {
        pix_x[i] = tile_x * TILE + (((b0 + i) & 1) * 8) + (lane & 7);
        pxf[i] = (float)pix_x[i];
    }
#pragma unroll
    for (int i = 0; i < NY; ++i) {
        pix_y[i] = tile_y * TILE + (((b0 + 2 * i) >> 1) * 4) + (lane >> 3);
        pyf[i] = (float)pix_y[i];
    }
    constexpr uint32_t ALL_DONE = (1u << PX) - 1u;
    uint32_t done = 0;
#pragma unroll
    for (int s = 0; s < PX; ++s)
        if (!(pix_x[s % NX] < p.W && pix_y[s / NX] < p.H)) done |= 1u << s;
    const uint32_t outside = done;

    float T[PX], Cr[PX], Cg[PX], Cb[PX];
    uint32_t last[PX];
#pragma unroll
    for (int s = 0; s < PX; ++s) { T[s] = 1.0f; Cr[s] = Cg[s] = Cb[s] = 0.0f; last[s] = 0; }

    const float X0 = (float)(tile_x * TILE), Y0 = (float)(tile_y * TILE);
    const float X1 = X0 + (float)(TILE - 1), Y1 = Y0 + (float)(TILE - 1);

    const int nb = (n + BATCH - 1) / BATCH;
    const uint32_t* list = p.point_list + range.x;
    // software pipeline: ids two batches ahead, records one batch ahead
    int id0 = (tid < n) ? (int)list[tid] : -1;
    int id1 = (BATCH + tid < n) ? (int)list[BATCH + tid] : -1;
    float4 q0 = make_float4(0, 0, 0, 0), q1 = q0, q2 = q0;
    if (id0 >= 0) { const SplatRec* r = p.rec + id0; q0 = r->q0; q1 = r->q1; q2 = r->q2; }

    for (int b = 0; b < nb; ++b) {
        const int buf = b & 1;
        const int nxt = (b + 2) * BATCH + tid;
        const int id2 = (nxt < n) ? (int)list[nxt] : -1;
        float4 n0 = make_float4(0, 0, 0, 0), n1 = n0, n2 = n0;
        if (id1 >= 0) { const SplatRec* r = p.rec + id1; n0 = r->q0; n1 = r->q1; n2 = r->q2; }

        float thr;
        const bool keep = (id0 >= 0) && splat_reaches_tile(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, X0, Y0, X1, Y1, thr);
        const uint32_t ballot = __ballot_sync(FULL_MASK, keep);
        if (keep) {
            const int pos = warp * 32 + __popc(ballot & lanemask_lt());
            s_a[buf][pos] = q0;
            s_b[buf][pos] = make_float4(q1.x, q1.y, thr, __uint_as_float((uint32_t)(b * BATCH + tid + 1)));
            s_c[buf][pos] = make_float4(q1.z, q1.w, q2.x, 0.0f);
        }
        if (lane == 0) s_cnt[buf][warp] = __popc(ballot);

        bool all_done;
        if (NW == 1) {
            __syncwarp();
            all_done = __all_sync(FULL_MASK, done == ALL_DONE);
        } else {
            all_done = __syncthreads_and(done == ALL_DONE);
        }
        if (all_done) break;

        for (int w = 0; w < NW; ++w) {
            const int cnt = s_cnt[buf][w];
            for (int j = 0; j < cnt; ++j) {
                const float4 A = s_a[buf][w * 32 + j];
                const float4 B = s_b[buf][w * 32 + j];
                float dx[NX], ax[NX], bx[NX], dy[NY], cy[NY];
#pragma unroll
                for (int i = 0; i < NX; ++i) {
                    dx[i] = fsub(A.x, pxf[i]);
                    ax[i] = fmul(dx[i], A.z);
                    bx[i] = fmul(dx[i], A.w);
                }
#pragma unroll
                for (int i = 0; i < NY; ++i) {
                    dy[i] = fsub(A.y, pyf[i]);
                    cy[i] = fmul(dy[i], fmul(dy[i], B.x));
                }
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    const int ix = s % NX, iy = s / NX;
                    const float sq = ffma(dx[ix], ax[ix], cy[iy]);
                    const float power = ffma(sq, -0.5f, -fmul(dy[iy], bx[ix]));
                    if ((done >> s) & 1u) continue;
                    if (power > 0.0f || power < B.z) continue;
                    const float alpha = fminf(0.99f, fmul(B.y, expf(power)));
                    if (alpha < ALPHA_MIN) continue;
                    const float test_T = fmul(T[s], fsub(1.0f, alpha));
                    if (test_T < 0.0001f) { done |= 1u << s; continue; }
                    const float4 C = s_c[buf][w * 32 + j];
                    Cr[s] = ffma(T[s], fmul(alpha, C.x), Cr[s]);
                    Cg[s] = ffma(T[s], fmul(alpha, C.y), Cg[s]);
                    Cb[s] = ffma(T[s], fmul(alpha, C.z), Cb[s]);
                    T[s] = test_T;
                    last[s] = __float_as_uint(B.w);
                }
            }
        }
        id0 = id1; id1 = id2; q0 = n0; q1 = n1; q2 = n2;
    }

    // epilogue
    const float bg0 = p.background[0], bg1 = p.background[1], bg2 = p.background[2];
    const size_t plane = (size_t)p.W * p.H;
    uint32_t lmax = 0;
#pragma unroll
    for (int s = 0; s < PX; ++s) {
        if ((outside >> s) & 1u) continue;
        const size_t pix = (size_t)pix_y[s / NX] * p.W + pix_x[s % NX];
        p.final_T[pix] = T[s];
        p.n_contrib[pix] = last[s];
        p.out_color[pix] = ffma(bg0, T[s], Cr[s]);
        p.out_color[plane + pix] = ffma(bg1, T[s], Cg[s]);
        p.out_color[2 * plane + pix] = ffma(bg2, T[s], Cb[s]);
        lmax = max(lmax, last[s]);
    }
    lmax = __reduce_max_sync(FULL_MASK, lmax);
    if (lane == 0) {
        if (NW == 1) p.tile_last[tile] = lmax;
        else atomicMax(&p.tile_last[tile], lmax);
    }
}

// Roofline accounting only: one thread per pixel replays the reference traversal
// (forward.cu:325-351) and counts traversed / contributing pairs, plus the backward's traversal.
__global__ void __launch_bounds__(256)
count_pairs_kernel(const __grid_constant__ BlendFwdParams p, unsigned long long* counts) {
    const int tile = blockIdx.x;
    const int tx = tile % p.grid_x, ty = tile / p.grid_x;
    const int x = tx * TILE + (threadIdx.x & 15), y = ty * TILE + (threadIdx.x >> 4);
    unsigned long long trav = 0, contrib = 0, bwd = 0;
    if (x < p.W && y < p.H) {
        const uint2 range = p.ranges[tile];
        const float pxf = (float)x, pyf = (float)y;
        float T = 1.0f;
        for (uint32_t k = range.x; k < range.y; ++k) {
            ++trav;
            const SplatRec* r = p.rec + p.point_list[k];
            const float4 q0 = r->q0;
            const float4 q1 = r->q1;
            const float power = power_ref(fsub(q0.x, pxf), fsub(q0.y, pyf), q0.z, q0.w, q1.x);
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, fmul(q1.y, expf(power)));
            if (alpha < ALPHA_MIN) continue;
            const float test_T = fmul(T, fsub(1.0f, alpha));
            if (test_T < 0.0001f) break;
            T = test_T;
            ++contrib;
        }
        bwd = p.n_contrib[(size_t)y * p.W + x];
    }
    for (int o = 16; o > 0; o >>= 1) {
        trav += __shfl_xor_sync(FULL_MASK, trav, o);
        contrib += __shfl_xor_sync(FULL_MASK, contrib, o);
        bwd += __shfl_xor_sync(FULL_MASK, bwd, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&counts[0], trav); atomicAdd(&counts[1], contrib); atomicAdd(&counts[2], bwd);
    }
}

int launch_count_pairs(const BlendFwdParams& p, unsigned long long* counts, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    if (tiles <= 0) return 0;
    count_launch();
    count_pairs_kernel<<<tiles, 256, 0, stream>>>(p, counts);
    return (int)cudaGetLastError();
}

template <int NW>
static int launch_blend_fwd_nw(const BlendFwdParams& p, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    count_launch();
    blend_fwd_kernel<NW><<<tiles, 32 * NW, 0, stream>>>(p);
    return (int)cudaGetLastError();
}

int launch_blend_fwd(const BlendFwdParams& p, int nw, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    if (tiles <= 0) return 0;
    if (nw != 1) {
        cudaError_t e = cudaMemsetAsync(p.tile_last, 0, sizeof(uint32_t) * (size_t)tiles, stream);
        if (e != cudaSuccess) return (int)e;
    }
    switch (nw) {
        case 1: return launch_blend_fwd_nw<1>(p, stream);
        case 2: return launch_blend_fwd_nw<2>(p, stream);
        case 4: return launch_blend_fwd_nw<4>(p, stream);
        default: return launch_blend_fwd_nw<8>(p, stream);
    }
}

}  // namespace b200s
