// K7 — backward of the alpha compositing, back-to-front, one CTA of NW warps per tile.
// Replaces renderCUDA<3> backward (reference backward.cu:399-557), whose cost is the nine
// global float atomics issued by every contributing (pixel, Gaussian) pair (:523,545-554).
//
// Here each lane owns PX = 8/NW pixels and first sums its own pixels' contributions in
// registers; the nine per-Gaussian partials are then reduced across the warp with shuffles
// and leave the SM as ONE 36-byte reduction per (warp, Gaussian) — or, in deterministic mode,
// as a plain store into the instance's own slot (no atomics at all; preprocess_bwd sums a
// Gaussian's slots in tile order). Staging applies the same exact tile culling as the
// forward, so entries no pixel of the tile can reach cost nothing.
//
// Accumulated raw sums per Gaussian (ACC_STRIDE floats):
//   [0..2] sum alpha*T*dL/dpix            = dL/dRGB
//   [3] S1 = sum u*dx   [4] S2 = sum u*dy                      u = opacity*G*dL/dalpha
//   [5] Sxx = sum u*dx*dx  [6] Sxy = sum u*dx*dy  [7] Syy = sum u*dy*dy
//   [8] sum G*dL/dalpha                   = dL/dopacity
// preprocess_bwd turns them into dL/dmean2D = -(a*S1+b*S2, c*S2+b*S1)*(W/2,H/2) and
// dL/dconic = -0.5*(Sxx, Sxy, Syy)   (backward.cu:538-554 factored).

#include "kernels.cuh"

namespace b200s {


constexpr float ALPHA_MIN_B = 1.0f / 255.0f;

__device__ __forceinline__ bool splat_reaches_tile_b(float x, float y, float a, float b, float c, float op,
                                                     float X0, float Y0, float X1, float Y1, float& thr) {
    thr = -logf(255.0f * op) - 1e-4f;
    if (op < ALPHA_MIN_B) return false;
    const float ex = fminf(fmaxf(x, X0), X1) - x;
    const float ey = fminf(fmaxf(y, Y0), Y1) - y;
    if (ex == 0.0f && ey == 0.0f) return true;
    if (!(a > 0.0f && c > 0.0f && a * c > b * b)) return true;
    const float dx0 = X0 - x, dx1 = X1 - x, dy0 = Y0 - y, dy1 = Y1 - y;
    float qmin = 3.0e38f;
    if (ex != 0.0f) {
        const float dy = fminf(fmaxf(__fdividef(-b * ex, c), dy0), dy1);
        qmin = fminf(qmin, a * ex * ex + 2.0f * b * ex * dy + c * dy * dy);
    }
    if (ey != 0.0f) {
        const float dx = fminf(fmaxf(__fdividef(-b * ey, a), dx0), dx1);
        qmin = fminf(qmin, a * dx * dx + 2.0f * b * dx * ey + c * ey * ey);
    }
    const float mx = fmaxf(fabsf(dx0), fabsf(dx1)), my = fmaxf(fabsf(dy0), fabsf(dy1));
    const float S = a * mx * mx + 2.0f * fabsf(b) * mx * my + c * my * my;
    return !(qmin > -2.0f * thr + 4e-6f * S + 1e-4f);
}

template <int NW>
__global__ void __launch_bounds__(32 * NW)
blend_bwd_kernel(const __grid_constant__ BlendBwdParams p) {
    constexpr int PX = 8 / NW;
    constexpr int NX = PX >= 2 ? 2 : 1;
    constexpr int NY = PX >= 2 ? PX / 2 : 1;
    constexpr int BATCH = 32 * NW;

    __shared__ float4 s_a[2][BATCH];           // x, y, conic a, conic b
    __shared__ float4 s_b[2][BATCH];           // conic c, opacity, power threshold, list pos
    __shared__ float4 s_c[2][BATCH];           // r, g, b, destination (Gaussian id | slot)
    __shared__ int s_cnt[2][NW];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tile_x = tile % p.grid_x, tile_y = tile / p.grid_x;
    const uint2 range = p.ranges[tile];
    // Entries at list positions >= tile_last contributed to no pixel of the tile.
    const int n = min((int)(range.y - range.x), (int)p.tile_last[tile]);
    if (n <= 0) return;

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

    const float bg0 = p.background[0], bg1 = p.background[1], bg2 = p.background[2];
    const size_t plane = (size_t)p.W * p.H;
    float T[PX], Tf_bg[PX], Ar[PX], Ag[PX], Ab[PX], gr[PX], gg[PX], gb[PX];
    uint32_t lastc[PX];
#pragma unroll
    for (int s = 0; s < PX; ++s) {
        const bool inside = pix_x[s % NX] < p.W && pix_y[s / NX] < p.H;
        const size_t pix = inside ? (size_t)pix_y[s / NX] * p.W + pix_x[s % NX] : 0;
        T[s] = inside ? p.final_T[pix] : 0.0f;
        lastc[s] = inside ? p.n_contrib[pix] : 0u;
        gr[s] = inside ? p.dL_dpix[pix] : 0.0f;
        gg[s] = inside ? p.dL_dpix[plane + pix] : 0.0f;
        gb[s] = inside ? p.dL_dpix[2 * plane + pix] : 0.0f;
        // -T_final * (bg . dL/dpix): the background term of dL/dalpha (backward.cu:531-534)
        Tf_bg[s] = -T[s] * (bg0 * gr[s] + bg1 * gg[s] + bg2 * gb[s]);
        Ar[s] = Ag[s] = Ab[s] = 0.0f;
    }

    const float X0 = (float)(tile_x * TILE), Y0 = (float)(tile_y * TILE);
    const float X1 = X0 + (float)(TILE - 1), Y1 = Y0 + (float)(TILE - 1);

    const int nb = (n + BATCH - 1) / BATCH;
    const uint32_t* list = p.point_list + range.x;
    // batch b covers list positions [n-(b+1)*BATCH, n-b*BATCH); thread t takes position
    // n-b*BATCH-BATCH+t so that warp/lane order stays list order.
    auto pos_of = [&](int b) { return n - (b + 1) * BATCH + tid; };
    int pos0 = pos_of(0);
    int id0 = (pos0 >= 0) ? (int)list[pos0] : -1;
    int pos1 = pos_of(1);
    int id1 = (pos1 >= 0) ? (int)list[pos1] : -1;
    float4 q0 = make_float4(0, 0, 0, 0), q1 = q0, q2 = q0;
    if (id0 >= 0) { const SplatRec* r = p.rec + id0; q0 = r->q0; q1 = r->q1; q2 = r->q2; }

    for (int b = 0; b < nb; ++b) {
        const int buf = b & 1;
        const int pos2 = pos_of(b + 2);
        const int id2 = (pos2 >= 0) ? (int)list[pos2] : -1;
        float4 n0 = make_float4(0, 0, 0, 0), n1 = n0, n2 = n0;
        if (id1 >= 0) { const SplatRec* r = p.rec + id1; n0 = r->q0; n1 = r->q1; n2 = r->q2; }

        float thr;
        const bool keep = (id0 >= 0) && splat_reaches_tile_b(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, X0, Y0, X1, Y1, thr);
        const uint32_t ballot = __ballot_sync(FULL_MASK, keep);
        if (keep) {
            uint32_t dest = (uint32_t)id0;
            if (p.deterministic) {
                // slot of this instance in the unsorted (Gaussian-major, tile-minor) order
                const uint32_t rmin = __float_as_uint(q2.z), rwh = __float_as_uint(q2.w);
                const uint32_t first = id0 == 0 ? 0u : p.point_offsets[id0 - 1];
                dest = first + ((uint32_t)tile_y - (rmin >> 16)) * (rwh & 0xffffu) + ((uint32_t)tile_x - (rmin & 0xffffu));
            }
            const int slot = warp * 32 + __popc(ballot & lanemask_lt());
            s_a[buf][slot] = q0;
            s_b[buf][slot] = make_float4(q1.x, q1.y, thr, __uint_as_float((uint32_t)pos0));
            s_c[buf][slot] = make_float4(q1.z, q1.w, q2.x, __uint_as_float(dest));
        }
        if (lane == 0) s_cnt[buf][warp] = __popc(ballot);
        if (NW == 1) __syncwarp(); else __syncthreads();

        for (int w = NW - 1; w >= 0; --w) {
            const int cnt = s_cnt[buf][w];
            for (int j = cnt - 1; j >= 0; --j) {
                const float4 A = s_a[buf][w * 32 + j];
                const float4 B = s_b[buf][w * 32 + j];
                const float4 C = s_c[buf][w * 32 + j];
                const uint32_t lpos = __float_as_uint(B.w);
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
                float acc[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) acc[i] = 0.0f;
                bool any = false;
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    const int ix = s % NX, iy = s / NX;
                    const float sq = ffma(dx[ix], ax[ix], cy[iy]);
                    const float power = ffma(sq, -0.5f, -fmul(dy[iy], bx[ix]));
                    if (lpos >= lastc[s]) continue;          // behind this pixel's last contributor
                    if (power > 0.0f || power < B.z) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, fmul(B.y, G));
                    if (alpha < ALPHA_MIN_B) continue;
                    any = true;
                    const float one_m = 1.0f - alpha;
                    const float inv = __frcp_rn(one_m);
                    T[s] = T[s] * inv;
                    const float wgt = alpha * T[s];
                    acc[0] += wgt * gr[s];
                    acc[1] += wgt * gg[s];
                    acc[2] += wgt * gb[s];
                    float dL_dalpha = ((C.x - Ar[s]) * gr[s] + (C.y - Ag[s]) * gg[s] + (C.z - Ab[s]) * gb[s]) * T[s];
                    dL_dalpha += Tf_bg[s] * inv;
                    Ar[s] = alpha * C.x + one_m * Ar[s];
                    Ag[s] = alpha * C.y + one_m * Ag[s];
                    Ab[s] = alpha * C.z + one_m * Ab[s];
                    const float gda = G * dL_dalpha;
                    const float u = B.y * gda;
                    const float udx = u * dx[ix], udy = u * dy[iy];
                    acc[3] += udx;
                    acc[4] += udy;
                    acc[5] += udx * dx[ix];
                    acc[6] += udx * dy[iy];
                    acc[7] += udy * dy[iy];
                    acc[8] += gda;
                }
                if (!__any_sync(FULL_MASK, any)) continue;
#pragma unroll
                for (int i = 0; i < 9; ++i) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(FULL_MASK, acc[i], o);
                }
                if (lane < 9) {
                    float v = acc[0];
#pragma unroll
                    for (int i = 1; i < 9; ++i) v = (lane == i) ? acc[i] : v;
                    const uint32_t dest = __float_as_uint(C.w);
                    if (p.deterministic) {
                        if (NW == 1) p.partial[(size_t)dest * ACC_STRIDE + lane] = v;
                        else atomicAdd(&p.partial[(size_t)dest * ACC_STRIDE + lane], v);
                    } else {
                        atomicAdd(&p.accum[(size_t)dest * ACC_STRIDE + lane], v);
                    }
                }
            }
        }
        id0 = id1; id1 = id2; pos0 = pos1; pos1 = pos2; q0 = n0; q1 = n1; q2 = n2;
    }
}

template <int NW>
static int launch_blend_bwd_nw(const BlendBwdParams& p, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    count_launch();
    blend_bwd_kernel<NW><<<tiles, 32 * NW, 0, stream>>>(p);
    return (int)cudaGetLastError();
}

int launch_blend_bwd(const BlendBwdParams& p, int nw, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    if (tiles <= 0) return 0;
    if (p.deterministic) nw = 1;   // one warp owns the whole (tile, Gaussian) sum: plain stores
    switch (nw) {
        case 1: return launch_blend_bwd_nw<1>(p, stream);
        case 2: return launch_blend_bwd_nw<2>(p, stream);
        case 4: return launch_blend_bwd_nw<4>(p, stream);
        default: return launch_blend_bwd_nw<8>(p, stream);
    }
}

}  // namespace b200s
