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

#include "blend_common.cuh"

namespace b200s {


constexpr float ALPHA_MIN_B = BLEND_ALPHA_MIN;

// Sum of nine per-lane partials over the warp. The first eight are reduced with a transposing
// butterfly (4 + 2 + 1 shuffles leave lane l with the sum, over the lanes that differ from it in
// bits 4..2, of value ((l >> 4) & 1) * 4 + ((l >> 3) & 1) * 2 + ((l >> 2) & 1)); the ninth takes
// the plain steps 16, 8, 4; the last two steps (2, 1) finish both at once, the ninth riding in the
// lanes with bit 1 set: 12 shuffles instead of 45. Returns this lane's value and its index in
// `which` (-1: nothing to write; lanes 0, 4, .., 28 write values 0..7, lane 2 writes the ninth).
__device__ __forceinline__ float reduce9(float (&acc)[9], int lane, int& which) {
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float send = hi ? acc[i] : acc[i + 4];
            const float keep = hi ? acc[i + 4] : acc[i];
            acc[i] = keep + __shfl_xor_sync(FULL_MASK, send, 16);
        }
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float send = hi ? acc[i] : acc[i + 2];
            const float keep = hi ? acc[i + 2] : acc[i];
            acc[i] = keep + __shfl_xor_sync(FULL_MASK, send, 8);
        }
    }
    {
        const bool hi = lane & 4;
        const float send = hi ? acc[0] : acc[1];
        const float keep = hi ? acc[1] : acc[0];
        acc[0] = keep + __shfl_xor_sync(FULL_MASK, send, 4);
    }
#pragma unroll
    for (int o = 16; o >= 4; o >>= 1) acc[8] += __shfl_xor_sync(FULL_MASK, acc[8], o);
    const bool hi = lane & 2;
    const float send = hi ? acc[0] : acc[8];
    const float keep = hi ? acc[8] : acc[0];
    float r = keep + __shfl_xor_sync(FULL_MASK, send, 2);
    r += __shfl_xor_sync(FULL_MASK, r, 1);
    which = -1;
    if ((lane & 3) == 0) which = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    if (lane == 2) which = 8;
    return r;
}

template <int NW, int SPT, bool FAST>
__global__ void __launch_bounds__(32 * NW, bwd_min_blocks(NW))
blend_bwd_kernel(const __grid_constant__ BlendBwdParams p) {
    using WB = WarpBlocks<NW>;
    constexpr int PX = WB::PX, NX = WB::NX, NY = WB::NY;
    constexpr int NT = 32 * NW;
    constexpr int BATCH = NT * SPT;
    constexpr int NSEG = NW * SPT;

    // staged records, [buffer][field][entry]: field 0 = x, y, conic a, conic b
    //   field 1 = conic c, opacity, r, g
    //   field 2 = b, power threshold, list pos, destination
    // One array, so that the three 16-byte loads of a survivor share one address register (constant offsets).
    __shared__ float4 s_q[2][3][BATCH];
    __shared__ uint16_t s_idx[BATCH];
    __shared__ int s_cnt[NSEG];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = p.tile_order ? (int)p.tile_order[blockIdx.x] : (int)blockIdx.x;
    const int tile_x = tile % p.grid_x, tile_y = tile / p.grid_x;
    const uint2 range = p.ranges[tile];
    // Entries at list positions >= tile_last contributed to no pixel of the tile.
    const int n = min((int)(range.y - range.x), (int)p.tile_last[tile]);
    B200S_CHECK(range.x <= range.y && (int64_t)range.y <= p.list_cap);
    if (n <= 0) return;

    float pxf[NX], pyf[NY], bx0[NX], by0[NY];
    int pix_x[NX], pix_y[NY];
#pragma unroll
    for (int i = 0; i < NX; ++i) {
        const int x0 = tile_x * TILE + WB::col_block(warp, i) * 8;
        bx0[i] = (float)x0;
        pix_x[i] = x0 + (lane & 7);
        pxf[i] = (float)pix_x[i];
    }
#pragma unroll
    for (int i = 0; i < NY; ++i) {
        const int y0 = tile_y * TILE + WB::row_block(warp, i) * 4;
        by0[i] = (float)y0;
        pix_y[i] = y0 + (lane >> 3);
        pyf[i] = (float)pix_y[i];
    }

    const float bg0 = p.background[0], bg1 = p.background[1], bg2 = p.background[2];
    const size_t plane = (size_t)p.W * p.H;
    float T[PX], Tf_bg[PX], Ar[PX], Ag[PX], Ab[PX], gr[PX], gg[PX], gb[PX];
    uint32_t lastc[PX], blocklast[PX];
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
        blocklast[s] = __reduce_max_sync(FULL_MASK, lastc[s]);   // no pixel of block s goes deeper
    }

    const float X0 = (float)(tile_x * TILE), Y0 = (float)(tile_y * TILE);
    const float X1 = X0 + (float)(TILE - 1), Y1 = Y0 + (float)(TILE - 1);

    const int nb = (n + BATCH - 1) / BATCH;
    const uint32_t* list = p.point_list + range.x;
    const uint32_t lt = lanemask_lt();
    // batch b covers list positions [n-(b+1)*BATCH, n-b*BATCH); element i of the batch is
    // position n-(b+1)*BATCH+i, so element order stays list order.
    int idc[SPT], idn[SPT];
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
        const int i = k * NT + tid;
        const int pos = n - BATCH + i;
        idc[k] = (pos >= 0) ? (int)list[pos] : -1;
        B200S_CHECK(idc[k] < p.P);
        if (idc[k] >= 0) {
            const SplatRec* r = p.rec + idc[k];
            cp_async16(&s_q[0][0][i], &r->q0);
            cp_async16(&s_q[0][1][i], &r->q1);
            cp_async16(&s_q[0][2][i], &r->q2);
        }
    }
    cp_async_commit();
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
        const int pos = n - 2 * BATCH + k * NT + tid;
        idn[k] = (pos >= 0) ? (int)list[pos] : -1;
    }

    for (int b = 0; b < nb; ++b) {
        const int buf = b & 1;
        const char* qbase = reinterpret_cast<const char*>(&s_q[buf][0][0]);
        cp_async_wait<0>();
        uint32_t ballot[SPT];
        bool keep[SPT];
#pragma unroll
        for (int k = 0; k < SPT; ++k) {
            const int i = k * NT + tid;
            const int pos = n - (b + 1) * BATCH + i;
            keep[k] = false;
            if (pos >= 0) {
                const float4 q0 = s_q[buf][0][i];
                const float4 q1 = s_q[buf][1][i];
                const float thr = power_threshold(q1.y);
                if (q1.y >= ALPHA_MIN_B) {
                    const ReachCtx rc = make_reach_ctx(q0.x, q0.y, q0.z, q0.w, q1.x, thr, X0, Y0, X1, Y1);
                    keep[k] = rect_reach(rc, X0, Y0, X1, Y1);
                }
                float4 q2 = s_q[buf][2][i];
                uint32_t dest = (uint32_t)idc[k];
                if (p.deterministic) {
                    // slot of this instance in the unsorted (Gaussian-major, tile-minor) order
                    const uint32_t rmin = __float_as_uint(q2.z), rwh = __float_as_uint(q2.w);
                    const uint32_t first = idc[k] == 0 ? 0u : p.point_offsets[idc[k] - 1];
                    dest = first + ((uint32_t)tile_y - (rmin >> 16)) * (rwh & 0xffffu) + ((uint32_t)tile_x - (rmin & 0xffffu));
                }
                q2.y = thr;
                q2.z = __uint_as_float((uint32_t)pos);
                q2.w = __uint_as_float(dest);
                s_q[buf][2][i] = q2;
            }
            ballot[k] = __ballot_sync(FULL_MASK, keep[k]);
            if (lane == 0) s_cnt[k * NW + warp] = __popc(ballot[k]);
        }
        if (NW == 1) __syncwarp(); else __syncthreads();

        // prefetch the next batch into the other buffer (its previous readers passed the barrier)
        if (b + 1 < nb) {
#pragma unroll
            for (int k = 0; k < SPT; ++k) {
                if (idn[k] >= 0) {
                    const int i = k * NT + tid;
                    const SplatRec* r = p.rec + idn[k];
                    cp_async16(&s_q[buf ^ 1][0][i], &r->q0);
                    cp_async16(&s_q[buf ^ 1][1][i], &r->q1);
                    cp_async16(&s_q[buf ^ 1][2][i], &r->q2);
                }
            }
        }
        cp_async_commit();
#pragma unroll
        for (int k = 0; k < SPT; ++k) {
            idc[k] = idn[k];
            const int pos = n - (b + 3) * BATCH + k * NT + tid;
            idn[k] = (pos >= 0) ? (int)list[pos] : -1;
        }

        int total = 0;
        {
            int base[SPT];
#pragma unroll
            for (int k = 0; k < SPT; ++k) base[k] = 0;
#pragma unroll
            for (int sgm = 0; sgm < NSEG; ++sgm) {
                const int c = s_cnt[sgm];
#pragma unroll
                for (int k = 0; k < SPT; ++k)
                    if (sgm < k * NW + warp) base[k] += c;
                total += c;
            }
#pragma unroll
            for (int k = 0; k < SPT; ++k)
                if (keep[k]) {
                    B200S_CHECK(base[k] + __popc(ballot[k] & lt) < BATCH);
                    s_idx[base[k] + __popc(ballot[k] & lt)] = (uint16_t)((k * NT + tid) << 4);
                }
        }
        if (NW == 1) __syncwarp(); else __syncthreads();

        for (int j0 = ((total - 1) >> 5) << 5; j0 >= 0; j0 -= 32) {
            // level 2: lane l tests survivor j0+l against each pixel block of this warp
            uint32_t packed = 0;
            if (j0 + lane < total) {
                const uint32_t idx = s_idx[j0 + lane];          // byte offset of the entry (index << 4)
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float4 Q = *reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16);
                const uint32_t lpos = __float_as_uint(Q.z);
                const ReachCtx rc = make_reach_ctx(A.x, A.y, A.z, A.w, B.x, Q.y, X0, Y0, X1, Y1);
                uint32_t bits = 0;
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    const float rx0 = bx0[s % NX], ry0 = by0[s / NX];
                    if (lpos < blocklast[s] && rect_reach(rc, rx0, ry0, rx0 + 7.0f, ry0 + 3.0f)) bits |= 1u << s;
                }
                packed = idx | (bits << 16);
            }
            uint32_t m = __ballot_sync(FULL_MASK, (packed >> 16) != 0u);
            while (m) {
                const int j = 31 - __clz(m);
                m &= ~(1u << j);
                const uint32_t pk = __shfl_sync(FULL_MASK, packed, j);
                const uint32_t idx = pk & 0xffffu, bj = pk >> 16;
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float4 Q = *reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16);
                const uint32_t lpos = __float_as_uint(Q.z);
                float dx[NX], ax[NX], bx[NX], dy[NY], cy[NY];
                // FAST (blend_mode 1): the forward's ex2.approx evaluation, conic pre-scaled into log2 units
                const float kA = FAST ? A.z * (-0.5f * LOG2E_F) : A.z;
                const float kB = FAST ? A.w * (-LOG2E_F) : A.w;
                const float kC = FAST ? B.x * (-0.5f * LOG2E_F) : B.x;
                const float thrq = FAST ? Q.y * LOG2E_F : Q.y;
#pragma unroll
                for (int i = 0; i < NX; ++i) {
                    dx[i] = fsub(A.x, pxf[i]);
                    ax[i] = fmul(dx[i], kA);
                    bx[i] = fmul(dx[i], kB);
                }
#pragma unroll
                for (int i = 0; i < NY; ++i) {
                    dy[i] = fsub(A.y, pyf[i]);
                    cy[i] = fmul(dy[i], fmul(dy[i], kC));
                }
                float acc[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) acc[i] = 0.0f;
                bool any = false;
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    if (!((bj >> s) & 1u)) continue;          // warp-uniform
                    const int ix = s % NX, iy = s / NX;
                    float power;
                    if (FAST) {
                        power = fmaf(dy[iy], bx[ix], fmaf(dx[ix], ax[ix], cy[iy]));      // log2 units
                    } else {
                        const float sq = ffma(dx[ix], ax[ix], cy[iy]);
                        power = ffma(sq, -0.5f, -fmul(dy[iy], bx[ix]));
                    }
                    if (lpos >= lastc[s]) continue;          // behind this pixel's last contributor
                    if (power > 0.0f || power < thrq) continue;
                    // Same accurate expf as the forward: T is recovered by dividing by (1 - alpha), which
                    // amplifies any mismatch with the forward's alpha by alpha / (1 - alpha) (up to 99x), so a
                    // MUFU.EX2-only G (2^-22 relative error) costs 2.4e-4 rel-L2 on dL/dopacity (measured, C1).
                    const float G = FAST ? ex2_approx(power) : expf(power);
                    const float alpha = fminf(0.99f, fmul(B.y, G));
                    if (alpha < ALPHA_MIN_B) continue;
                    any = true;
                    const float one_m = 1.0f - alpha;
                    const float inv = rcp_approx(one_m);
                    T[s] = T[s] * inv;
                    const float wgt = alpha * T[s];
                    acc[0] += wgt * gr[s];
                    acc[1] += wgt * gg[s];
                    acc[2] += wgt * gb[s];
                    // c - accumulated colour behind this Gaussian, per channel (backward.cu:505-512); the same
                    // differences update the accumulators: A += alpha * (c - A)  ==  alpha*c + (1-alpha)*A
                    const float dr_ = B.z - Ar[s], dg_ = B.w - Ag[s], db_ = Q.x - Ab[s];
                    float dL_dalpha = (dr_ * gr[s] + dg_ * gg[s] + db_ * gb[s]) * T[s];
                    dL_dalpha += Tf_bg[s] * inv;
                    Ar[s] = fmaf(alpha, dr_, Ar[s]);
                    Ag[s] = fmaf(alpha, dg_, Ag[s]);
                    Ab[s] = fmaf(alpha, db_, Ab[s]);
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
                int which;
                const float v = reduce9(acc, lane, which);
                if (which >= 0) {
                    const uint32_t dest = __float_as_uint(Q.w);
                    if (p.deterministic) {
                        if (NW == 1) p.partial[(size_t)dest * ACC_STRIDE + which] = v;
                        else atomicAdd(&p.partial[(size_t)dest * ACC_STRIDE + which], v);
                    } else {
                        atomicAdd(&p.accum[(size_t)dest * ACC_STRIDE + which], v);
                    }
                }
            }
        }
    }
    cp_async_wait<0>();
}

template <int NW>
static int launch_blend_bwd_nw(const BlendBwdParams& p, cudaStream_t stream) {
    const int tiles = p.grid_x * p.grid_y;
    count_launch();
    if (p.fast) blend_bwd_kernel<NW, (NW >= 8 ? 1 : BLEND_SPT), true><<<tiles, 32 * NW, 0, stream>>>(p);
    else blend_bwd_kernel<NW, (NW >= 8 ? 1 : BLEND_SPT), false><<<tiles, 32 * NW, 0, stream>>>(p);
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
