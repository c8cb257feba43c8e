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

#include "blend_common.cuh"

#ifndef B200S_FWD_BLOCKLISTS
#define B200S_FWD_BLOCKLISTS 0
#endif

namespace b200s {


constexpr float ALPHA_MIN = BLEND_ALPHA_MIN;

// FAST (blend_mode 1): the falloff exp(power) is evaluated as ex2.approx(power * log2 e) with the conic pre-scaled by
// -0.5*log2(e) / -log2(e) once per staged Gaussian, and the quadratic form is left to FMA contraction (2 instead of 3
// instructions per pixel, 2 instead of ~9 for the exponential). Not bit-exact: alpha differs from expf's by ~1e-7
// relative, so n_contrib can differ on isolated threshold pixels; the backward uses the same evaluation.
template <int NW, int SPT, bool FAST>
__global__ void __launch_bounds__(32 * NW, fwd_min_blocks(NW))
blend_fwd_kernel(const __grid_constant__ BlendFwdParams p) {
    using WB = WarpBlocks<NW>;
    constexpr int PX = WB::PX, NX = WB::NX, NY = WB::NY;
    constexpr int NT = 32 * NW;                // threads per tile
    constexpr int BATCH = NT * SPT;            // list entries staged per round
    constexpr int NSEG = NW * SPT;             // 32-entry segments of a batch, in list order

    // staged records, [buffer][field][entry]: field 0 = x, y, conic a, conic b
    //   field 1 = conic c, opacity, r, g
    //   field 2 = b, power threshold, list pos + 1, -
    // One array, so that the three 16-byte loads of a survivor share one address register (constant offsets).
    __shared__ float4 s_q[2][3][BATCH];
    __shared__ uint16_t s_idx[BATCH];          // tile-level survivors of the batch, list order
    __shared__ int s_cnt[NSEG];
#if B200S_FWD_BLOCKLISTS
    static_assert(BATCH <= 256, "s_blk stores entry indices in 8 bits");
    __shared__ uint8_t s_blk[NW][PX][BATCH];   // per pixel block: entry indices of the survivors that reach it
#endif

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = p.tile_order ? (int)p.tile_order[blockIdx.x] : (int)blockIdx.x;
    const int tile_x = tile % p.grid_x, tile_y = tile / p.grid_x;
    const uint2 range = p.ranges[tile];
    const int n = (int)(range.y - range.x);
    B200S_CHECK(range.x <= range.y && (int64_t)range.y <= p.list_cap);

    // pixel ownership: block s = (ix, iy) of this warp, pixel (lane & 7, lane >> 3) inside it. Only the FLOAT
    // coordinates stay live across the list loop (the integer ones are re-derived in the epilogue): with both live
    // the compiler keeps the integers and re-converts them per survivor (3 I2FP per (warp, Gaussian) in the SASS).
    float pxf[NX], pyf[NY];
    constexpr uint32_t ALL_DONE = (1u << PX) - 1u;
    uint32_t done = 0;
    {
        int pix_x[NX], pix_y[NY];
#pragma unroll
        for (int i = 0; i < NX; ++i) {
            pix_x[i] = tile_x * TILE + WB::col_block(warp, i) * 8 + (lane & 7);
            pxf[i] = (float)pix_x[i];
        }
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            pix_y[i] = tile_y * TILE + WB::row_block(warp, i) * 4 + (lane >> 3);
            pyf[i] = (float)pix_y[i];
        }
#pragma unroll
        for (int s = 0; s < PX; ++s)
            if (!(pix_x[s % NX] < p.W && pix_y[s / NX] < p.H)) done |= 1u << s;
    }
    const uint32_t outside = done;
    const float lane_x = (float)(lane & 7), lane_y = (float)(lane >> 3);

    float T[PX], Cr[PX], Cg[PX], Cb[PX];
    uint32_t last[PX];
#pragma unroll
    for (int s = 0; s < PX; ++s) { T[s] = 1.0f; Cr[s] = Cg[s] = Cb[s] = 0.0f; last[s] = 0; }

    const float X0 = (float)(tile_x * TILE), Y0 = (float)(tile_y * TILE);
    const float X1 = X0 + (float)(TILE - 1), Y1 = Y0 + (float)(TILE - 1);

    const int nb = (n + BATCH - 1) / BATCH;
    const uint32_t* list = p.point_list + range.x;
    const uint32_t lt = lanemask_lt();

    // pipeline: ids two batches ahead (registers), records one batch ahead (cp.async -> smem)
    int idn[SPT];
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
        const int pos = k * NT + tid;
        if (pos < n) {
            B200S_CHECK(list[pos] < (uint32_t)p.P);
            const SplatRec* r = p.rec + list[pos];
            cp_async16(&s_q[0][0][pos], &r->q0);
            cp_async16(&s_q[0][1][pos], &r->q1);
            cp_async16(&s_q[0][2][pos], &r->q2);
        }
    }
    cp_async_commit();
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
        const int pos = BATCH + k * NT + tid;
        idn[k] = (pos < n) ? (int)list[pos] : -1;
    }

    for (int b = 0; b < nb; ++b) {
        const int buf = b & 1;
        const char* qbase = reinterpret_cast<const char*>(&s_q[buf][0][0]);
        cp_async_wait<0>();
        // level 1: every thread tests ITS entries against the tile
        uint32_t ballot[SPT];
        bool keep[SPT];
#pragma unroll
        for (int k = 0; k < SPT; ++k) {
            const int i = k * NT + tid;
            const int pos = b * BATCH + i;
            keep[k] = false;
            if (pos < n) {
                const float4 q0 = s_q[buf][0][i];
                const float4 q1 = s_q[buf][1][i];
                const float thr = power_threshold(q1.y);
                if (q1.y >= ALPHA_MIN) {
                    const ReachCtx rc = make_reach_ctx(q0.x, q0.y, q0.z, q0.w, q1.x, thr, X0, Y0, X1, Y1);
                    keep[k] = rect_reach(rc, X0, Y0, X1, Y1);
                }
                float4 q2 = s_q[buf][2][i];
                q2.y = thr;
                q2.z = __uint_as_float((uint32_t)(pos + 1));
                s_q[buf][2][i] = q2;
            }
            ballot[k] = __ballot_sync(FULL_MASK, keep[k]);
            if (lane == 0) s_cnt[k * NW + warp] = __popc(ballot[k]);
        }
        bool all_done;
        if (NW == 1) {
            __syncwarp();
            all_done = __all_sync(FULL_MASK, done == ALL_DONE);
        } else {
            all_done = __syncthreads_and(done == ALL_DONE);
        }
        if (all_done) break;

        // prefetch the next batch into the other buffer (its previous readers passed the barrier)
        if (b + 1 < nb) {
#pragma unroll
            for (int k = 0; k < SPT; ++k) {
                if (idn[k] >= 0) {
                    B200S_CHECK(idn[k] < p.P);
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
            const int pos = (b + 2) * BATCH + k * NT + tid;
            idn[k] = (pos < n) ? (int)list[pos] : -1;
        }

        // compact the survivors' indices in list order
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

        // pixel blocks whose 32 pixels are all finished take no more entries
        uint32_t blockdone = 0;
#pragma unroll
        for (int s = 0; s < PX; ++s)
            if (__all_sync(FULL_MASK, (done >> s) & 1u)) blockdone |= 1u << s;

#if B200S_FWD_BLOCKLISTS
        // level 2 -> one survivor list PER PIXEL BLOCK (list order kept): lane l tests survivor j0+l against each block
        // of this warp, the hits are compacted per block with a ballot, and every block then walks only its own
        // list. Against the per-warp walk this drops the 4 uniform block tests, the find-first-set / shuffle
        // broadcast and the unused row / column factors of every (warp, survivor) visit.
        int bcnt[PX];
#pragma unroll
        for (int s = 0; s < PX; ++s) bcnt[s] = 0;
        for (int j0 = 0; j0 < total; j0 += 32) {
            uint32_t bits = 0, ent = 0;
            if (j0 + lane < total) {
                const uint32_t idx = s_idx[j0 + lane];          // byte offset of the entry (index << 4)
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float thr = reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16)->y;
                const ReachCtx rc = make_reach_ctx(A.x, A.y, A.z, A.w, B.x, thr, X0, Y0, X1, Y1);
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    const float rx0 = pxf[s % NX] - lane_x, ry0 = pyf[s / NX] - lane_y;   // exact: small integers
                    if (rect_reach(rc, rx0, ry0, rx0 + 7.0f, ry0 + 3.0f)) bits |= 1u << s;
                }
                bits &= ~blockdone;
                ent = idx >> 4;
            }
#pragma unroll
            for (int s = 0; s < PX; ++s) {
                const bool hit = (bits >> s) & 1u;
                const uint32_t bal = __ballot_sync(FULL_MASK, hit);
                if (hit) {
                    B200S_CHECK(bcnt[s] + __popc(bal & lt) < BATCH);
                    s_blk[warp][s][bcnt[s] + __popc(bal & lt)] = (uint8_t)ent;
                }
                bcnt[s] += __popc(bal);
            }
        }
        __syncwarp();
#pragma unroll
        for (int s = 0; s < PX; ++s) {
            const int ix = s % NX, iy = s / NX;
            const uint8_t* lst = s_blk[warp][s];
            for (int j = 0; j < bcnt[s]; ++j) {
                const uint32_t idx = (uint32_t)lst[j] << 4;
                B200S_CHECK(idx < (uint32_t)BATCH * 16u);
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float4 Q = *reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16);
                const float kA = FAST ? A.z * (-0.5f * LOG2E_F) : A.z;
                const float kB = FAST ? A.w * (-LOG2E_F) : A.w;
                const float kC = FAST ? B.x * (-0.5f * LOG2E_F) : B.x;
                const float thrq = FAST ? Q.y * LOG2E_F : Q.y;
                const float dx = fsub(A.x, pxf[ix]), dy = fsub(A.y, pyf[iy]);
                const float ax = fmul(dx, kA), bx = fmul(dx, kB), cy = fmul(dy, fmul(dy, kC));
                float power;
                if (FAST) {
                    power = fmaf(dy, bx, fmaf(dx, ax, cy));      // log2 units
                } else {
                    const float sq = ffma(dx, ax, cy);
                    power = ffma(sq, -0.5f, -fmul(dy, bx));
                }
                if ((done >> s) & 1u) continue;
                if (power > 0.0f || power < thrq) continue;
                const float alpha = fminf(0.99f, fmul(B.y, FAST ? ex2_approx(power) : expf(power)));
                if (alpha < ALPHA_MIN) continue;
                const float test_T = fmul(T[s], fsub(1.0f, alpha));
                if (test_T < 0.0001f) { done |= 1u << s; continue; }
                Cr[s] = ffma(T[s], fmul(alpha, B.z), Cr[s]);
                Cg[s] = ffma(T[s], fmul(alpha, B.w), Cg[s]);
                Cb[s] = ffma(T[s], fmul(alpha, Q.x), Cb[s]);
                T[s] = test_T;
                last[s] = __float_as_uint(Q.z);
            }
        }
        __syncwarp();
#else
        for (int j0 = 0; j0 < total; j0 += 32) {
            // level 2: lane l tests survivor j0+l against each pixel block of this warp
            uint32_t packed = 0;
            if (j0 + lane < total) {
                const uint32_t idx = s_idx[j0 + lane];          // byte offset of the entry (index << 4)
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float thr = reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16)->y;
                const ReachCtx rc = make_reach_ctx(A.x, A.y, A.z, A.w, B.x, thr, X0, Y0, X1, Y1);
                uint32_t bits = 0;
#pragma unroll
                for (int s = 0; s < PX; ++s) {
                    const float rx0 = pxf[s % NX] - lane_x, ry0 = pyf[s / NX] - lane_y;   // exact: small integers
                    if (rect_reach(rc, rx0, ry0, rx0 + 7.0f, ry0 + 3.0f)) bits |= 1u << s;
                }
                bits &= ~blockdone;
                packed = idx | (bits << 16);
            }
            uint32_t m = __ballot_sync(FULL_MASK, (packed >> 16) != 0u);
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t pk = __shfl_sync(FULL_MASK, packed, j);
                const uint32_t idx = pk & 0xffffu, bj = pk >> 16;
                B200S_CHECK(idx < (uint32_t)BATCH * 16u && (idx & 15u) == 0u && bj < (1u << PX));
                const float4 A = *reinterpret_cast<const float4*>(qbase + idx);
                const float4 B = *reinterpret_cast<const float4*>(qbase + idx + BATCH * 16);
                const float4 Q = *reinterpret_cast<const float4*>(qbase + idx + 2 * BATCH * 16);
                float dx[NX], ax[NX], bx[NX], dy[NY], cy[NY];
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
                    if ((done >> s) & 1u) continue;
                    if (power > 0.0f || power < thrq) continue;
                    const float alpha = fminf(0.99f, fmul(B.y, FAST ? ex2_approx(power) : expf(power)));
                    if (alpha < ALPHA_MIN) continue;
                    const float test_T = fmul(T[s], fsub(1.0f, alpha));
                    if (test_T < 0.0001f) { done |= 1u << s; continue; }
                    Cr[s] = ffma(T[s], fmul(alpha, B.z), Cr[s]);
                    Cg[s] = ffma(T[s], fmul(alpha, B.w), Cg[s]);
                    Cb[s] = ffma(T[s], fmul(alpha, Q.x), Cb[s]);
                    T[s] = test_T;
                    last[s] = __float_as_uint(Q.z);
                }
            }
        }
#endif
    }
    cp_async_wait<0>();

    // epilogue
    const float bg0 = p.background[0], bg1 = p.background[1], bg2 = p.background[2];
    const size_t plane = (size_t)p.W * p.H;
    uint32_t lmax = 0;
#pragma unroll
    for (int s = 0; s < PX; ++s) {
        if ((outside >> s) & 1u) continue;
        const size_t pix = (size_t)__float2int_rn(pyf[s / NX]) * p.W + __float2int_rn(pxf[s % NX]);
        B200S_CHECK(pix < plane && last[s] <= (uint32_t)n);
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
    if (p.fast) blend_fwd_kernel<NW, (NW >= 8 ? 1 : BLEND_SPT), true><<<tiles, 32 * NW, 0, stream>>>(p);
    else blend_fwd_kernel<NW, (NW >= 8 ? 1 : BLEND_SPT), false><<<tiles, 32 * NW, 0, stream>>>(p);
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
