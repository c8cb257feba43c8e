// K2..K5 — binning: inclusive prefix sum of tiles_touched, key/value expansion, the sort and
// the per-tile ranges. Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys +
// cub::DeviceRadixSort::SortPairs + identifyTileRanges (reference rasterizer_impl.cu:70-138,277-317).
//
// The reference sorts N instances (N ~ 6 per visible Gaussian) on a 45-bit tile|depth key with a
// stable LSD radix sort, 6 passes over 12-byte pairs. All of this is integer work, so ANY
// procedure that yields "instances grouped by tile, inside a tile ordered by (depth bits,
// Gaussian index)" reproduces point_list / point_list_keys / ranges bit for bit. Here:
//
//  * scan_compact   one chained-look-back pass computes the reference's point_offsets (inclusive
//                   sum in Gaussian-index order; also num_rendered) AND compacts the visible
//                   Gaussians' (depth bits, id) in index order.
//  * depth sort     the P_v ~ N/6 visible Gaussians are sorted by depth with 4 stable onesweep
//                   passes. Compaction order + stability give (depth, index) order: exactly the
//                   order the reference's stable sort leaves inside every tile.
//  * expand         walks the Gaussians front to back and emits their instances (tile id, Gaussian
//                   id); block-level load-balanced (consecutive output slots -> consecutive
//                   threads, binary search in shared memory), so a Gaussian covering 2,000 tiles
//                   does not serialise one thread (the reference loops per thread,
//                   rasterizer_impl.cu:98-109). The prefix sum of tiles_touched in sorted order is
//                   fused in as a chained look-back between blocks, and so is the digit histogram
//                   of the tile ids.
//  * tile sort      a STABLE onesweep over the tile position only (2 passes for up to 65,536 tiles):
//                   the instances of a tile keep their emission (= depth) order. The key is row << 8 | column
//                   (grids up to 256 x 256 tiles; the tile id beyond): same order as the tile id, but both
//                   digits are tile coordinates, so their histograms are range-adds per Gaussian (sorted_scan)
//                   instead of shared-memory atomics per instance (expand), and a pass sees fewer distinct
//                   digits. Opt-in alternatives, bit-identical (tests/test_binning_modes_gpu.py): counting-rank
//                   passes without MATCH / VOTE, and the expansion fused into the first tile pass.
//  * ranges         as the reference.
//  Traffic: ~16 B x 2 passes per instance + 16 B x 4 passes per visible Gaussian, instead of
//  ~152 B per instance. The 64-bit keys are never materialised (the parity export rebuilds them).
//  * N (num_rendered) is read from device memory by every kernel: no host round trip is
//    needed between the stages (the reference blocks on a 4-byte D2H, rasterizer_impl.cu:281).

#include <cstdlib>
#include <atomic>
#include "kernels.cuh"

namespace b200s {

constexpr uint32_t FLAG_AGG = 1u, FLAG_INCL = 2u;
constexpr int SPIN_LIMIT = 1 << 22;

// Spins until the look-back word at `p` carries a flag; returns the word. A bounded spin:
// on timeout raises ERR_LOOKBACK_TIMEOUT and returns an "inclusive 0" word so that the
// grid always terminates.
__device__ __forceinline__ uint32_t wait_status(const uint32_t* p, B200sHeader* header) {
    uint32_t v = ld_volatile_u32(p);
    int spins = 0;
    while ((v & 3u) == 0u) {
        if (++spins > SPIN_LIMIT) {
            atomicOr(&header->error, (uint32_t)ERR_LOOKBACK_TIMEOUT);
            return FLAG_INCL;
        }
        __nanosleep(20);
        v = ld_volatile_u32(p);
    }
    return v;
}

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long wait_status64(const unsigned long long* p, B200sHeader* header) {
    unsigned long long v = ld_volatile_u64(p);
    int spins = 0;
    while ((v & 3ull) == 0ull) {
        if (++spins > SPIN_LIMIT) {
            atomicOr(&header->error, (uint32_t)ERR_LOOKBACK_TIMEOUT);
            return (unsigned long long)FLAG_INCL;
        }
        __nanosleep(20);
        v = ld_volatile_u64(p);
    }
    return v;
}


// Warp-cooperative decoupled look-back for a scan with ONE value per tile: the 32 lanes read the
// 32 predecessors' status words at once and consume them up to the first inclusive word / first
// word not yet published. Returns the exclusive prefix of `tile` (same value in every lane).
// Packed u64 words (two fields) are summed as they are: the caller guarantees no carry between fields.
template <typename T>
__device__ __forceinline__ T warp_lookback(const T* status, int64_t tile, B200sHeader* header) {
    const int lane = threadIdx.x & 31;
    T excl = 0;
    int64_t t = tile - 1;
    int spins = 0;
    while (t >= 0) {
        const int64_t idx = t - lane;
        T w = (T)FLAG_INCL;                       // virtual inclusive zero in front of tile 0
        if (idx >= 0) {
            if (sizeof(T) == 8) w = (T)ld_volatile_u64(reinterpret_cast<const unsigned long long*>(status + idx));
            else w = (T)ld_volatile_u32(reinterpret_cast<const uint32_t*>(status + idx));
        }
        const uint32_t flag = (uint32_t)(w & (T)3);
        const uint32_t nr = __ballot_sync(FULL_MASK, flag == 0u);
        const uint32_t inc = __ballot_sync(FULL_MASK, flag == FLAG_INCL);
        const int first_nr = nr ? __ffs(nr) - 1 : 32;
        const int first_inc = inc ? __ffs(inc) - 1 : 32;
        const bool done = first_inc < first_nr;
        const int take = done ? first_inc + 1 : first_nr;
        T v = lane < take ? (w >> 2) : (T)0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
        excl += v;
        if (done) break;
        t -= take;
        if (take == 0) {
            if (++spins > SPIN_LIMIT) {
                if (lane == 0) atomicOr(&header->error, (uint32_t)ERR_LOOKBACK_TIMEOUT);
                break;
            }
            __nanosleep(20);
        }
    }
    return excl;
}

// Per-thread look-back (one chain per radix digit) with LB independent loads in flight per step.
//
// Measured and rejected (round 2, profiles/r02j_*): a GROUPED look-back (one extra row per 16 tiles holding the
// group's sum / inclusive prefix, so that a tile walks <= 15 in-group words + one word per earlier group instead of
// the ~100-150 words of the flat walk). The protocol is correct (bit-exact suite green) but slower: C3 sort 0.278 ->
// 0.317 ms, C4 0.516 -> 0.608 ms. A group's row only appears when its last tile has reached its own look-back, so
// later tiles WAIT on the progress of their neighbours; the flat walk never waits for progress, only for counts that
// every resident tile publishes right after ranking.
#ifndef B200S_LB_DEPTH
#define B200S_LB_DEPTH 8
#endif
constexpr int LB_DEPTH = B200S_LB_DEPTH;
// The window depth is a template parameter, separately selectable for the depth-sort passes (B200S_LB_DEPTH_DS): all of
// their tiles (314 at C3) are resident at once, so a deeper window looked attractive there. Measured: it is not — C3
// sort 0.263 (8) / 0.270 (16) / 0.280 ms (32), C4 0.484 / 0.493 / 0.505 (profiles/r02r_lb_depth_ds.log). Both stay at 8.
#ifndef B200S_LB_DEPTH_DS
#define B200S_LB_DEPTH_DS 8
#endif
constexpr int LB_DEPTH_DS = B200S_LB_DEPTH_DS;
template <int DEPTH>
__device__ __forceinline__ uint32_t thread_lookback(const uint32_t* col, size_t stride, int64_t tile, B200sHeader* header) {
    uint32_t excl = 0;
    int64_t t = tile - 1;
    int spins = 0;
    while (t >= 0) {
        uint32_t w[DEPTH];
#pragma unroll
        for (int i = 0; i < DEPTH; ++i)
            w[i] = (t - i >= 0) ? ld_volatile_u32(col + (size_t)(t - i) * stride) : FLAG_INCL;
        bool done = false, stop = false;
        int take = 0;
#pragma unroll
        for (int i = 0; i < DEPTH; ++i) {
            if (!done && !stop) {
                const uint32_t f = w[i] & 3u;
                if (f == 0u) stop = true;
                else { excl += w[i] >> 2; ++take; if (f == FLAG_INCL) done = true; }
            }
        }
        if (done) break;
        t -= take;
        if (take == 0) {
            if (++spins > SPIN_LIMIT) { atomicOr(&header->error, (uint32_t)ERR_LOOKBACK_TIMEOUT); break; }
            __nanosleep(20);
        }
    }
    return excl;
}

// Publishes this tile's digit count (unless the caller already has), finds its exclusive prefix and publishes the
// inclusive one; `col` = the pass's look-back array + this thread's digit. Returns the exclusive prefix.
template <int DEPTH>
__device__ __forceinline__ uint32_t lookback_exchange(uint32_t* col, size_t stride, int64_t tile, uint32_t count,
                                                      bool already_published, B200sHeader* header) {
    uint32_t* mine = col + (size_t)tile * stride;
    if (!already_published) st_volatile_u32(mine, (count << 2) | (tile == 0 ? FLAG_INCL : FLAG_AGG));
    uint32_t excl = 0;
    if (tile != 0) {
        excl = thread_lookback<DEPTH>(col, stride, tile, header);
        st_volatile_u32(mine, ((excl + count) << 2) | FLAG_INCL);
    }
    return excl;
}

// ---------------------------------------------------------------------- scan + compaction
// Look-back word: bits [0,2) flag, [2,32) sum of tiles_touched, [32,64) number of visible.
__global__ void __launch_bounds__(SCAN_THREADS)
scan_compact_kernel(const uint32_t* __restrict__ in, const uint32_t* __restrict__ depth_bits,
                    uint32_t* __restrict__ out, uint32_t* __restrict__ vis_keys, uint32_t* __restrict__ vis_ids,
                    int P, unsigned long long* status, uint32_t* counters, B200sHeader* header) {
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_prefix;
    __shared__ unsigned long long s_warp[SCAN_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&counters[CTR_SCAN_TICKET], 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)tid * SCAN_ITEMS;

    uint32_t v[SCAN_ITEMS], dep[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= P) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
        const uint4* dsrc = reinterpret_cast<const uint4*>(depth_bits + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; ++i) {
            const uint4 t = src[i];
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
            const uint4 d = dsrc[i];
            dep[4 * i] = d.x; dep[4 * i + 1] = d.y; dep[4 * i + 2] = d.z; dep[4 * i + 3] = d.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i) {
            v[i] = (base + i < P) ? in[base + i] : 0u;
            dep[i] = (base + i < P) ? depth_bits[base + i] : 0u;
        }
    }
    uint32_t tsum = 0, vmask = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (v[i] != 0u) vmask |= 1u << i;
        tsum += v[i]; v[i] = tsum;                     // thread-inclusive
    }
    const unsigned long long mine = (unsigned long long)tsum | ((unsigned long long)__popc(vmask) << 32);

    unsigned long long incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned long long warp_excl = 0, total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const unsigned long long t = s_warp[w];
        if (w < warp) warp_excl += t;
        total += t;
    }
    if (warp == 0) {
        // packed word << 2: the tile sum (< 2^32) moves to bits [2,34), the visible count (< 2^30,
        // checked by the host) to [34,64)
        if (lane == 0) st_volatile_u64(&status[tile], (total << 2) | (tile == 0 ? FLAG_INCL : FLAG_AGG));
        unsigned long long excl = 0;
        if (tile != 0) {
            excl = warp_lookback<unsigned long long>(status, (int64_t)tile, header);
            if (lane == 0) st_volatile_u64(&status[tile], ((excl + total) << 2) | FLAG_INCL);
        }
        if (lane == 0) s_prefix = excl;
    }
    __syncthreads();
    const unsigned long long pre = s_prefix + warp_excl + (incl - mine);   // exclusive of this thread
    const uint32_t off = (uint32_t)pre;
    uint32_t ci = (uint32_t)(pre >> 32);
    if (base + SCAN_ITEMS <= P) {
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; ++i)
            dst[i] = make_uint4(v[4 * i] + off, v[4 * i + 1] + off, v[4 * i + 2] + off, v[4 * i + 3] + off);
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i)
            if (base + i < P) out[base + i] = v[i] + off;
    }
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if ((vmask >> i) & 1u) {
            vis_keys[ci] = dep[i];
            vis_ids[ci] = (uint32_t)(base + i);
            ++ci;
        }
    }
    // The thread that owns element P-1 publishes num_rendered and num_visible.
    if (base <= (int64_t)P - 1 && (int64_t)P - 1 < base + SCAN_ITEMS) {
        header->num_rendered = v[(P - 1) - base] + off;
        header->num_visible = ci;
    }
}

// ------------------------------------------------------------------------------- sort
// Digit histograms of all passes in one read of the keys (used for the depth sort; the tile
// sort's histogram is accumulated by the expansion kernel).
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ n_ptr, int passes,
                      uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_h[SORT_MAX_PASSES * SORT_RADIX];
    const uint32_t n = *n_ptr;
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += 256) s_h[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const uint32_t key = keys[i];
        for (int p = 0; p < passes; ++p)
            atomicAdd(&s_h[p * SORT_RADIX + ((key >> (p * SORT_RADIX_BITS)) & (SORT_RADIX - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += 256) {
        const uint32_t c = s_h[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// One block per pass: in-place exclusive scan of the 256 digit counts. With from_diff the input is a
// DIFFERENCE array (+c where a range of digits starts, -c behind its end; see sorted_scan_kernel): it is
// integrated into counts first (modulo 2^32, so the negative entries are exact).
__global__ void __launch_bounds__(SORT_RADIX)
sort_scan_hist_kernel(uint32_t* hist, int from_diff) {
    __shared__ uint32_t s_warp[2][SORT_RADIX / 32];
    uint32_t* h = hist + blockIdx.x * SORT_RADIX;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t c = h[tid];
    if (from_diff) {
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[1][warp] = incl;
        __syncthreads();
        uint32_t wex = 0;
        for (int w = 0; w < warp; ++w) wex += s_warp[1][w];
        c = wex + incl;
    }
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[0][warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
    for (int w = 0; w < warp; ++w) wex += s_warp[0][w];
    h[tid] = wex + incl - c;
}

#ifndef B200S_SORT_MINB
#define B200S_SORT_MINB 3
#endif
#ifndef B200S_RANK_COUNT_DEFAULT
#define B200S_RANK_COUNT_DEFAULT 0
#endif
#ifndef B200S_XY_KEYS_DEFAULT
#define B200S_XY_KEYS_DEFAULT 1
#endif
#ifndef B200S_FUSE_EXPAND_DEFAULT
#define B200S_FUSE_EXPAND_DEFAULT 0
#endif
constexpr int SORT_MINB = B200S_SORT_MINB;   // resident CTAs per SM requested from ptxas (register cap)
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_CNT_STRIDE = SORT_RADIX + 1;      // + sentinel digit for padding items
constexpr size_t SORT_SMEM = (size_t)SORT_TILE * 4 + (size_t)SORT_TILE * 4 +
                             (size_t)SORT_WARPS * SORT_CNT_STRIDE * 4 + (size_t)(SORT_RADIX + 1) * 4 * 2 + 64;

// One stable 8-bit pass over (u32 key, u32 value) pairs: chained look-back over SORT_TILE-key
// blocks ("onesweep"), in-block ranking by warp match. `n` is read from device memory; a count
// above `capacity` (binning overflow) makes the pass a no-op.
template <bool USE_MATCH, int LBD>
__global__ void __launch_bounds__(SORT_THREADS, SORT_MINB)
sort_onesweep_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                     B200sHeader* header, const uint32_t* __restrict__ n_ptr, int64_t capacity, int shift,
                     const uint32_t* __restrict__ bin_base, uint32_t* lookback, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* s_keys = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_vals = s_keys + SORT_TILE;
    uint32_t* s_cnt = s_vals + SORT_TILE;                         // [WARPS][257]
    uint32_t* s_dstart = s_cnt + SORT_WARPS * SORT_CNT_STRIDE;    // [257] tile-local digit start
    uint32_t* s_gbase = s_dstart + (SORT_RADIX + 1);              // [257] global dst - local pos
    uint32_t* s_misc = s_gbase + (SORT_RADIX + 1);                // [0] tile id, [1..8] warp sums

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t n = *n_ptr;
    if ((int64_t)n > capacity) n = 0;
    const uint32_t num_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (tid == 0) s_misc[0] = atomicAdd(ticket, 1u);
    for (int i = tid; i < SORT_WARPS * SORT_CNT_STRIDE; i += SORT_THREADS) s_cnt[i] = 0;
    __syncthreads();
    const uint32_t tile = s_misc[0];
    if (tile >= num_tiles) return;
    const uint32_t tile_base = tile * SORT_TILE;
    const uint32_t tile_n = min((uint32_t)SORT_TILE, n - tile_base);

    // warp-striped load: item k of lane l in warp w is element w*512 + k*32 + l of the tile,
    // so (k, lane) order inside a warp and warp order inside the tile equal memory order.
    uint32_t key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t dr[SORT_ITEMS];       // digit | rank << 16
    const uint32_t wbase = tile_base + warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t i = wbase + k * 32;
        const bool ok = i < n;
        key[k] = ok ? keys_in[i] : ~0u;
        val[k] = ok ? vals_in[i] : 0u;
        dr[k] = ok ? ((key[k] >> shift) & (SORT_RADIX - 1)) : (uint32_t)SORT_RADIX;
    }
    uint32_t* wc = s_cnt + warp * SORT_CNT_STRIDE;
    const uint32_t lt = lanemask_lt();
    {
        uint32_t peers[SORT_ITEMS];   // independent MATCHes issued back to back (long latency)
#pragma unroll
        for (int k = 0; k < SORT_ITEMS; ++k) peers[k] = USE_MATCH ? __match_any_sync(FULL_MASK, dr[k]) : digit_peers_ballot(dr[k]);
#pragma unroll
        for (int k = 0; k < SORT_ITEMS; ++k) {
            // every lane reads its digit's running count (same-address reads broadcast), then the
            // lowest peer lane alone bumps it: no FLO/BREV/SHFL on the critical path
            const uint32_t d = dr[k];
            const uint32_t before = peers[k] & lt;
            const uint32_t old = wc[d];
            __syncwarp();
            if (before == 0u) wc[d] = old + __popc(peers[k]);
            dr[k] = d | ((old + __popc(before)) << 16);
            __syncwarp();
        }
    }
    __syncthreads();

    // Thread d owns digit d: exclusive prefix over warps, tile count, chained look-back.
    uint32_t count = 0;
    {
        const int d = tid;   // SORT_THREADS == SORT_RADIX
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            const uint32_t t = s_cnt[w * SORT_CNT_STRIDE + d];
            s_cnt[w * SORT_CNT_STRIDE + d] = count;
            count += t;
        }
        if (tid == 0) {   // sentinel digit: padding items go after every real item
            uint32_t c = 0;
            for (int w = 0; w < SORT_WARPS; ++w) {
                const uint32_t t = s_cnt[w * SORT_CNT_STRIDE + SORT_RADIX];
                s_cnt[w * SORT_CNT_STRIDE + SORT_RADIX] = c;
                c += t;
            }
            s_dstart[SORT_RADIX] = tile_n;
        }
    }
    // block exclusive scan of `count` over digits
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_misc[1 + warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
    for (int w = 0; w < warp; ++w) wex += s_misc[1 + w];
    const uint32_t dstart = wex + incl - count;

    // Publish this tile's digit counts first, scatter locally while the predecessors make progress, and only then
    // look back: key / val / dr registers are dead by then, so the look-back can keep LB_DEPTH loads in flight.
    st_volatile_u32(lookback + (size_t)tile * SORT_RADIX + tid, (count << 2) | (tile == 0 ? FLAG_INCL : FLAG_AGG));
    s_dstart[tid] = dstart;
    __syncthreads();

    // local scatter into digit order
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t d = dr[k] & 0xffffu;
        const uint32_t pos = s_dstart[d] + wc[d] + (dr[k] >> 16);
        B200S_CHECK(d <= (uint32_t)SORT_RADIX && pos < (uint32_t)SORT_TILE);
        s_keys[pos] = key[k];
        s_vals[pos] = val[k];
    }
    const uint32_t excl = lookback_exchange<LBD>(lookback + tid, SORT_RADIX, (int64_t)tile, count, true, header);
    s_gbase[tid] = bin_base[tid] + excl - dstart;
    __syncthreads();
    for (uint32_t i = tid; i < tile_n; i += SORT_THREADS) {
        const uint32_t kk = s_keys[i];
        const uint32_t d = (kk >> shift) & (SORT_RADIX - 1);
        const uint32_t dst = s_gbase[d] + i;
        B200S_CHECK(dst < n);
        keys_out[dst] = kk;
        vals_out[dst] = s_vals[i];
    }
}

// ---------------------------------------------------------------- sort, counting rank
// The same stable 8-bit pass ranked WITHOUT match.any / vote. One MATCH costs ~70 SM cycles on the ADU
// pipe and a tile needs 128 of them, which bounds the pass above (ncu: ADU 62 %, issue slots 25 % busy).
// Here the tile is stably partitioned in shared memory by the two nibbles of the digit. Each round:
// every thread owns 16 CONSECUTIVE keys (blocked order), counts their nibbles in a private row of byte
// counters (remembering, per key, how many equal nibbles it had seen: the key's rank inside the thread),
// the counts are scanned in (nibble-major, thread-minor) order — shuffles inside a warp on 16-bit packed
// fields, 16 threads for the 8 x 16 warp totals — and the keys move to their ranks. After the second
// round the tile is in digit order; digit boundaries come from a binary search, and the chained
// look-back + global scatter are those of the kernel above. Only LDS / STS / SHFL / integer ALU.
constexpr int CR_STRIDE = SORT_ITEMS + 1;   // padded row of a thread's 16 keys: conflict-free blocked access
constexpr int CR_ROW = 9;                   // words of a thread's private counter / base row (odd: conflict-free)
constexpr size_t CR_ARRAY_BYTES = (size_t)SORT_THREADS * CR_STRIDE * 4 * 2 + (size_t)SORT_THREADS * CR_ROW * 4;
constexpr size_t CR_TAIL_BYTES = (size_t)SORT_WARPS * 8 * 4 * 2 + (size_t)(SORT_RADIX + 1) * 4 * 2 + 64;
constexpr size_t SORT_CR_SMEM = CR_ARRAY_BYTES + CR_TAIL_BYTES;
static_assert(SORT_ITEMS == 16 && SORT_THREADS == 256 && SORT_RADIX_BITS == 8, "counting-rank geometry");
static_assert(CR_ARRAY_BYTES % 16 == 0, "tail alignment");

struct CrSmem {
    uint32_t* keys;     // [256][17], logical position i at cr_phys(i)
    uint32_t* vals;
    uint32_t* row;      // [256][9] private rows
    uint32_t* wtot;     // [warps][8] warp totals, 2 nibble values per word
    uint32_t* comb;     // [warps][8] nibble start + warp prefix, packed
    uint32_t* dstart;   // [257] tile-local digit start
    uint32_t* gbase;    // [257] global dst - local pos
    uint32_t* misc;     // [0] tile id
};
__device__ __forceinline__ CrSmem cr_carve(unsigned char* arrays, unsigned char* tail) {
    CrSmem sm;
    sm.keys = reinterpret_cast<uint32_t*>(arrays);
    sm.vals = sm.keys + SORT_THREADS * CR_STRIDE;
    sm.row = sm.vals + SORT_THREADS * CR_STRIDE;
    sm.wtot = reinterpret_cast<uint32_t*>(tail);
    sm.comb = sm.wtot + SORT_WARPS * 8;
    sm.dstart = sm.comb + SORT_WARPS * 8;
    sm.gbase = sm.dstart + (SORT_RADIX + 1);
    sm.misc = sm.gbase + (SORT_RADIX + 1);
    return sm;
}

__device__ __forceinline__ uint32_t cr_phys(uint32_t i) { return i + (i >> 4); }

// Stable partition of the tile's 4,096 (key, value) pairs by the 8-bit digit at `shift`. Input: blocked order in
// registers (thread t holds tile positions [16t, 16t + 16)); result at cr_phys() positions of sm.keys / sm.vals.
// Ends with a __syncthreads().
__device__ __forceinline__ void cr_sort_tile(uint32_t (&key)[SORT_ITEMS], uint32_t (&val)[SORT_ITEMS], const CrSmem& sm,
                                             int shift) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* row = sm.row + tid * CR_ROW;
    unsigned char* row8 = reinterpret_cast<unsigned char*>(row);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int sh = shift + 4 * r;
        if (r == 1) {
#pragma unroll
            for (int k = 0; k < SORT_ITEMS; ++k) {
                key[k] = sm.keys[tid * CR_STRIDE + k];
                val[k] = sm.vals[tid * CR_STRIDE + k];
            }
        }
        row[0] = 0u; row[1] = 0u; row[2] = 0u; row[3] = 0u;
        uint32_t prev_lo = 0, prev_hi = 0;     // 4-bit fields: equal nibbles among this thread's earlier keys
#pragma unroll
        for (int k = 0; k < SORT_ITEMS; ++k) {
            const uint32_t v = (key[k] >> sh) & 15u;
            const uint32_t c = row8[v];
            row8[v] = (unsigned char)(c + 1u);
            if (k < 8) prev_lo |= c << (4 * k); else prev_hi |= c << (4 * (k - 8));
        }
        // own counts as 16-bit fields: word m holds nibble values 2m (low half) and 2m + 1 (high half)
        uint32_t W[8], own[8];
        {
            const uint32_t c0 = row[0], c1 = row[1], c2 = row[2], c3 = row[3];
            W[0] = __byte_perm(c0, 0u, 0x4140); W[1] = __byte_perm(c0, 0u, 0x4342);
            W[2] = __byte_perm(c1, 0u, 0x4140); W[3] = __byte_perm(c1, 0u, 0x4342);
            W[4] = __byte_perm(c2, 0u, 0x4140); W[5] = __byte_perm(c2, 0u, 0x4342);
            W[6] = __byte_perm(c3, 0u, 0x4140); W[7] = __byte_perm(c3, 0u, 0x4342);
        }
#pragma unroll
        for (int m = 0; m < 8; ++m) own[m] = W[m];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                const uint32_t t = __shfl_up_sync(FULL_MASK, W[m], o);
                if (lane >= o) W[m] += t;          // fields stay <= 512: no carry between them
            }
        }
        if (lane == 31) {
#pragma unroll
            for (int m = 0; m < 8; ++m) sm.wtot[warp * 8 + m] = W[m];
        }
        __syncthreads();
        if (tid < 16) {
            // thread v: prefix over warps of nibble value v, then the exclusive scan over the 16 values
            const uint32_t v = (uint32_t)tid;
            uint32_t run = 0, pre[SORT_WARPS];
#pragma unroll
            for (int w = 0; w < SORT_WARPS; ++w) {
                const uint32_t word = sm.wtot[w * 8 + (v >> 1)];
                pre[w] = run;
                run += (word >> ((v & 1u) * 16u)) & 0xffffu;
            }
            uint32_t incl = run;
#pragma unroll
            for (int o = 1; o < 16; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const uint32_t start = incl - run;
            unsigned short* comb16 = reinterpret_cast<unsigned short*>(sm.comb);
#pragma unroll
            for (int w = 0; w < SORT_WARPS; ++w) comb16[w * 16 + v] = (unsigned short)(start + pre[w]);   // <= 4096
        }
        __syncthreads();
        // base of (nibble value, this thread) = start + earlier warps + earlier lanes; packed, fields <= 4096
#pragma unroll
        for (int m = 0; m < 8; ++m) row[m] = sm.comb[warp * 8 + m] + W[m] - own[m];
#pragma unroll
        for (int k = 0; k < SORT_ITEMS; ++k) {
            const uint32_t v = (key[k] >> sh) & 15u;
            const uint32_t word = row[v >> 1];
            const uint32_t base = (word >> ((v & 1u) * 16u)) & 0xffffu;
            const uint32_t c = ((k < 8 ? prev_lo : prev_hi) >> (4 * (k & 7))) & 15u;
            const uint32_t pos = base + c;
            B200S_CHECK(pos < (uint32_t)SORT_TILE);
            const uint32_t ph = cr_phys(pos);
            sm.keys[ph] = key[k];
            sm.vals[ph] = val[k];
        }
        __syncthreads();
    }
}

// The sorted tile leaves for global memory: digit boundaries by binary search (thread d finds the first
// position whose digit is >= d), chained look-back per digit, coalesced scatter.
__device__ __forceinline__ void cr_scatter_tile(const CrSmem& sm, uint32_t tile, uint32_t tile_n, uint32_t n, int shift,
                                                const uint32_t* __restrict__ bin_base, uint32_t* lookback,
                                                B200sHeader* header, uint32_t* __restrict__ keys_out,
                                                uint32_t* __restrict__ vals_out) {
    const int tid = threadIdx.x;
    {
        const uint32_t d = (uint32_t)tid;
        uint32_t lo = 0, hi = tile_n;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t kd = (sm.keys[cr_phys(mid)] >> shift) & (SORT_RADIX - 1);
            if (kd < d) lo = mid + 1; else hi = mid;
        }
        sm.dstart[d] = lo;
        if (tid == 0) sm.dstart[SORT_RADIX] = tile_n;
    }
    __syncthreads();
    const uint32_t dstart = sm.dstart[tid];
    const uint32_t count = sm.dstart[tid + 1] - dstart;
    const uint32_t excl = lookback_exchange<LB_DEPTH>(lookback + tid, SORT_RADIX, (int64_t)tile, count, false, header);
    sm.gbase[tid] = bin_base[tid] + excl - dstart;
    __syncthreads();
    for (uint32_t i = tid; i < tile_n; i += SORT_THREADS) {
        const uint32_t ph = cr_phys(i);
        const uint32_t kk = sm.keys[ph];
        const uint32_t d = (kk >> shift) & (SORT_RADIX - 1);
        const uint32_t dst = sm.gbase[d] + i;
        B200S_CHECK(dst < n);
        keys_out[dst] = kk;
        vals_out[dst] = sm.vals[ph];
    }
}

__global__ void __launch_bounds__(SORT_THREADS, SORT_MINB)
sort_onesweep_count_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                           uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                           B200sHeader* header, const uint32_t* __restrict__ n_ptr, int64_t capacity, int shift,
                           const uint32_t* __restrict__ bin_base, uint32_t* lookback, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const CrSmem sm = cr_carve(smem_raw, smem_raw + CR_ARRAY_BYTES);
    const int tid = threadIdx.x;
    uint32_t n = *n_ptr;
    if ((int64_t)n > capacity) n = 0;
    const uint32_t num_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (tid == 0) sm.misc[0] = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = sm.misc[0];
    if (tile >= num_tiles) return;
    const uint32_t tile_base = tile * SORT_TILE;
    const uint32_t tile_n = min((uint32_t)SORT_TILE, n - tile_base);

    // blocked load: thread t owns tile positions [16t, 16t + 16). Padding items (key ~0: digit 255 in every
    // pass, behind every real item by stability) are never written out.
    uint32_t key[SORT_ITEMS], val[SORT_ITEMS];
    {
        const uint32_t i0 = tile_base + (uint32_t)tid * SORT_ITEMS;
#pragma unroll
        for (int q = 0; q < SORT_ITEMS / 4; ++q) {
            const uint32_t i = i0 + 4 * q;
            if (i + 4 <= n) {
                const uint4 kk = __ldg(reinterpret_cast<const uint4*>(keys_in + i));
                const uint4 vv = __ldg(reinterpret_cast<const uint4*>(vals_in + i));
                key[4 * q] = kk.x; key[4 * q + 1] = kk.y; key[4 * q + 2] = kk.z; key[4 * q + 3] = kk.w;
                val[4 * q] = vv.x; val[4 * q + 1] = vv.y; val[4 * q + 2] = vv.z; val[4 * q + 3] = vv.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const bool ok = i + j < n;
                    key[4 * q + j] = ok ? keys_in[i + j] : ~0u;
                    val[4 * q + j] = ok ? vals_in[i + j] : 0u;
                }
            }
        }
    }
    cr_sort_tile(key, val, sm, shift);
    cr_scatter_tile(sm, tile, tile_n, n, shift, bin_base, lookback, header, keys_out, vals_out);
}

// ----------------------------------------------------------------------------- expand
// (1) sorted_scan: inclusive prefix sum of tiles_touched in FRONT-TO-BACK order (chained look-back
//     between blocks of EXP_TILE Gaussians). Writes, per sorted Gaussian j, the end of its run of
//     instances and a dense copy of its tile rectangle, and records for every OUT_TILE-th output
//     slot which Gaussian's run contains it (block_first).
// (2) expand: every block emits exactly OUT_TILE consecutive instances, whatever the footprint of
//     the Gaussians involved. (Depth order puts the nearest — largest — Gaussians next to each
//     other, so a per-Gaussian partition would leave a few blocks with 100x the average work; the
//     reference loops per thread over a Gaussian's rectangle, rasterizer_impl.cu:98-109.) Each thread
//     owns OUT_ITEMS consecutive slots: one binary search, then a walk. The digit histograms of the
//     tile ids for the stable tile sort are accumulated here as well.
constexpr int OUT_ITEMS = 8;
constexpr int OUT_TILE = EXP_OUT_TILE;                 // 2,048 instances per block
static_assert(OUT_TILE == EXP_THREADS * OUT_ITEMS, "expansion geometry");

__global__ void __launch_bounds__(SS_THREADS)
sorted_scan_kernel(const uint32_t* __restrict__ sorted_ids, const SplatRec* __restrict__ rec,
                   uint32_t* __restrict__ run_end, uint32_t* __restrict__ run_rmin, uint32_t* __restrict__ run_rwh,
                   uint32_t* __restrict__ block_first, B200sHeader* header, int64_t capacity,
                   uint32_t* status, uint32_t* ticket, uint32_t* __restrict__ hist, int xy_keys) {
    __shared__ uint32_t s_warp[SS_THREADS / 32];
    __shared__ uint32_t s_block, s_prefix;
    __shared__ uint32_t s_diff[2 * SORT_RADIX];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_vis = header->num_visible;
    if ((int64_t)header->num_rendered > capacity) {
        if (blockIdx.x == 0 && tid == 0) header->overflow = 1u;
        return;
    }
    if (tid == 0) s_block = atomicAdd(ticket, 1u);
    if (xy_keys) for (int i = tid; i < 2 * SORT_RADIX; i += SS_THREADS) s_diff[i] = 0u;
    __syncthreads();
    const uint32_t block = s_block;
    const uint32_t g0 = block * EXP_TILE + tid * EXP_ITEMS;
    if (block * EXP_TILE >= n_vis) return;

    uint32_t cnt[EXP_ITEMS], rmin[EXP_ITEMS], rwh[EXP_ITEMS];
    uint32_t tsum = 0;
#pragma unroll
    for (int k = 0; k < EXP_ITEMS; ++k) {
        cnt[k] = 0; rmin[k] = 0; rwh[k] = 0;
        if (g0 + k < n_vis) {
            const float4 q2 = rec[sorted_ids[g0 + k]].q2;
            rmin[k] = __float_as_uint(q2.z); rwh[k] = __float_as_uint(q2.w);
            cnt[k] = (rwh[k] & 0xffffu) * (rwh[k] >> 16);
            if (xy_keys && cnt[k] != 0u) {
                // digit histograms of the (x, y) tile keys as range-adds: the w x h rectangle puts h instances on
                // every column digit x in [xmin, xmin + w) and w on every row digit y in [ymin, ymin + h):
                // 4 shared-memory atomics per Gaussian instead of 2 per instance in the expansion
                const uint32_t w = rwh[k] & 0xffffu, h = rwh[k] >> 16;
                const uint32_t x0 = rmin[k] & 0xffffu, y0 = rmin[k] >> 16;
                B200S_CHECK(x0 + w <= (uint32_t)SORT_RADIX && y0 + h <= (uint32_t)SORT_RADIX);
                atomicAdd(&s_diff[x0], h);
                if (x0 + w < (uint32_t)SORT_RADIX) atomicAdd(&s_diff[x0 + w], 0u - h);
                atomicAdd(&s_diff[SORT_RADIX + y0], w);
                if (y0 + h < (uint32_t)SORT_RADIX) atomicAdd(&s_diff[SORT_RADIX + y0 + h], 0u - w);
            }
        }
        tsum += cnt[k];
    }
    uint32_t incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t wex = 0, total = 0;
#pragma unroll
    for (int w = 0; w < SS_THREADS / 32; ++w) {
        const uint32_t t = s_warp[w];
        if (w < warp) wex += t;
        total += t;
    }
    if (warp == 0) {
        if (lane == 0) st_volatile_u32(&status[block], (total << 2) | (block == 0 ? FLAG_INCL : FLAG_AGG));
        uint32_t excl = 0;
        if (block != 0) {
            excl = warp_lookback<uint32_t>(status, (int64_t)block, header);
            if (lane == 0) st_volatile_u32(&status[block], ((excl + total) << 2) | FLAG_INCL);
        }
        if (lane == 0) s_prefix = excl;
    }
    __syncthreads();
    uint32_t run = s_prefix + wex + incl - tsum;     // global start of this thread's first run
#pragma unroll
    for (int k = 0; k < EXP_ITEMS; ++k) {
        if (g0 + k < n_vis) {
            const uint32_t start = run, end = run + cnt[k];
            run_end[g0 + k] = end;
            run_rmin[g0 + k] = rmin[k];
            run_rwh[g0 + k] = rwh[k];
            for (uint32_t m = (start + OUT_TILE - 1) / OUT_TILE; (uint64_t)m * OUT_TILE < end; ++m)
                block_first[m] = g0 + k;
            run = end;
        }
    }
    if (xy_keys) {
        // every thread's shared-memory atomics precede the __syncthreads() of the scan above
        for (int i = tid; i < 2 * SORT_RADIX; i += SS_THREADS) {
            const uint32_t c = s_diff[i];
            if (c) atomicAdd(&hist[i], c);
        }
    }
}

__global__ void __launch_bounds__(EXP_THREADS)
expand_kernel(const uint32_t* __restrict__ sorted_ids, const uint32_t* __restrict__ run_end,
              const uint32_t* __restrict__ run_rmin, const uint32_t* __restrict__ run_rwh,
              const uint32_t* __restrict__ block_first, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
              int grid_x, int passes, const B200sHeader* header, int64_t capacity, uint32_t* __restrict__ hist,
              int xy_keys) {
    __shared__ uint32_t s_end[OUT_TILE + 1], s_id[OUT_TILE + 1], s_rmin[OUT_TILE + 1], s_rwh[OUT_TILE + 1];
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
    const int tid = threadIdx.x;
    const uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) return;
    const uint32_t o0 = blockIdx.x * (uint32_t)OUT_TILE;
    if (o0 >= n) return;
    const uint32_t o1 = min(n, o0 + (uint32_t)OUT_TILE);
    for (int i = tid; i < passes * SORT_RADIX; i += EXP_THREADS) s_hist[i] = 0;
    // Gaussians whose runs intersect [o0, o1): j0 .. j1 (every run has >= 1 instance, so at most
    // OUT_TILE + 1 of them... the run containing o0 plus one per further slot)
    const uint32_t j0 = block_first[blockIdx.x];
    const uint32_t n_vis = header->num_visible;
    const uint32_t jmax = min(n_vis - 1, j0 + (uint32_t)OUT_TILE);
    for (uint32_t j = j0 + tid; j <= jmax; j += EXP_THREADS) {
        const uint32_t l = j - j0;
        s_end[l] = run_end[j];
        s_id[l] = sorted_ids[j];
        s_rmin[l] = run_rmin[j];
        s_rwh[l] = run_rwh[j];
    }
    __syncthreads();
    const int cntj = (int)(jmax - j0) + 1;
    uint32_t o = o0 + tid * OUT_ITEMS;
    if (o < o1) {
        // smallest l with s_end[l] > o
        int lo = 0, hi = cntj - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > o) hi = mid; else lo = mid + 1;
        }
        int l = lo;
        uint32_t end = s_end[l];
        uint32_t rwh = s_rwh[l], rmin = s_rmin[l], id = s_id[l];
        uint32_t w = rwh & 0xffffu;
        uint32_t k = o - (end - w * (rwh >> 16));
        uint32_t ty = k / w, tx = k - ty * w;
        uint32_t tk[OUT_ITEMS], tv[OUT_ITEMS];
        uint32_t hd = 0xffffffffu, hc = 0;     // run-length aggregation of the high digit
#pragma unroll
        for (int i = 0; i < OUT_ITEMS; ++i) {
            tk[i] = 0; tv[i] = 0;
            if (o + i < o1) {
                if (o + i >= end) {
                    ++l;
                    B200S_CHECK(l < cntj);
                    end = s_end[l]; rwh = s_rwh[l]; rmin = s_rmin[l]; id = s_id[l];
                    w = rwh & 0xffffu; tx = 0; ty = 0;
                }
                const uint32_t tile = ((rmin >> 16) + ty) * (uint32_t)grid_x + (rmin & 0xffffu) + tx;
                B200S_CHECK(w != 0 && ty < (rwh >> 16) && (int64_t)(o + i) < capacity &&
                            tile < (passes >= 4 ? 0xffffffffu : (1u << (passes * SORT_RADIX_BITS))));
                // xy_keys: key = row << 8 | column (same order as the tile id for grids up to 256 x 256); its digit
                // histograms were accumulated per Gaussian by sorted_scan_kernel
                tk[i] = xy_keys ? ((((rmin >> 16) + ty) << SORT_RADIX_BITS) | ((rmin & 0xffffu) + tx)) : tile;
                tv[i] = id;
                if (++tx == w) { tx = 0; ++ty; }
                if (!xy_keys) atomicAdd(&s_hist[tile & (SORT_RADIX - 1)], 1u);
                if (!xy_keys && passes > 1) {
                    const uint32_t hi_d = tile >> SORT_RADIX_BITS;     // all higher digits at once
                    if (hi_d != hd) {
                        if (hc) {
                            for (int p = 1; p < passes; ++p)
                                atomicAdd(&s_hist[p * SORT_RADIX + ((hd >> ((p - 1) * SORT_RADIX_BITS)) & (SORT_RADIX - 1))], hc);
                        }
                        hd = hi_d; hc = 0;
                    }
                    ++hc;
                }
            }
        }
        if (!xy_keys && hc) {
            for (int p = 1; p < passes; ++p)
                atomicAdd(&s_hist[p * SORT_RADIX + ((hd >> ((p - 1) * SORT_RADIX_BITS)) & (SORT_RADIX - 1))], hc);
        }
        if (o + OUT_ITEMS <= o1) {
            uint4* kd = reinterpret_cast<uint4*>(keys + o);
            uint4* vd = reinterpret_cast<uint4*>(vals + o);
#pragma unroll
            for (int i = 0; i < OUT_ITEMS / 4; ++i) {
                kd[i] = make_uint4(tk[4 * i], tk[4 * i + 1], tk[4 * i + 2], tk[4 * i + 3]);
                vd[i] = make_uint4(tv[4 * i], tv[4 * i + 1], tv[4 * i + 2], tv[4 * i + 3]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < OUT_ITEMS; ++i)
                if (o + i < o1) { keys[o + i] = tk[i]; vals[o + i] = tv[i]; }
        }
    }
    if (xy_keys) return;
    __syncthreads();
    for (int i = tid; i < passes * SORT_RADIX; i += EXP_THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// Expansion fused into the first tile pass ((x, y) keys + counting rank): the block GENERATES its 4,096 instances in
// blocked order (thread t: slots [16t, 16t + 16), one binary search then a walk, as expand_kernel does) and ranks
// them by column right away, so the unsorted instances never travel through HBM (16 B written + 16 B read per
// instance) and one launch disappears. The column histogram is already known (sorted_scan_kernel's range-adds).
constexpr int XP_RUNS = SORT_TILE + 1;                              // Gaussians whose runs can intersect one tile
constexpr size_t XP_RUN_BYTES = ((size_t)XP_RUNS * 4 + 15) / 16 * 16;
constexpr size_t XP_UNION_BYTES = 3 * XP_RUN_BYTES > CR_ARRAY_BYTES ? 3 * XP_RUN_BYTES : CR_ARRAY_BYTES;
constexpr size_t SORT_XP_SMEM = XP_RUN_BYTES + XP_UNION_BYTES + CR_TAIL_BYTES;
static_assert(SORT_TILE % OUT_TILE == 0, "block_first granularity");

__global__ void __launch_bounds__(SORT_THREADS, SORT_MINB)
expand_sort_count_kernel(const uint32_t* __restrict__ sorted_ids, const uint32_t* __restrict__ run_end,
                         const uint32_t* __restrict__ run_rmin, const uint32_t* __restrict__ run_rwh,
                         const uint32_t* __restrict__ block_first, uint32_t* __restrict__ keys_out,
                         uint32_t* __restrict__ vals_out, B200sHeader* header, int64_t capacity,
                         const uint32_t* __restrict__ bin_base, uint32_t* lookback, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* s_end = reinterpret_cast<uint32_t*>(smem_raw);
    unsigned char* uni = smem_raw + XP_RUN_BYTES;                   // run arrays first, the sort's arrays afterwards
    uint32_t* s_id = reinterpret_cast<uint32_t*>(uni);
    uint32_t* s_rmin = reinterpret_cast<uint32_t*>(uni + XP_RUN_BYTES);
    uint32_t* s_rwh = reinterpret_cast<uint32_t*>(uni + 2 * XP_RUN_BYTES);
    const CrSmem sm = cr_carve(uni, uni + XP_UNION_BYTES);
    const int tid = threadIdx.x;
    uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) n = 0;
    const uint32_t num_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (tid == 0) sm.misc[0] = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = sm.misc[0];
    if (tile >= num_tiles) return;
    const uint32_t o0 = tile * SORT_TILE;
    const uint32_t o1 = min(n, o0 + (uint32_t)SORT_TILE);
    const uint32_t tile_n = o1 - o0;

    const uint32_t j0 = block_first[tile * (SORT_TILE / OUT_TILE)];
    const uint32_t n_vis = header->num_visible;
    const uint32_t jmax = min(n_vis - 1, j0 + (uint32_t)SORT_TILE);
    for (uint32_t j = j0 + tid; j <= jmax; j += SORT_THREADS) {
        const uint32_t l = j - j0;
        s_end[l] = run_end[j];
        s_id[l] = sorted_ids[j];
        s_rmin[l] = run_rmin[j];
        s_rwh[l] = run_rwh[j];
    }
    __syncthreads();
    const int cntj = (int)(jmax - j0) + 1;
    uint32_t key[SORT_ITEMS], val[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) { key[i] = ~0u; val[i] = 0u; }
    const uint32_t o = o0 + (uint32_t)tid * SORT_ITEMS;
    if (o < o1) {
        int lo = 0, hi = cntj - 1;          // smallest l with s_end[l] > o
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > o) hi = mid; else lo = mid + 1;
        }
        int l = lo;
        uint32_t end = s_end[l];
        uint32_t rwh = s_rwh[l], rmin = s_rmin[l], id = s_id[l];
        uint32_t w = rwh & 0xffffu;
        const uint32_t k0 = o - (end - w * (rwh >> 16));
        uint32_t ty = k0 / w, tx = k0 - ty * w;
#pragma unroll
        for (int i = 0; i < SORT_ITEMS; ++i) {
            if (o + i < o1) {
                if (o + i >= end) {
                    ++l;
                    B200S_CHECK(l < cntj);
                    end = s_end[l]; rwh = s_rwh[l]; rmin = s_rmin[l]; id = s_id[l];
                    w = rwh & 0xffffu; tx = 0; ty = 0;
                }
                B200S_CHECK(w != 0 && ty < (rwh >> 16) && (rmin >> 16) + ty < (uint32_t)SORT_RADIX &&
                            (rmin & 0xffffu) + tx < (uint32_t)SORT_RADIX);
                key[i] = (((rmin >> 16) + ty) << SORT_RADIX_BITS) | ((rmin & 0xffffu) + tx);
                val[i] = id;
                if (++tx == w) { tx = 0; ++ty; }
            }
        }
    }
    __syncthreads();                                                // the run arrays are dead: their space is the sort's
    cr_sort_tile(key, val, sm, 0);
    cr_scatter_tile(sm, tile, tile_n, n, 0, bin_base, lookback, header, keys_out, vals_out);
}

// ----------------------------------------------------------------------------- ranges
// identifyTileRanges (rasterizer_impl.cu:116-138); `ranges` was cleared beforehand.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ keys, const B200sHeader* header, int64_t capacity,
                   uint2* __restrict__ ranges, int xy_keys, int grid_x) {
    uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) n = 0;
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        uint32_t cur = keys[i];
        if (xy_keys) cur = (cur >> SORT_RADIX_BITS) * (uint32_t)grid_x + (cur & (SORT_RADIX - 1));
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            uint32_t prev = keys[i - 1];
            if (xy_keys) prev = (prev >> SORT_RADIX_BITS) * (uint32_t)grid_x + (prev & (SORT_RADIX - 1));
            if (cur != prev) {
                ranges[prev].y = (uint32_t)i;
                ranges[cur].x = (uint32_t)i;
            }
        }
        if (i == (int64_t)n - 1) ranges[cur].y = n;
    }
}

// ----------------------------------------------------------------------------- tile launch order
// The blend kernels run one CTA per tile and a CTA lives as long as its list is long: with ~2 waves of CTAs per SM a
// long tile that starts late leaves SMs idle at the end (ncu, C3: 13 % of SM cycles with no resident warp in
// blend_fwd, 12 % in blend_bwd). Longest-processing-time-first: a single-block counting sort of the tile ids by list
// length (32-entry buckets, descending) gives the CTA launch order; results do not depend on it.
constexpr int TO_THREADS = 1024;
constexpr int TO_BUCKETS = 1024;
__device__ __forceinline__ void tile_order_block(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_last,
                                                 int tiles, uint32_t* __restrict__ order) {
    __shared__ uint32_t s_hist[TO_BUCKETS];
    __shared__ uint32_t s_warp[TO_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    s_hist[tid] = 0;
    __syncthreads();
    auto bucket = [&](int t) {
        const uint2 r = ranges[t];
        uint32_t len = r.y - r.x;
        if (tile_last != nullptr) len = min(len, tile_last[t]);
        return (int)min(len >> 5, (uint32_t)(TO_BUCKETS - 1));
    };
    for (int t = tid; t < tiles; t += TO_THREADS) atomicAdd(&s_hist[bucket(t)], 1u);
    __syncthreads();
    // thread tid owns bucket B-1-tid: exclusive prefix in that (descending length) order
    const uint32_t c = s_hist[TO_BUCKETS - 1 - tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
    for (int w = 0; w < warp; ++w) wex += s_warp[w];
    __syncthreads();
    s_hist[TO_BUCKETS - 1 - tid] = wex + incl - c;
    __syncthreads();
    for (int t = tid; t < tiles; t += TO_THREADS) {
        const uint32_t pos = atomicAdd(&s_hist[bucket(t)], 1u);
        B200S_CHECK(pos < (uint32_t)tiles);
        order[pos] = (uint32_t)t;
    }
}

__global__ void __launch_bounds__(TO_THREADS)
tile_order_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_last, int tiles,
                  uint32_t* __restrict__ order) {
    tile_order_block(ranges, tile_last, tiles, order);
}

// Backward prologue: the zero-fill of the raw-sum scratch (the reference's torch::zeros) with block 0 doing the tile
// ordering meanwhile, so the ordering costs no time of its own (the fill is HBM-bound and takes longer). In atomic
// mode only the rows of visible Gaussians are cleared (radii > 0: the only rows blend_bwd adds to and preprocess_bwd
// reads): 4 B read per Gaussian + 48 B written per visible one instead of 48 B per Gaussian.
__global__ void __launch_bounds__(TO_THREADS)
bwd_prologue_kernel(float4* __restrict__ zero, size_t n16, const int* __restrict__ radii, int P,
                    const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_last, int tiles,
                    uint32_t* __restrict__ order) {
    if (order != nullptr) {
        if (blockIdx.x == 0) { tile_order_block(ranges, tile_last, tiles, order); return; }
    }
    const size_t first = order != nullptr ? 1 : 0;
    const size_t stride = (size_t)(gridDim.x - first) * TO_THREADS;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (radii != nullptr) {
        static_assert(ACC_STRIDE == 12, "three float4 per row");
        for (size_t i = (size_t)(blockIdx.x - first) * TO_THREADS + threadIdx.x; i < (size_t)P; i += stride) {
            if (radii[i] > 0) { zero[3 * i] = z; zero[3 * i + 1] = z; zero[3 * i + 2] = z; }
        }
        return;
    }
    for (size_t i = (size_t)(blockIdx.x - first) * TO_THREADS + threadIdx.x; i < n16; i += stride) zero[i] = z;
}

bool tile_order_enabled() {
    static int on = [] { const char* v = getenv("B200S_TILE_ORDER"); return v && *v ? atoi(v) : 1; }();
    return on != 0;
}

bool tile_order_fwd_enabled() {
    static int on = [] { const char* v = getenv("B200S_TILE_ORDER"); return v && *v ? atoi(v) : 1; }();
    return on >= 2;
}

int launch_tile_order(const uint2* ranges, const uint32_t* tile_last, int tiles, uint32_t* order, cudaStream_t stream) {
    if (tiles <= 0) return 0;
    count_launch();
    tile_order_kernel<<<1, TO_THREADS, 0, stream>>>(ranges, tile_last, tiles, order);
    return (int)cudaGetLastError();
}

int launch_bwd_prologue(void* zero, size_t zero_bytes, const int* radii, int P, const uint2* ranges,
                        const uint32_t* tile_last, int tiles, uint32_t* order, cudaStream_t stream) {
    // zero_bytes is a multiple of 16 and `zero` 256-byte aligned (workspace carving); radii != NULL: rows of
    // ACC_STRIDE floats indexed by Gaussian, only the visible ones are cleared
    count_launch();
    // two fill blocks per SM of the current device (148 SMs on B200), asked for once per device
    static std::atomic<int> sms_of[64];
    int dev = 0;
    B200S_CUDA_OK(cudaGetDevice(&dev));
    int sms = dev >= 0 && dev < 64 ? sms_of[dev].load(std::memory_order_relaxed) : 0;
    if (sms <= 0) {
        B200S_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        if (sms <= 0) sms = 148;
        if (dev >= 0 && dev < 64) sms_of[dev].store(sms, std::memory_order_relaxed);
    }
    const int blocks = sms * 2 + (order != nullptr ? 1 : 0);
    bwd_prologue_kernel<<<blocks, TO_THREADS, 0, stream>>>(reinterpret_cast<float4*>(zero), zero_bytes / 16, radii, P,
                                                           ranges, tile_last, tiles, order);
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------ host launchers
// Ranking primitive per pass: match.any costs grow with the number of DISTINCT digits in a warp, the
// ballot form is flat. B200S_RANK_MATCH = bit mask over (which * 4 + pass), which: 0 depth, 1 tile.
static bool rank_by_match(int which, int pass) {
    static int mask = [] { const char* v = getenv("B200S_RANK_MATCH"); return v && *v ? (int)strtol(v, nullptr, 0) : 0xff; }();
    return (mask >> (which * 4 + pass)) & 1;
}
// B200S_RANK_COUNT = bit mask over (which * 4 + pass): passes ranked by the counting kernel (no MATCH / VOTE).
static bool rank_by_count(int which, int pass) {
    static int mask = [] { const char* v = getenv("B200S_RANK_COUNT"); return v && *v ? (int)strtol(v, nullptr, 0) : B200S_RANK_COUNT_DEFAULT; }();
    return (mask >> (which * 4 + pass)) & 1;
}
static int set_sort_attr() {
    B200S_SET_FUNC_ATTR_ONCE((sort_onesweep_kernel<true, LB_DEPTH>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM);
    B200S_SET_FUNC_ATTR_ONCE((sort_onesweep_kernel<false, LB_DEPTH>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM);
    if (LB_DEPTH_DS != LB_DEPTH) {
        B200S_SET_FUNC_ATTR_ONCE((sort_onesweep_kernel<true, LB_DEPTH_DS>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM);
        B200S_SET_FUNC_ATTR_ONCE((sort_onesweep_kernel<false, LB_DEPTH_DS>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM);
    }
    B200S_SET_FUNC_ATTR_ONCE(sort_onesweep_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_CR_SMEM);
    B200S_SET_FUNC_ATTR_ONCE(expand_sort_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_XP_SMEM);
    return 0;
}
using SortPassFn = void (*)(const uint32_t*, const uint32_t*, uint32_t*, uint32_t*, B200sHeader*, const uint32_t*, int64_t,
                            int, const uint32_t*, uint32_t*, uint32_t*);
static void sort_pass_kernel(int which, int pass, SortPassFn* fn, size_t* smem) {
    if (rank_by_count(which, pass)) { *fn = sort_onesweep_count_kernel; *smem = SORT_CR_SMEM; }
    else if (which == 0) {
        *fn = rank_by_match(which, pass) ? sort_onesweep_kernel<true, LB_DEPTH_DS> : sort_onesweep_kernel<false, LB_DEPTH_DS>;
        *smem = SORT_SMEM;
    } else {
        *fn = rank_by_match(which, pass) ? sort_onesweep_kernel<true, LB_DEPTH> : sort_onesweep_kernel<false, LB_DEPTH>;
        *smem = SORT_SMEM;
    }
}

// Stage 1b: point_offsets + num_rendered, compaction of the visible Gaussians, depth sort.
// Leaves the front-to-back Gaussian ids in g.vis_ids[0] (DEPTH_PASSES is even).
int launch_scan_and_depth_sort(int P, const GeomWS& g, bool debug, cudaStream_t stream) {
    if (P <= 0) return 0;
    { const int rc_attr = set_sort_attr(); if (rc_attr) return rc_attr; }
    {
        ProfScope ps(B200S_STAGE_SCAN, stream);
        const int tiles = (P + SCAN_TILE - 1) / SCAN_TILE;
        count_launch();
        scan_compact_kernel<<<tiles, SCAN_THREADS, 0, stream>>>(g.tiles_touched, g.depth_bits, g.point_offsets,
                                                               g.vis_keys[0], g.vis_ids[0], P, g.scan_status,
                                                               g.counters, g.header);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    {
        ProfScope ps(B200S_STAGE_SORT, stream);
        const uint32_t* n_ptr = &g.header->num_visible;
        int hb = (P + 256 * 16 - 1) / (256 * 16);
        if (hb > 148 * 8) hb = 148 * 8;
        count_launch(2);
        sort_histogram_kernel<<<hb, 256, 0, stream>>>(g.vis_keys[0], n_ptr, DEPTH_PASSES, g.depth_hist);
        B200S_AFTER_LAUNCH(stream, debug);
        sort_scan_hist_kernel<<<DEPTH_PASSES, SORT_RADIX, 0, stream>>>(g.depth_hist, 0);
        B200S_AFTER_LAUNCH(stream, debug);
        int cur = 0;
        for (int p = 0; p < DEPTH_PASSES; ++p) {
            count_launch();
            SortPassFn kern; size_t smem;
            sort_pass_kernel(0, p, &kern, &smem);
            kern<<<(unsigned)g.depth_tiles, SORT_THREADS, smem, stream>>>(
                g.vis_keys[cur], g.vis_ids[cur], g.vis_keys[cur ^ 1], g.vis_ids[cur ^ 1], g.header, n_ptr,
                (int64_t)P, p * SORT_RADIX_BITS, g.depth_hist + p * SORT_RADIX,
                g.depth_lookback + (size_t)p * g.depth_tiles * SORT_RADIX, g.counters + CTR_DEPTH_TICKET0 + p);
            B200S_AFTER_LAUNCH(stream, debug);
            cur ^= 1;
        }
    }
    return 0;
}

// getHigherMsb (rasterizer_impl.cu:35-50): number of bits needed for tile ids.
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// Number of onesweep passes over the TILE bits; the sorted run lives in ping-pong buffer (passes & 1).
int sort_num_passes(int tiles) {
    const int bits = (int)higher_msb((uint32_t)tiles);
    const int p = (bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
    return p < 1 ? 1 : p;
}

// Stage 2: expand + tile sort + ranges. `n_hint` bounds the number of instances for grid sizing
// (capacity when the host does not know N). Returns in *final_buf which ping-pong buffer
// holds the sorted result.
int launch_binning(int P, const GeomWS& g, const BinWS& bw, int64_t capacity, int64_t n_hint, int grid_x,
                   int tiles, uint2* ranges, int* final_buf, bool debug, cudaStream_t stream) {
    { const int rc_attr = set_sort_attr(); if (rc_attr) return rc_attr; }
    const int passes = sort_num_passes(tiles);
    // (x, y) keys: row << 8 | column instead of the tile id, when that is still a 2-pass sort. Same order, but the
    // two digits are the tile coordinates themselves, so their histograms are range-adds per Gaussian.
    static const int xy_env = [] { const char* v = getenv("B200S_XY_KEYS"); return v && *v ? atoi(v) : B200S_XY_KEYS_DEFAULT; }();
    const int xy_keys = (xy_env && passes == 2 && grid_x <= SORT_RADIX && tiles / grid_x <= SORT_RADIX) ? 1 : 0;
    static const int fuse_env = [] { const char* v = getenv("B200S_FUSE_EXPAND"); return v && *v ? atoi(v) : B200S_FUSE_EXPAND_DEFAULT; }();
    const bool fuse_expand = xy_keys && fuse_env && rank_by_count(1, 0);
    if (n_hint > capacity) n_hint = capacity;
    if (n_hint < 1) n_hint = 1;
    {
        ProfScope ps(B200S_STAGE_MISC, stream);
        B200S_CUDA_OK(cudaMemsetAsync(bw.hist, 0, bw.zero_bytes, stream));
        B200S_CUDA_OK(cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, stream));
    }
    if (P > 0) {
        // the depth sort's spare buffers (and its no longer needed sorted keys) hold the runs
        ProfScope ps(B200S_STAGE_EXPAND, stream);
        count_launch(2);
        sorted_scan_kernel<<<(P + EXP_TILE - 1) / EXP_TILE, SS_THREADS, 0, stream>>>(
            g.vis_ids[0], g.rec, g.vis_keys[1], g.vis_ids[1], g.vis_keys[0], bw.block_first, g.header, capacity,
            g.exp_status, g.counters + CTR_EXPAND_TICKET, bw.hist, xy_keys);
        B200S_AFTER_LAUNCH(stream, debug);
        if (!fuse_expand) {
            expand_kernel<<<(unsigned)((n_hint + OUT_TILE - 1) / OUT_TILE), EXP_THREADS, 0, stream>>>(
                g.vis_ids[0], g.vis_keys[1], g.vis_ids[1], g.vis_keys[0], bw.block_first, bw.keys[0], bw.vals[0],
                grid_x, passes, g.header, capacity, bw.hist, xy_keys);
            B200S_AFTER_LAUNCH(stream, debug);
        }
    }
    const int64_t sort_tiles = (n_hint + SORT_TILE - 1) / SORT_TILE;
    int cur = 0;
    {
        ProfScope ps(B200S_STAGE_SORT, stream);
        count_launch();
        sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, stream>>>(bw.hist, xy_keys);
        B200S_AFTER_LAUNCH(stream, debug);
        for (int p = 0; p < passes; ++p) {
            count_launch();
            if (p == 0 && fuse_expand && P > 0) {
                expand_sort_count_kernel<<<(unsigned)sort_tiles, SORT_THREADS, SORT_XP_SMEM, stream>>>(
                    g.vis_ids[0], g.vis_keys[1], g.vis_ids[1], g.vis_keys[0], bw.block_first, bw.keys[1], bw.vals[1],
                    g.header, capacity, bw.hist, bw.lookback, g.counters + CTR_SORT_TICKET0);
                B200S_AFTER_LAUNCH(stream, debug);
                cur ^= 1;
                continue;
            }
            SortPassFn kern; size_t smem;
            sort_pass_kernel(1, p, &kern, &smem);
            kern<<<(unsigned)sort_tiles, SORT_THREADS, smem, stream>>>(
                bw.keys[cur], bw.vals[cur], bw.keys[cur ^ 1], bw.vals[cur ^ 1], g.header, &g.header->num_rendered,
                capacity, p * SORT_RADIX_BITS, bw.hist + p * SORT_RADIX,
                bw.lookback + (size_t)p * bw.sort_tiles * SORT_RADIX, g.counters + CTR_SORT_TICKET0 + p);
            B200S_AFTER_LAUNCH(stream, debug);
            cur ^= 1;
        }
    }
    *final_buf = cur;
    {
        ProfScope ps(B200S_STAGE_RANGES, stream);
        count_launch();
        int rb = (int)((n_hint + 255) / 256);
        if (rb > 148 * 16) rb = 148 * 16;
        tile_ranges_kernel<<<rb, 256, 0, stream>>>(bw.keys[cur], g.header, capacity, ranges, xy_keys, grid_x);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    return 0;
}

}  // namespace b200s
