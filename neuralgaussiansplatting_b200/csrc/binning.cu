// K2..K5 — binning: inclusive prefix sum of tiles_touched, key/value expansion,
// stable LSD radix sort of (u64 tile|depth key, u32 Gaussian id), per-tile ranges.
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs
// + identifyTileRanges (reference rasterizer_impl.cu:70-138,277-317).
//
// Everything here is integer work: any correct inclusive sum and any correct STABLE sort on
// bits [0, 32+bit) reproduce the reference's point_offsets / point_list_keys / point_list /
// ranges bit for bit.
//
//  * scan:    single-pass chained scan with decoupled look-back (one launch, 8 B/Gaussian).
//  * expand:  block-level load-balanced expansion — each block turns its 256 Gaussians'
//             counts into one contiguous run of instances and assigns consecutive output
//             slots to consecutive threads (binary search in shared memory), so a Gaussian
//             covering 2,000 tiles no longer serialises one thread (the reference loops
//             per thread, rasterizer_impl.cu:98-109) and all 12-B stores coalesce.
//  * sort:    two levels. (1) onesweep over the TILE bits only (bits [32, 32+bit), two 8-bit
//             passes for up to 65,536 tiles): one histogram pass, then one read+write pass per
//             digit with a chained look-back over 4,096-key blocks; in-block ranking by warp
//             match, stable by construction. (2) every tile's run — now contiguous, in
//             insertion (= Gaussian index) order — is sorted on the 32 depth bits by ONE CTA
//             entirely in shared memory (4 stable 8-bit passes, no HBM traffic besides one read
//             of the run and one 4-byte write per instance). Runs longer than the shared-memory
//             capacity take the same stable passes through global memory. The result is
//             bit-identical to the reference's single 6-pass sort of the 45-bit key
//             (any stable sort on those bits is), at ~64 instead of ~152 B/instance.
//  * N (num_rendered) is read from device memory by every kernel: no host round trip is
//    needed between the stages (the reference blocks on a 4-byte D2H, rasterizer_impl.cu:281).

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

// ------------------------------------------------------------------------------- scan
__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int P,
                  uint32_t* status, uint32_t* counters, B200sHeader* header) {
    __shared__ uint32_t s_tile, s_prefix;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&counters[CTR_SCAN_TICKET], 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)tid * SCAN_ITEMS;

    uint32_t v[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= P) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; ++i) {
            const uint4 t = src[i];
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < P) ? in[base + i] : 0u;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) { tsum += v[i]; v[i] = tsum; }   // thread-inclusive

    uint32_t incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_excl = 0, total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const uint32_t t = s_warp[w];
        if (w < warp) warp_excl += t;
        total += t;
    }
    if (tid == 0) {
        uint32_t excl = 0;
        if (tile == 0) {
            st_volatile_u32(&status[0], (total << 2) | FLAG_INCL);
        } else {
            st_volatile_u32(&status[tile], (total << 2) | FLAG_AGG);
            int64_t t = (int64_t)tile - 1;
            while (true) {
                const uint32_t w = wait_status(&status[t], header);
                excl += w >> 2;
                if ((w & 3u) == FLAG_INCL || t == 0) break;
                --t;
            }
            st_volatile_u32(&status[tile], ((excl + total) << 2) | FLAG_INCL);
        }
        s_prefix = excl;
    }
    __syncthreads();
    const uint32_t off = s_prefix + warp_excl + (incl - tsum);
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
    // The thread that owns element P-1 publishes num_rendered.
    if (base <= (int64_t)P - 1 && (int64_t)P - 1 < base + SCAN_ITEMS) {
        header->num_rendered = v[(P - 1) - base] + off;
        header->num_visible = ld_volatile_u32(&counters[CTR_VISIBLE]);
    }
}

// ----------------------------------------------------------------------------- expand
__global__ void __launch_bounds__(256)
expand_keys_kernel(int P, const SplatRec* __restrict__ rec, const uint32_t* __restrict__ offsets,
                   uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, int grid_x,
                   B200sHeader* header, int64_t capacity) {
    __shared__ uint32_t s_end[256];
    __shared__ uint32_t s_depth[256], s_rmin[256], s_w[256];
    __shared__ uint32_t s_start0;
    const uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) {
        if (blockIdx.x == 0 && threadIdx.x == 0) header->overflow = 1u;
        return;
    }
    const int tid = threadIdx.x;
    const int64_t g0 = (int64_t)blockIdx.x * 256;
    const int64_t g = g0 + tid;
    const int64_t gc = g < P ? g : (int64_t)P - 1;
    const uint32_t end = offsets[gc];
    const uint32_t start = (gc == 0) ? 0u : offsets[gc - 1];
    const bool has = (g < P) && end > start;
    s_end[tid] = end;
    if (tid == 0) s_start0 = start;
    if (has) {
        const float4 q2 = rec[g].q2;
        s_depth[tid] = __float_as_uint(q2.y);
        s_rmin[tid] = __float_as_uint(q2.z);
        s_w[tid] = __float_as_uint(q2.w) & 0xffffu;
    }
    __syncthreads();
    const uint32_t block_start = s_start0;
    const uint32_t total = s_end[255] - block_start;
    for (uint32_t o = tid; o < total; o += 256) {
        const uint32_t target = block_start + o;
        // smallest j with s_end[j] > target
        int lo = 0, hi = 255;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > target) hi = mid; else lo = mid + 1;
        }
        const int j = lo;
        const uint32_t first = (j == 0) ? block_start : s_end[j - 1];
        const uint32_t k = target - first;
        const uint32_t w = s_w[j];
        const uint32_t ty = k / w;
        const uint32_t tx = k - ty * w;
        const uint32_t rmin = s_rmin[j];
        const uint32_t tile = ((rmin >> 16) + ty) * (uint32_t)grid_x + (rmin & 0xffffu) + tx;
        keys[target] = ((uint64_t)tile << 32) | (uint64_t)s_depth[j];
        vals[target] = (uint32_t)(g0 + j);
    }
}

// ------------------------------------------------------------------------------- sort
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const uint64_t* __restrict__ keys, const B200sHeader* header, int64_t capacity,
                      int first_shift, int passes, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_h[SORT_MAX_PASSES * SORT_RADIX];
    uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) n = 0;
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += 256) s_h[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const uint64_t key = keys[i];
        for (int p = 0; p < passes; ++p)
            atomicAdd(&s_h[p * SORT_RADIX + (uint32_t)((key >> (first_shift + p * SORT_RADIX_BITS)) & (SORT_RADIX - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += 256) {
        const uint32_t c = s_h[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// One block per pass: in-place exclusive scan of the 256 digit counts.
__global__ void __launch_bounds__(SORT_RADIX)
sort_scan_hist_kernel(uint32_t* hist) {
    __shared__ uint32_t s_warp[SORT_RADIX / 32];
    uint32_t* h = hist + blockIdx.x * SORT_RADIX;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t c = h[tid];
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
    for (int w = 0; w < warp; ++w) wex += s_warp[w];
    h[tid] = wex + incl - c;
}

constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_CNT_STRIDE = SORT_RADIX + 1;      // + sentinel digit for padding items
constexpr size_t SORT_SMEM = (size_t)SORT_TILE * 8 + (size_t)SORT_TILE * 4 +
                             (size_t)SORT_WARPS * SORT_CNT_STRIDE * 4 + (size_t)(SORT_RADIX + 1) * 4 * 2 + 64;

__global__ void __launch_bounds__(SORT_THREADS)
sort_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                     B200sHeader* header, int64_t capacity, int shift,
                     const uint32_t* __restrict__ bin_base, uint32_t* lookback, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem_raw + (size_t)SORT_TILE * 8);
    uint32_t* s_cnt = s_vals + SORT_TILE;                         // [WARPS][257]
    uint32_t* s_dstart = s_cnt + SORT_WARPS * SORT_CNT_STRIDE;    // [257] tile-local digit start
    uint32_t* s_gbase = s_dstart + (SORT_RADIX + 1);              // [257] global dst - local pos
    uint32_t* s_misc = s_gbase + (SORT_RADIX + 1);                // [0] tile id, [1..8] warp sums

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t n = header->num_rendered;
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
    uint64_t key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t dr[SORT_ITEMS];       // digit | rank << 16
    const uint32_t wbase = tile_base + warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t i = wbase + k * 32;
        const bool ok = i < n;
        key[k] = ok ? keys_in[i] : ~0ull;
        val[k] = ok ? vals_in[i] : 0u;
        dr[k] = ok ? (uint32_t)((key[k] >> shift) & (SORT_RADIX - 1)) : (uint32_t)SORT_RADIX;
    }
    uint32_t* wc = s_cnt + warp * SORT_CNT_STRIDE;
    const uint32_t lt = lanemask_lt();
    {
        uint32_t peers[SORT_ITEMS];   // independent MATCHes issued back to back (long latency)
#pragma unroll
        for (int k = 0; k < SORT_ITEMS; ++k) peers[k] = digit_peers(dr[k]);
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

    uint32_t excl = 0;
    {
        uint32_t* lb = lookback + (size_t)tile * SORT_RADIX + tid;
        if (tile == 0) {
            st_volatile_u32(lb, (count << 2) | FLAG_INCL);
        } else {
            st_volatile_u32(lb, (count << 2) | FLAG_AGG);
            int64_t t = (int64_t)tile - 1;
            while (true) {
                const uint32_t w = wait_status(lookback + (size_t)t * SORT_RADIX + tid, header);
                excl += w >> 2;
                if ((w & 3u) == FLAG_INCL || t == 0) break;
                --t;
            }
            st_volatile_u32(lb, ((excl + count) << 2) | FLAG_INCL);
        }
    }
    s_dstart[tid] = dstart;
    s_gbase[tid] = bin_base[tid] + excl - dstart;
    __syncthreads();

    // local scatter into digit order
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t d = dr[k] & 0xffffu;
        const uint32_t pos = s_dstart[d] + wc[d] + (dr[k] >> 16);
        s_keys[pos] = key[k];
        s_vals[pos] = val[k];
    }
    __syncthreads();
    for (uint32_t i = tid; i < tile_n; i += SORT_THREADS) {
        const uint64_t kk = s_keys[i];
        const uint32_t d = (uint32_t)((kk >> shift) & (SORT_RADIX - 1));
        const uint32_t dst = s_gbase[d] + i;
        keys_out[dst] = kk;
        vals_out[dst] = s_vals[i];
    }
}

// ----------------------------------------------------------------------------- ranges
// identifyTileRanges (rasterizer_impl.cu:116-138); `ranges` was cleared beforehand.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint64_t* __restrict__ keys, const B200sHeader* header, int64_t capacity,
                   uint2* __restrict__ ranges) {
    uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) n = 0;
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) {
                ranges[prev].y = (uint32_t)i;
                ranges[cur].x = (uint32_t)i;
            }
        }
        if (i == (int64_t)n - 1) ranges[cur].y = n;
    }
}

// -------------------------------------------------------------------- per-tile depth sort
// One CTA per tile sorts the tile's run (contiguous after the tile-bit passes, in insertion
// order) on the 32 depth bits with 4 stable 8-bit LSD passes.
constexpr int TS_THREADS = 256;
constexpr int TS_WARPS = TS_THREADS / 32;
constexpr int TS_ITEMS = 16;
constexpr int TS_CAP = TS_THREADS * TS_ITEMS;          // 4,096 entries sorted entirely in shared memory
constexpr int TS_CNT_STRIDE = 257;                     // + sentinel digit for padding lanes
constexpr size_t TS_SMEM = (size_t)TS_CAP * 4 * 4 + (size_t)TS_WARPS * TS_CNT_STRIDE * 4 + 257 * 4 + 64;

// Stable ranking of up to TS_ITEMS items per thread (warp-striped: item k of lane l of warp w is
// element w*chunk + k*32 + l) by the 8-bit digit; leaves digit | rank << 16 in dr[] and the
// per-warp digit counts in s_cnt[warp][*].
__device__ __forceinline__ void ts_rank(uint32_t (&dr)[TS_ITEMS], int kmax, uint32_t* wc, uint32_t lt) {
    // all matches first: MATCH has a long latency (the dominant stall of this kernel when each one
    // was consumed immediately), and they are independent of the counter updates below
    uint32_t peers[TS_ITEMS];
#pragma unroll
    for (int k = 0; k < TS_ITEMS; ++k) peers[k] = (k < kmax) ? digit_peers(dr[k]) : 0u;
#pragma unroll
    for (int k = 0; k < TS_ITEMS; ++k) {
        if (k < kmax) {
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
}

// Thread d (< 256): exclusive prefix of digit d over the warps (written back to s_cnt) and the
// digit's total; then the block-exclusive scan over digits -> returns this digit's start.
__device__ __forceinline__ uint32_t ts_digit_scan(uint32_t* s_cnt, uint32_t* s_wsum, uint32_t& count) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    count = 0;
#pragma unroll
    for (int w = 0; w < TS_WARPS; ++w) {
        const uint32_t t = s_cnt[w * TS_CNT_STRIDE + tid];
        s_cnt[w * TS_CNT_STRIDE + tid] = count;
        count += t;
    }
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t wex = 0;
    for (int w = 0; w < warp; ++w) wex += s_wsum[w];
    return wex + incl - count;
}

__global__ void __launch_bounds__(TS_THREADS)
tile_sort_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint64_t* __restrict__ keys_tmp,
                 uint32_t* __restrict__ vals_tmp, const uint2* __restrict__ ranges) {
    extern __shared__ __align__(16) unsigned char ts_raw[];
    uint32_t* s_key = reinterpret_cast<uint32_t*>(ts_raw);          // [2][TS_CAP] depth bits
    uint32_t* s_val = s_key + 2 * TS_CAP;                           // [2][TS_CAP]
    uint32_t* s_cnt = s_val + 2 * TS_CAP;                           // [TS_WARPS][257]
    uint32_t* s_dstart = s_cnt + TS_WARPS * TS_CNT_STRIDE;          // [257]
    uint32_t* s_wsum = s_dstart + 257;                              // [8]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint2 range = ranges[blockIdx.x];
    const uint32_t n = range.y - range.x;
    if (n <= 1) return;
    const uint32_t lt = lanemask_lt();
    uint32_t* wc = s_cnt + warp * TS_CNT_STRIDE;
    uint32_t dr[TS_ITEMS];

    if (n <= (uint32_t)TS_CAP) {
        // ---- shared-memory path
        for (uint32_t i = tid; i < n; i += TS_THREADS) {
            s_key[i] = (uint32_t)keys[range.x + i];
            s_val[i] = vals[range.x + i];
        }
        const uint32_t chunk = ((n + TS_THREADS - 1) / TS_THREADS) * 32;   // items per warp
        const int kmax = (int)(chunk / 32);
        const uint32_t wbase = warp * chunk + lane;
        int cur = 0;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = pass * 8;
            for (int i = tid; i < TS_WARPS * TS_CNT_STRIDE; i += TS_THREADS) s_cnt[i] = 0;
            __syncthreads();
            const uint32_t* kin = s_key + cur * TS_CAP;
#pragma unroll
            for (int k = 0; k < TS_ITEMS; ++k) {
                const uint32_t i = wbase + k * 32;
                dr[k] = (k < kmax && i < n) ? ((kin[i] >> shift) & 255u) : 256u;
            }
            ts_rank(dr, kmax, wc, lt);
            __syncthreads();
            uint32_t count;
            const uint32_t dstart = ts_digit_scan(s_cnt, s_wsum, count);
            s_dstart[tid] = dstart;
            // a pass whose digit is the same for every entry is the identity permutation: skip it
            const int uniform = __syncthreads_or(count == n);
            if (!uniform) {
                const uint32_t* vin = s_val + cur * TS_CAP;
                uint32_t* kout = s_key + (cur ^ 1) * TS_CAP;
                uint32_t* vout = s_val + (cur ^ 1) * TS_CAP;
#pragma unroll
                for (int k = 0; k < TS_ITEMS; ++k) {
                    const uint32_t i = wbase + k * 32;
                    if (k < kmax && i < n) {
                        const uint32_t d = dr[k] & 0xffffu;
                        const uint32_t pos = s_dstart[d] + wc[d] + (dr[k] >> 16);
                        kout[pos] = kin[i];
                        vout[pos] = vin[i];
                    }
                }
                cur ^= 1;
            }
            __syncthreads();
        }
        const uint32_t* vfin = s_val + cur * TS_CAP;
        for (uint32_t i = tid; i < n; i += TS_THREADS) vals[range.x + i] = vfin[i];
        return;
    }

    // ---- long runs: the same stable passes through global memory (this CTA only), ping-pong
    // between the run's own slots and the same slots of the spare buffers; 4 passes end in place.
    uint64_t* kin = keys + range.x;  uint32_t* vin = vals + range.x;
    uint64_t* kout = keys_tmp + range.x;  uint32_t* vout = vals_tmp + range.x;
    uint32_t* s_base = s_key;      // [256] running destination of each digit (shared-memory reuse)
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = pass * 8;
        s_base[tid] = 0;
        __syncthreads();
        for (uint32_t i = tid; i < n; i += TS_THREADS) atomicAdd(&s_base[(uint32_t)(kin[i] >> shift) & 255u], 1u);
        __syncthreads();
        {
            const uint32_t c = s_base[tid];
            uint32_t incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            uint32_t wex = 0;
            for (int w = 0; w < warp; ++w) wex += s_wsum[w];
            s_base[tid] = wex + incl - c;
        }
        __syncthreads();
        for (uint32_t c0 = 0; c0 < n; c0 += TS_CAP) {
            for (int i = tid; i < TS_WARPS * TS_CNT_STRIDE; i += TS_THREADS) s_cnt[i] = 0;
            __syncthreads();
            const uint32_t wbase = c0 + warp * (32 * TS_ITEMS) + lane;
#pragma unroll
            for (int k = 0; k < TS_ITEMS; ++k) {
                const uint32_t i = wbase + k * 32;
                dr[k] = i < n ? ((uint32_t)(kin[i] >> shift) & 255u) : 256u;
            }
            ts_rank(dr, TS_ITEMS, wc, lt);
            __syncthreads();
            uint32_t count;
            (void)ts_digit_scan(s_cnt, s_wsum, count);
#pragma unroll
            for (int k = 0; k < TS_ITEMS; ++k) {
                const uint32_t i = wbase + k * 32;
                if (i < n) {
                    const uint32_t d = dr[k] & 0xffffu;
                    const uint32_t pos = s_base[d] + wc[d] + (dr[k] >> 16);
                    kout[pos] = kin[i];
                    vout[pos] = vin[i];
                }
            }
            __syncthreads();
            s_base[tid] += count;
            __syncthreads();
        }
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
}

// ------------------------------------------------------------------------ host launchers
int launch_scan(const uint32_t* in, uint32_t* out, int P, uint32_t* status, uint32_t* counters,
                B200sHeader* header, cudaStream_t stream) {
    if (P <= 0) return 0;
    const int tiles = (P + SCAN_TILE - 1) / SCAN_TILE;
    count_launch();
    scan_tiles_kernel<<<tiles, SCAN_THREADS, 0, stream>>>(in, out, P, status, counters, header);
    return (int)cudaGetLastError();
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

// Number of onesweep passes over the TILE bits [32, 32+bit); the sorted run lives in ping-pong
// buffer (passes & 1).
int sort_num_passes(int tiles) {
    const int bits = (int)higher_msb((uint32_t)tiles);
    const int p = (bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
    return p < 1 ? 1 : p;
}

// Runs expand + sort + ranges. `n_hint` bounds the number of instances for grid sizing
// (capacity when the host does not know N). Returns in *final_buf which ping-pong buffer
// holds the sorted result.
int launch_binning(int P, const SplatRec* rec, const uint32_t* offsets, const BinWS& bw,
                   int64_t capacity, int64_t n_hint, int grid_x, int tiles, uint2* ranges,
                   B200sHeader* header, uint32_t* counters, int* final_buf, bool debug,
                   cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        B200S_CUDA_OK(cudaFuncSetAttribute(sort_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)SORT_SMEM));
        B200S_CUDA_OK(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TS_SMEM));
        attr_set = true;
    }
    const int passes = sort_num_passes(tiles);
    if (n_hint > capacity) n_hint = capacity;
    if (n_hint < 1) n_hint = 1;
    {
        ProfScope ps(B200S_STAGE_MISC, stream);
        B200S_CUDA_OK(cudaMemsetAsync(bw.hist, 0, bw.zero_bytes, stream));
        B200S_CUDA_OK(cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, stream));
    }
    if (P > 0) {
        ProfScope ps(B200S_STAGE_EXPAND, stream);
        count_launch();
        expand_keys_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, rec, offsets, bw.keys[0], bw.vals[0],
                                                               grid_x, header, capacity);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    const int64_t sort_tiles = (n_hint + SORT_TILE - 1) / SORT_TILE;
    prof_begin(B200S_STAGE_SORT, stream);
    {
        count_launch(2);
        int hb = (int)((n_hint + 256 * 16 - 1) / (256 * 16));
        if (hb > 148 * 8) hb = 148 * 8;
        if (hb < 1) hb = 1;
        sort_histogram_kernel<<<hb, 256, 0, stream>>>(bw.keys[0], header, capacity, 32, passes, bw.hist);
        B200S_AFTER_LAUNCH(stream, debug);
        sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, stream>>>(bw.hist);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    int cur = 0;
    for (int p = 0; p < passes; ++p) {
        count_launch();
        sort_onesweep_kernel<<<(unsigned)sort_tiles, SORT_THREADS, SORT_SMEM, stream>>>(
            bw.keys[cur], bw.vals[cur], bw.keys[cur ^ 1], bw.vals[cur ^ 1], header, capacity,
            32 + p * SORT_RADIX_BITS, bw.hist + p * SORT_RADIX,
            bw.lookback + (size_t)p * bw.sort_tiles * SORT_RADIX, counters + CTR_SORT_TICKET0 + p);
        B200S_AFTER_LAUNCH(stream, debug);
        cur ^= 1;
    }
    prof_end(B200S_STAGE_SORT, stream);
    *final_buf = cur;
    {
        ProfScope ps(B200S_STAGE_RANGES, stream);
        count_launch();
        int rb = (int)((n_hint + 255) / 256);
        if (rb > 148 * 16) rb = 148 * 16;
        tile_ranges_kernel<<<rb, 256, 0, stream>>>(bw.keys[cur], header, capacity, ranges);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    {
        // depth order inside every tile (ranges are all-zero when the binning overflowed: no work)
        ProfScope ps(B200S_STAGE_SORT, stream);
        count_launch();
        tile_sort_kernel<<<tiles, TS_THREADS, TS_SMEM, stream>>>(bw.keys[cur], bw.vals[cur], bw.keys[cur ^ 1],
                                                                bw.vals[cur ^ 1], ranges);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    return 0;
}

}  // namespace b200s
