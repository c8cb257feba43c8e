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
//  * sort:    onesweep: one histogram pass over the keys for all digits, then one
//             read+write pass per 8-bit digit with a chained look-back over 4,096-key tiles;
//             in-tile ranking by warp match, stable by construction.
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
                      int passes, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_h[SORT_MAX_PASSES * SORT_RADIX];
    uint32_t n = header->num_rendered;
    if ((int64_t)n > capacity) n = 0;
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += 256) s_h[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        const uint64_t key = keys[i];
        for (int p = 0; p < passes; ++p)
            atomicAdd(&s_h[p * SORT_RADIX + (uint32_t)((key >> (p * SORT_RADIX_BITS)) & (SORT_RADIX - 1))], 1u);
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
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const uint32_t d = dr[k];
        const uint32_t peers = __match_any_sync(FULL_MASK, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader) { old = wc[d]; wc[d] = old + __popc(peers); }
        old = __shfl_sync(FULL_MASK, old, leader);
        dr[k] = d | ((old + __popc(peers & lt)) << 16);
        __syncwarp();
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

int sort_num_passes(int tiles) {
    const int bits = 32 + (int)higher_msb((uint32_t)tiles);
    return (bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
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
        sort_histogram_kernel<<<hb, 256, 0, stream>>>(bw.keys[0], header, capacity, passes, bw.hist);
        B200S_AFTER_LAUNCH(stream, debug);
        sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, stream>>>(bw.hist);
        B200S_AFTER_LAUNCH(stream, debug);
    }
    int cur = 0;
    for (int p = 0; p < passes; ++p) {
        count_launch();
        sort_onesweep_kernel<<<(unsigned)sort_tiles, SORT_THREADS, SORT_SMEM, stream>>>(
            bw.keys[cur], bw.vals[cur], bw.keys[cur ^ 1], bw.vals[cur ^ 1], header, capacity,
            p * SORT_RADIX_BITS, bw.hist + p * SORT_RADIX,
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
    return 0;
}

}  // namespace b200s
