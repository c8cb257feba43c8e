// Decoupled look-back helpers shared by the fused preprocess scan (preprocess.cu) and the binning kernels.
#pragma once
#include "common.cuh"

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


}  // namespace b200s
