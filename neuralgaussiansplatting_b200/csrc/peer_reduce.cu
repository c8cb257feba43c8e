// Gradient all-reduce over NVLink peer memory (the one collective of the path —
// a SUM of the 59 gradient floats per Gaussian across the view-parallel ranks). The reference has
// no multi-GPU code (utils/general_utils.py:133); the baseline this replaces is ncclAllReduce.
//
// Every rank holds the gradient buffer at the same offset of a symmetric allocation whose peer
// mappings (and, on NVSwitch, one multicast mapping) are passed in. preprocess_bwd writes its
// outputs straight into that buffer; this kernel then runs the whole collective in ONE launch:
//
//   start barrier   CTA b of every rank tells CTA b of every peer "my gradients are complete"
//   two-shot sum    rank r owns elements [r*n/world, (r+1)*n/world): it reads that slice from every
//                   rank in rank order (fixed-order fp32 sum -> every replica gets bit-identical
//                   values, run to run), and stores the total into every rank's buffer
//                   (multimem mode: one multimem.ld_reduce + one multimem.st per 16 bytes instead;
//                   the NVSwitch adds and replicates)
//   end barrier     CTA b waits until CTA b of every peer has stored its slice here
//
// Per GPU and direction this moves (world-1)/world * n bytes over NVLink, the minimum for an
// all-reduce without in-switch reduction, and it occupies only `ctas` CTAs of 512 threads, so the
// next view's rasterization keeps (almost) all SMs while the gradients of this one are summed.
// Flags are monotonically increasing epochs in a second symmetric allocation; no reset is needed.
//
// Visibility-sparse tail. A Gaussian culled in a rank's view has an all-zero gradient row there
// (preprocess_bwd zero-fills rows with radii <= 0), and a view sees well under half of a large
// scene. When the caller passes its radii, each rank first publishes one visibility byte per
// Gaussian next to its buffer; for the last (row-structured) tensor of the buffer — dL/dSH, 192 of
// the 236 bytes per Gaussian — the owner then reads a row only from the ranks that saw it and
// skips rows nobody saw. Adding the skipped +0.0 terms would not change a bit of the fixed-order
// sum, so the result is identical to the dense one; NVLink read traffic drops to the visible
// fraction (measured at 8 ranks: the collective is NVLink-bound, so this is time saved).

#include "kernels.cuh"

namespace b200s {

constexpr int PR_THREADS = 512;

struct PeerReduceParams {
    float4* bufs[B200S_MAX_PEERS];
    uint32_t* flags[B200S_MAX_PEERS];
    float4* mc;            // multicast mapping of the same buffer, or NULL
    size_t n4;             // float4 elements
    int rank, world;
    uint32_t epoch;
    // visibility-sparse tail (radii == NULL: the whole buffer is reduced densely)
    const int* radii;      // local radii of the view these gradients belong to
    int rows;              // Gaussians
    int rows_per;          // rows owned per rank, a multiple of PR_CHUNK
    size_t sparse_off4;    // first float4 of the row-structured tensor (PR_ROW4 float4 per row)
    size_t vis_off;        // byte offset of the visibility array from bufs[r]
    // all-gather region: slot r (gslot4 float4) at goff4 + r*gslot4 of every buffer ends up holding rank r's data
    size_t goff4, gslot4;
    int gather_push;       // 1: every rank multimem.st's its own slot through the NVSwitch multicast mapping
};

constexpr unsigned long long PR_BARRIER_TIMEOUT_NS = 60ull * 1000000000ull;   // peers start within a minute or never
constexpr int PR_CHUNK = 1024;   // rows per work unit of the sparse tail
constexpr int PR_ROW4 = 12;      // float4 per row: 16 SH coefficients x 3 channels

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Data accesses are WEAK (plain LDG/STG; .STRONG.SYS accesses measured 2x slower over NVLink): every address is
// read at most once per launch and only after the start barrier (acquire + bar.sync), and the stores are made
// visible by the end barrier's fence.sys + release.
__device__ __forceinline__ float4 ld_peer(const float4* p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_peer(float4* p, const float4& v) {
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce(const float4* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* p, const float4& v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Flag slot written by (phase, cta, source rank) in the DESTINATION rank's flag array.
__device__ __forceinline__ size_t flag_slot(int phase, int cta, int ctas, int src, int world) {
    return ((size_t)phase * ctas + cta) * world + src;
}

// All ranks' CTA `blockIdx.x` meet. Threads 0..world-1 each handle one peer.
__device__ __forceinline__ void peer_barrier(const PeerReduceParams& p, int phase) {
    __syncthreads();
    if ((int)threadIdx.x < p.world) {
        const int peer = threadIdx.x;
        __threadfence_system();
        st_release_sys(p.flags[peer] + flag_slot(phase, blockIdx.x, gridDim.x, p.rank, p.world), p.epoch);
        const uint32_t* mine = p.flags[p.rank] + flag_slot(phase, blockIdx.x, gridDim.x, peer, p.world);
        // Failure detection: a peer that died (or never launched this collective) must not hang this GPU for good.
        // After PR_BARRIER_TIMEOUT_NS without the peer's flag the kernel traps, which surfaces as a CUDA error on the
        // host's next synchronisation instead of a silent hang.
        uint32_t spins = 0;
        unsigned long long t0 = 0;
        while ((int32_t)(ld_acquire_sys(mine) - p.epoch) < 0) {
            __nanosleep(64);
            if ((++spins & 0x3fffu) == 0u) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > PR_BARRIER_TIMEOUT_NS) __trap();
            }
        }
    }
    __syncthreads();
}

// All-gather of the per-rank slots (factored SH gradient: 3 floats per Gaussian instead of the 48-float row):
// this rank pulls slot r from rank r's buffer into slot r of its own, for every peer r. Loads of up to eight
// peers x UG elements are in flight per thread before the first store.
__device__ __forceinline__ void peer_gather_slots(const PeerReduceParams& p, int world) {
    constexpr int UG = 2;
    const size_t stride = (size_t)gridDim.x * PR_THREADS;
    float4* mine = p.bufs[p.rank] + p.goff4;
    for (int r0 = 0; r0 < world; r0 += 8) {
        for (size_t i0 = (size_t)blockIdx.x * PR_THREADS + threadIdx.x; i0 < p.gslot4; i0 += stride * UG) {
            float4 v[UG][8];
#pragma unroll
            for (int u = 0; u < UG; ++u) {
                const size_t i = i0 + u * stride;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int r = r0 + q;
                    if (r < world && r != p.rank && i < p.gslot4) v[u][q] = ld_peer(p.bufs[r] + p.goff4 + (size_t)r * p.gslot4 + i);
                }
            }
#pragma unroll
            for (int u = 0; u < UG; ++u) {
                const size_t i = i0 + u * stride;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int r = r0 + q;
                    if (r < world && r != p.rank && i < p.gslot4) st_peer(mine + (size_t)r * p.gslot4 + i, v[u][q]);
                }
            }
        }
    }
}

// The same all-gather as a PUSH through the NVSwitch multicast mapping: each rank streams its own slot from local
// memory into `multimem.st`, the switch replicates every 16 bytes to all peers. Per GPU 1 x slot leaves and
// (world-1) x slot arrives, no read round trips over NVLink and 1/(world-1) of the pull's SM work.
__device__ __forceinline__ void peer_gather_push(const PeerReduceParams& p) {
    constexpr int UP = 8;
    const size_t stride = (size_t)gridDim.x * PR_THREADS;
    const size_t off = p.goff4 + (size_t)p.rank * p.gslot4;
    const float4* src = p.bufs[p.rank] + off;
    float4* dst = p.mc + off;
    for (size_t i0 = (size_t)blockIdx.x * PR_THREADS + threadIdx.x; i0 < p.gslot4; i0 += stride * UP) {
        float4 v[UP];
#pragma unroll
        for (int u = 0; u < UP; ++u) {
            const size_t i = i0 + u * stride;
            if (i < p.gslot4) v[u] = ld_peer(src + i);
        }
#pragma unroll
        for (int u = 0; u < UP; ++u) {
            const size_t i = i0 + u * stride;
            if (i < p.gslot4) multimem_st(dst + i, v[u]);
        }
    }
}

template <int WORLD, bool MULTIMEM, bool SPARSE>
__global__ void __launch_bounds__(PR_THREADS)
peer_allreduce_kernel(const __grid_constant__ PeerReduceParams p) {
    const int world = WORLD > 0 ? WORLD : p.world;
    if (SPARSE) {
        // Publish this rank's visibility bytes. CTA b writes exactly the rows CTA b of each owner will read,
        // so the per-CTA start barrier below orders them.
        uint8_t* myvis = reinterpret_cast<uint8_t*>(p.bufs[p.rank]) + p.vis_off;
        const int chunks_per = p.rows_per / PR_CHUNK;
        for (int o = 0; o < world; ++o)
            for (int c = blockIdx.x; c < chunks_per; c += gridDim.x) {
                const int row0 = o * p.rows_per + c * PR_CHUNK;
                for (int j = threadIdx.x; j < PR_CHUNK; j += PR_THREADS) {
                    const int g = row0 + j;
                    if (g < p.rows) myvis[g] = p.radii[g] > 0 ? 1 : 0;
                }
            }
    }
    peer_barrier(p, 0);
    // the pushed slots travel while the dense part is summed (stores are fire-and-forget)
    if (p.gslot4 != 0 && p.gather_push) peer_gather_push(p);

    const size_t n4 = SPARSE ? p.sparse_off4 : p.n4;     // densely reduced part
    const size_t per = (n4 + world - 1) / world;
    const size_t begin = min(n4, per * (size_t)p.rank);
    const size_t end = min(n4, begin + per);
    const size_t stride = (size_t)gridDim.x * PR_THREADS;
    constexpr int U = MULTIMEM ? 8 : (WORLD > 0 && WORLD <= 2 ? 8 : WORLD > 0 && WORLD <= 4 ? 4 : 2);

    for (size_t i0 = begin + (size_t)blockIdx.x * PR_THREADS + threadIdx.x; i0 < end; i0 += stride * U) {
        float4 acc[U];
        if (MULTIMEM) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                if (i < end) acc[u] = multimem_ld_reduce(p.mc + i);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                if (i < end) multimem_st(p.mc + i, acc[u]);
            }
        } else {
            float4 v[U][WORLD > 0 ? WORLD : 1];
            if (WORLD > 0) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const size_t i = i0 + u * stride;
#pragma unroll
                    for (int r = 0; r < WORLD; ++r)
                        if (i < end) v[u][r] = ld_peer(p.bufs[r] + i);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    acc[u] = v[u][0];
#pragma unroll
                    for (int r = 1; r < WORLD; ++r) {
                        acc[u].x += v[u][r].x; acc[u].y += v[u][r].y; acc[u].z += v[u][r].z; acc[u].w += v[u][r].w;
                    }
                }
            } else {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const size_t i = i0 + u * stride;
                    if (i >= end) continue;
                    acc[u] = ld_peer(p.bufs[0] + i);
                    for (int r = 1; r < world; ++r) {
                        const float4 t = ld_peer(p.bufs[r] + i);
                        acc[u].x += t.x; acc[u].y += t.y; acc[u].z += t.z; acc[u].w += t.w;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                if (i < end) {
                    if (WORLD > 0) {
#pragma unroll
                        for (int r = 0; r < WORLD; ++r) st_peer(p.bufs[r] + i, acc[u]);
                    } else {
                        for (int r = 0; r < world; ++r) st_peer(p.bufs[r] + i, acc[u]);
                    }
                }
            }
        }
    }

    if (p.gslot4 != 0 && !p.gather_push) peer_gather_slots(p, world);

    if constexpr (SPARSE && WORLD > 0) {
        __shared__ uint32_t smask[PR_CHUNK / 4];        // one byte per row: bit r = visible on rank r
        constexpr int US = WORLD <= 4 ? 4 : 2;
        const int chunks_per = p.rows_per / PR_CHUNK;
        for (int c = blockIdx.x; c < chunks_per; c += gridDim.x) {
            const int row0 = p.rank * p.rows_per + c * PR_CHUNK;
            if (row0 >= p.rows) break;
            __syncthreads();
            if (threadIdx.x < PR_CHUNK / 4) {
                uint32_t m = 0;
#pragma unroll
                for (int r = 0; r < WORLD; ++r) {
                    const uint8_t* vis = reinterpret_cast<const uint8_t*>(p.bufs[r]) + p.vis_off + row0;
                    m |= ld_relaxed_sys_u32(reinterpret_cast<const uint32_t*>(vis) + threadIdx.x) << r;
                }
                smask[threadIdx.x] = m;
            }
            __syncthreads();
            const uint8_t* mask = reinterpret_cast<const uint8_t*>(smask);
            const int n_el = min(PR_CHUNK, p.rows - row0) * PR_ROW4;
            const size_t base = p.sparse_off4 + (size_t)row0 * PR_ROW4;
            for (int e0 = threadIdx.x; e0 < n_el; e0 += PR_THREADS * US) {
                float4 v[US][WORLD];
                uint32_t m[US];
#pragma unroll
                for (int u = 0; u < US; ++u) {
                    const int e = e0 + u * PR_THREADS;
                    m[u] = e < n_el ? mask[e / PR_ROW4] : 0u;
#pragma unroll
                    for (int r = 0; r < WORLD; ++r)
                        if ((m[u] >> r) & 1u) v[u][r] = ld_peer(p.bufs[r] + base + e);
                }
#pragma unroll
                for (int u = 0; u < US; ++u) {
                    if (m[u] == 0u) continue;            // seen by no rank: every buffer already holds zeros
                    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
                    for (int r = 0; r < WORLD; ++r)
                        if ((m[u] >> r) & 1u) {
                            acc.x += v[u][r].x; acc.y += v[u][r].y; acc.z += v[u][r].z; acc.w += v[u][r].w;
                        }
                    const int e = e0 + u * PR_THREADS;
#pragma unroll
                    for (int r = 0; r < WORLD; ++r) st_peer(p.bufs[r] + base + e, acc);
                }
            }
        }
    }
    peer_barrier(p, 1);
}

template <int WORLD, bool MM, bool SP = false>
static int launch_pr(const PeerReduceParams& p, int ctas, cudaStream_t stream) {
    count_launch();
    peer_allreduce_kernel<WORLD, MM, SP><<<ctas, PR_THREADS, 0, stream>>>(p);
    return (int)cudaGetLastError();
}

int launch_peer_allreduce(const B200sPeerReduceArgs& a, cudaStream_t stream) {
    PeerReduceParams p{};
    for (int r = 0; r < a.world; ++r) {
        p.bufs[r] = (float4*)a.peer_bufs[r];
        p.flags[r] = (uint32_t*)a.peer_flags[r];
    }
    p.mc = (float4*)a.multicast_buf;
    p.n4 = (size_t)(a.n_floats / 4);
    p.rank = a.rank; p.world = a.world; p.epoch = a.epoch;
    p.goff4 = (size_t)(a.gather_offset_floats / 4);
    p.gslot4 = (size_t)(a.gather_floats / 4);
    p.gather_push = (a.use_multimem & 2) && a.multicast_buf != nullptr;
    const int ctas = a.ctas;
    if (a.use_multimem & 1) return launch_pr<0, true>(p, ctas, stream);
    if (a.radii && a.sparse_rows > 0) {
        p.radii = (const int*)a.radii;
        p.rows = a.sparse_rows;
        p.rows_per = (int)b200s_peer_allreduce_rows_per_rank(a.sparse_rows, a.world);
        p.sparse_off4 = (size_t)(a.sparse_offset_floats / 4);
        p.vis_off = (size_t)a.vis_offset_bytes;
        switch (a.world) {
            case 2: return launch_pr<2, false, true>(p, ctas, stream);
            case 4: return launch_pr<4, false, true>(p, ctas, stream);
            case 8: return launch_pr<8, false, true>(p, ctas, stream);
            default: break;   // other world sizes: dense
        }
    }
    switch (a.world) {
        case 2: return launch_pr<2, false>(p, ctas, stream);
        case 4: return launch_pr<4, false>(p, ctas, stream);
        case 8: return launch_pr<8, false>(p, ctas, stream);
        default: return launch_pr<0, false>(p, ctas, stream);
    }
}

}  // namespace b200s
