// Factored SH gradient: rebuild (and sum over views) dL/dsh from the 3-float per-Gaussian factors.
//
// Within one view dL/dsh[k][c] = basis_k(dir) * dL/dRGB[c] (computeColorFromSH backward, reference
// backward.cu:20-139): a rank-1 48-float row determined by 3 floats and the view direction, which every rank can
// recompute from the replicated means and the view's camera centre. A view-parallel step therefore gathers
// 3 floats per Gaussian and view over NVLink instead of all-reducing 48 (the dense
// ncclAllReduce of the 59 gradient floats is the baseline), and this kernel forms
//     dL_dsh[i][k][c] = sum_v fl( basis_k(dir_v(i)) * g_v[i][c] )      (views in order, one rounding per op)
// locally: same products (sh_grad.cuh, shared with preprocess_bwd) and the same fixed summation order as the
// two-shot all-reduce of the materialised rows, so the bits are the same. 12 B x views + 12 B in, 192 B out per
// Gaussian; rows leave the SM as full 16-byte vectors staged through shared memory.
//
// The exact form costs ~170 instructions per (Gaussian, view) — IEEE sqrt and three IEEE divides for the direction,
// 48 products and 48 sums that must not fuse — and a warp runs a view as soon as ONE of its 32 Gaussians saw it, so at
// 8 views the kernel is issue bound (127 M warp instructions, 0.26 ms at P = 3 M for 0.13 ms of HBM time), and it is the
// part of the multi-GPU step that does not hide behind the next view (the blend kernels leave no SM slots free).
// FAST (args.fast = 1, what the collective uses when its dense part is summed inside the NVSwitch, i.e. when the
// step is not a rank-order sum anyway): rsqrt-based direction and 48 fused multiply-adds, ~90 instructions; same
// sum to ~1e-7 relative, identical on every rank. The SH degree 3 case is compiled without the per-coefficient test.

#include "kernels.cuh"
#include "sh_grad.cuh"

namespace b200s {

constexpr int SR_THREADS = 128;
constexpr int SR_MAX_VIEWS = B200S_MAX_PEERS;

struct ShReconParams {
    int P, D, M, views;
    const float* means3D;
    const float* factors;
    size_t stride;
    float* dL_dsh;
};

#ifndef B200S_SR_MINB
#define B200S_SR_MINB 5
#endif
template <bool SH16, bool FAST, int DEG>     // DEG < 0: degree read from the parameters
__global__ void __launch_bounds__(SR_THREADS, B200S_SR_MINB)
sh_grad_reconstruct_kernel(const __grid_constant__ ShReconParams p) {
    __shared__ __align__(16) float tile[SR_THREADS * SH_ROW_PAD];
    __shared__ float s_cam[SR_MAX_VIEWS][4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < p.views * 4) s_cam[tid >> 2][tid & 3] = p.factors[(size_t)(tid >> 2) * p.stride + 3 * (size_t)p.P + (tid & 3)];
    __syncthreads();
    const int64_t idx = (int64_t)blockIdx.x * SR_THREADS + tid;
    const bool in_range = idx < p.P;
    const int deg = DEG >= 0 ? DEG : p.D;
    const int ncoef = (deg + 1) * (deg + 1);

    float acc[SH_ROW_FLOATS];
#pragma unroll
    for (int i = 0; i < SH_ROW_FLOATS; ++i) acc[i] = 0.0f;
    if (in_range) {
        const float mx = p.means3D[3 * idx], my = p.means3D[3 * idx + 1], mz = p.means3D[3 * idx + 2];
        // the factor loads of eight views are issued back to back before any of them is consumed (the kernel is
        // HBM-latency bound otherwise: one dependent 12-byte load per view and thread)
        for (int v0 = 0; v0 < p.views; v0 += 8) {
            float gx[8], gy[8], gz[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                gx[q] = gy[q] = gz[q] = 0.0f;
                if (v0 + q < p.views) {
                    const float* g = p.factors + (size_t)(v0 + q) * p.stride + 3 * idx;
                    gx[q] = __ldg(g); gy[q] = __ldg(g + 1); gz[q] = __ldg(g + 2);
                }
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (gx[q] == 0.0f && gy[q] == 0.0f && gz[q] == 0.0f) continue;   // culled / clamped in that view: an all-zero row
                float3 dir_orig, dir;
                if (FAST) {
                    const float dx = mx - s_cam[v0 + q][0], dy = my - s_cam[v0 + q][1], dz = mz - s_cam[v0 + q][2];
                    const float inv = rsqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
                    dir = make_float3(dx * inv, dy * inv, dz * inv);
                } else {
                    sh_view_dir(mx, my, mz, s_cam[v0 + q][0], s_cam[v0 + q][1], s_cam[v0 + q][2], dir_orig, dir);
                }
                float bas[16];
                sh_basis(deg, dir.x, dir.y, dir.z, bas);
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    if (k < ncoef) {
                        if (FAST) {
                            acc[3 * k] = fmaf(bas[k], gx[q], acc[3 * k]);
                            acc[3 * k + 1] = fmaf(bas[k], gy[q], acc[3 * k + 1]);
                            acc[3 * k + 2] = fmaf(bas[k], gz[q], acc[3 * k + 2]);
                        } else {
                            acc[3 * k] = fadd(acc[3 * k], fmul(bas[k], gx[q]));
                            acc[3 * k + 1] = fadd(acc[3 * k + 1], fmul(bas[k], gy[q]));
                            acc[3 * k + 2] = fadd(acc[3 * k + 2], fmul(bas[k], gz[q]));
                        }
                    }
                }
            }
        }
    }
    if (SH16) {
        float* row = tile + tid * SH_ROW_PAD;
#pragma unroll
        for (int k = 0; k < 12; ++k)
            reinterpret_cast<float4*>(row)[k] = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
        __syncwarp();
        const int64_t gbase = (int64_t)blockIdx.x * SR_THREADS + warp * 32;
        float* gdst = p.dL_dsh + gbase * SH_ROW_FLOATS;
        const float* wt = tile + warp * 32 * SH_ROW_PAD;
#pragma unroll
        for (int it = 0; it < 12; ++it) {
            const int c = it * 32 + lane;
            const int r = c / 12;
            const int ch = c - r * 12;
            if (gbase + r < p.P)
                reinterpret_cast<float4*>(gdst)[c] = *reinterpret_cast<const float4*>(wt + r * SH_ROW_PAD + ch * 4);
        }
    } else if (in_range) {
        float* out = p.dL_dsh + idx * p.M * 3;
        for (int k = 0; k < p.M * 3; ++k) out[k] = k < SH_ROW_FLOATS ? acc[k] : 0.0f;
    }
}

int launch_sh_grad_reconstruct(const B200sShGradArgs& a, cudaStream_t stream) {
    if (a.P <= 0) return 0;
    ShReconParams p;
    p.P = a.P; p.D = a.D; p.M = a.M; p.views = a.views;
    p.means3D = a.means3D; p.factors = a.factors; p.stride = (size_t)a.stride_floats; p.dL_dsh = a.dL_dsh;
    const int blocks = (a.P + SR_THREADS - 1) / SR_THREADS;
    count_launch();
    if (a.M == 16 && a.D == 3) {
        if (a.fast) sh_grad_reconstruct_kernel<true, true, 3><<<blocks, SR_THREADS, 0, stream>>>(p);
        else sh_grad_reconstruct_kernel<true, false, 3><<<blocks, SR_THREADS, 0, stream>>>(p);
    } else if (a.M == 16) {
        if (a.fast) sh_grad_reconstruct_kernel<true, true, -1><<<blocks, SR_THREADS, 0, stream>>>(p);
        else sh_grad_reconstruct_kernel<true, false, -1><<<blocks, SR_THREADS, 0, stream>>>(p);
    } else {
        if (a.fast) sh_grad_reconstruct_kernel<false, true, -1><<<blocks, SR_THREADS, 0, stream>>>(p);
        else sh_grad_reconstruct_kernel<false, false, -1><<<blocks, SR_THREADS, 0, stream>>>(p);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200s
