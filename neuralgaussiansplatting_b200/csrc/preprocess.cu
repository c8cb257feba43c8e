// K1 — fused forward preprocess: near-plane cull, EWA 3D->2D covariance / conic,
// eigen-radius, tile rectangle + tiles_touched, SH degree<=3 -> RGB.
// Replaces preprocessCUDA + helpers (reference forward.cu:20-256, auxiliary.h:41-164)
// and checkFrustum (rasterizer_impl.cu:54-66).
//
// One thread per Gaussian for the geometry chain (bit-exact, see common.cuh). The 192-B SH
// rows — 62 % of the bytes this kernel reads — are not fetched by the owning thread with
// strided scalar loads (the reference: 48 LDG.32 per thread, 32 lines per warp
// instruction); instead each warp streams the rows of its VISIBLE Gaussians with
// fully-coalesced 16-B cp.async (LDGSTS) into a padded shared-memory tile and every lane
// then reads its own row conflict-free. Culled Gaussians cost 12 B read + 8 B written.

#include "kernels.cuh"

namespace b200s {


constexpr int PRE_THREADS = 256;
constexpr int SH_ROW_FLOATS = 48;   // 16 coefficients x RGB
constexpr int SH_ROW_PAD = 52;      // 208 B rows: 13 x 16 B => LDS.128 conflict-free
constexpr size_t PRE_SMEM_SH16 = (size_t)PRE_THREADS * SH_ROW_PAD * sizeof(float);

__device__ const float kSH_C0 = 0.28209479177387814f;
__device__ const float kSH_C1 = 0.4886025119029199f;
__device__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                    -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                    0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                    -0.5900435899266435f};

// SH basis evaluation (computeColorFromSH, forward.cu:20-71). `sh(k)` returns coefficient k (RGB).
// The operation order — one FMA per coefficient, accumulated in coefficient order, with the
// polynomial factors contracted exactly as below — mirrors the SASS of the reference
// preprocessCUDA built for sm_100, so the RGB (and therefore the `clamped` flags that gate
// the SH gradient) come out bit-identical.
template <typename LoadF3>
__device__ __forceinline__ float3 sh_to_rgb(int deg, float x, float y, float z, LoadF3 sh) {
    const float3 c0 = sh(0);
    float3 res = make_float3(fmul(c0.x, kSH_C0), fmul(c0.y, kSH_C0), fmul(c0.z, kSH_C0));
#define B200S_SH_ACC(b, k)                                                        \
    {                                                                             \
        const float3 s_ = sh(k);                                                  \
        res.x = ffma((b), s_.x, res.x); res.y = ffma((b), s_.y, res.y); res.z = ffma((b), s_.z, res.z); \
    }
    if (deg > 0) {
        const float t1 = fmul(y, kSH_C1), t2 = fmul(z, kSH_C1), t3 = fmul(x, kSH_C1);
        B200S_SH_ACC(-t1, 1) B200S_SH_ACC(t2, 2) B200S_SH_ACC(-t3, 3)
        if (deg > 1) {
            const float xx = fmul(x, x), yy = fmul(y, y), zz = fmul(z, z);
            const float xy = fmul(y, x), yz = fmul(z, y), xz = fmul(z, x);
            const float zz2 = fadd(zz, zz);
            const float xx_yy = fsub(xx, yy);
            B200S_SH_ACC(fmul(xy, kSH_C2[0]), 4)
            B200S_SH_ACC(fmul(yz, kSH_C2[1]), 5)
            B200S_SH_ACC(fmul(fsub(fsub(zz2, xx), yy), kSH_C2[2]), 6)
            B200S_SH_ACC(fmul(xz, kSH_C2[3]), 7)
            B200S_SH_ACC(fmul(xx_yy, kSH_C2[4]), 8)
            if (deg > 2) {
                const float q = fsub(ffma(zz, 4.0f, -xx), yy);              // 4zz - xx - yy
                B200S_SH_ACC(fmul(fmul(y, kSH_C3[0]), ffma(xx, 3.0f, -yy)), 9)
                B200S_SH_ACC(fmul(fmul(xy, kSH_C3[1]), z), 10)
                B200S_SH_ACC(fmul(fmul(y, kSH_C3[2]), q), 11)
                B200S_SH_ACC(fmul(fmul(z, kSH_C3[3]), ffma(yy, -3.0f, ffma(xx, -3.0f, zz2))), 12)
                B200S_SH_ACC(fmul(q, fmul(x, kSH_C3[4])), 13)
                B200S_SH_ACC(fmul(xx_yy, fmul(z, kSH_C3[5])), 14)
                B200S_SH_ACC(fmul(fmul(x, kSH_C3[6]), ffma(yy, -3.0f, xx)), 15)
            }
        }
    }
#undef B200S_SH_ACC
    res.x = fadd(res.x, 0.5f); res.y = fadd(res.y, 0.5f); res.z = fadd(res.z, 0.5f);
    return res;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

template <bool SH16>
__global__ void __launch_bounds__(PRE_THREADS)
preprocess_fwd_kernel(const __grid_constant__ PreprocessParams p) {
    extern __shared__ __align__(16) float sh_tile[];   // [PRE_THREADS][SH_ROW_PAD] when SH16
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int64_t idx = (int64_t)blockIdx.x * PRE_THREADS + tid;
    const bool in_range = idx < p.P;

    bool visible = false;
    int radius_i = 0;
    uint32_t tiles = 0;
    float px = 0.f, py = 0.f, pz = 0.f;
    float depth = 0.f, img_x = 0.f, img_y = 0.f, con_a = 0.f, con_b = 0.f, con_c = 0.f;
    TileRect rect = {0, 0, 0, 0};

    if (in_range) {
        px = p.means3D[3 * idx + 0];
        py = p.means3D[3 * idx + 1];
        pz = p.means3D[3 * idx + 2];
        // in_frustum (auxiliary.h:139-164): near-plane only
        depth = affine_row_ref(p.viewmatrix, 2, px, py, pz);
        if (!(depth <= 0.2f)) {
            const float hx = affine_row_ref(p.projmatrix, 0, px, py, pz);
            const float hy = affine_row_ref(p.projmatrix, 1, px, py, pz);
            const float hw = affine_row_ref(p.projmatrix, 3, px, py, pz);
            const float p_w = __frcp_rn(fadd(hw, 0.0000001f));
            const float proj_x = fmul(hx, p_w);
            const float proj_y = fmul(hy, p_w);

            float cov3D[6];
            if (p.cov3D_precomp != nullptr) {
#pragma unroll
                for (int i = 0; i < 6; ++i) cov3D[i] = p.cov3D_precomp[6 * idx + i];
            } else {
                const float4 q = *reinterpret_cast<const float4*>(p.rotations + 4 * idx);
                cov3d_ref(p.scales[3 * idx + 0], p.scales[3 * idx + 1], p.scales[3 * idx + 2],
                          p.scale_modifier, q.x, q.y, q.z, q.w, cov3D);
            }
            const Cov2D c2 = cov2d_ref(px, py, pz, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov3D,
                                       p.viewmatrix);
            // forward.cu:219-232
            const float det = ffma(c2.a, c2.c, -fmul(c2.b, c2.b));
            if (det != 0.0f) {
                const float det_inv = __frcp_rn(det);
                con_a = fmul(c2.c, det_inv);
                con_b = fmul(c2.b, -det_inv);
                con_c = fmul(c2.a, det_inv);
                const float mid = fmul(fadd(c2.a, c2.c), 0.5f);
                const float disc = __fsqrt_rn(fmaxf(ffma(mid, mid, -det), 0.1f));
                const float lam = fmaxf(fadd(mid, disc), fsub(mid, disc));
                radius_i = __float2int_ru(fmul(__fsqrt_rn(lam), 3.0f));
                img_x = ndc2pix_ref(proj_x, p.W);
                img_y = ndc2pix_ref(proj_y, p.H);
                rect = get_rect_ref(img_x, img_y, radius_i, p.grid_x, p.grid_y);
                tiles = (uint32_t)((rect.x1 - rect.x0) * (rect.y1 - rect.y0));
                visible = tiles != 0;
            }
        } else if (p.prefiltered) {
            atomicOr(&p.header->error, (uint32_t)ERR_PREFILTERED);
        }
    }

    const uint32_t vis_mask = __ballot_sync(FULL_MASK, visible);

    float3 rgb = make_float3(0.f, 0.f, 0.f);
    uint32_t clamp_bits = 0;
    if (p.colors_precomp != nullptr) {
        if (visible) {
            rgb.x = p.colors_precomp[3 * idx + 0];
            rgb.y = p.colors_precomp[3 * idx + 1];
            rgb.z = p.colors_precomp[3 * idx + 2];
        }
    } else {
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (visible) {
            // glm::length as compiled: sqrt(fma(dz,dz, fma(dx,dx, dy*dy))), then IEEE divisions
            dx = fsub(px, p.campos[0]); dy = fsub(py, p.campos[1]); dz = fsub(pz, p.campos[2]);
            const float len = __fsqrt_rn(ffma(dz, dz, ffma(dx, dx, fmul(dy, dy))));
            dx = __fdiv_rn(dx, len); dy = __fdiv_rn(dy, len); dz = __fdiv_rn(dz, len);
        }
        float3 res;
        if (SH16) {
            // Stream the visible rows of this warp's 32 Gaussians: 384 16-B chunks.
            const int warp = tid >> 5;
            float* tile = sh_tile + (size_t)warp * 32 * SH_ROW_PAD;
            const float* gsrc = p.shs + ((int64_t)blockIdx.x * PRE_THREADS + warp * 32) * SH_ROW_FLOATS;
            if (vis_mask != 0) {
#pragma unroll
                for (int it = 0; it < 12; ++it) {
                    const int c = it * 32 + lane;
                    const int row = c / 12;
                    const int ch = c - row * 12;
                    if ((vis_mask >> row) & 1u)
                        cp_async16(tile + row * SH_ROW_PAD + ch * 4, gsrc + (size_t)c * 4);
                }
                cp_async_commit_wait_all();
            }
            __syncwarp();
            if (visible) {
                const float4* rowp = reinterpret_cast<const float4*>(tile + lane * SH_ROW_PAD);
                float v[SH_ROW_FLOATS];
#pragma unroll
                for (int k = 0; k < 12; ++k) {
                    const float4 t = rowp[k];
                    v[4 * k + 0] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
                }
                res = sh_to_rgb(p.D, dx, dy, dz,
                                [&](int k) { return make_float3(v[3 * k], v[3 * k + 1], v[3 * k + 2]); });
            }
        } else if (visible) {
            const float* sh = p.shs + (int64_t)idx * p.M * 3;
            res = sh_to_rgb(p.D, dx, dy, dz,
                            [&](int k) { return make_float3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]); });
        }
        if (visible) {
            clamp_bits = (res.x < 0.f ? 1u : 0u) | (res.y < 0.f ? 2u : 0u) | (res.z < 0.f ? 4u : 0u);
            rgb.x = fmaxf(res.x, 0.f); rgb.y = fmaxf(res.y, 0.f); rgb.z = fmaxf(res.z, 0.f);
        }
    }

    if (in_range) {
        p.radii[idx] = visible ? radius_i : 0;
        p.tiles_touched[idx] = tiles;
        if (visible) {
            SplatRec r;
            r.q0 = make_float4(img_x, img_y, con_a, con_b);
            r.q1 = make_float4(con_c, p.opacities[idx], rgb.x, rgb.y);
            r.q2 = make_float4(rgb.z, depth,
                               __uint_as_float((uint32_t)rect.x0 | ((uint32_t)rect.y0 << 16)),
                               __uint_as_float((uint32_t)(rect.x1 - rect.x0) |
                                               ((uint32_t)(rect.y1 - rect.y0) << 16)));
            p.rec[idx] = r;
            p.depth_bits[idx] = __float_as_uint(depth);
            p.clamped[idx] = (uint8_t)clamp_bits;
        }
    }
}

// K10 — markVisible / checkFrustum (rasterizer_impl.cu:54-66,141-153)
__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                    uint8_t* __restrict__ present) {
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float z = affine_row_ref(view, 2, means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    present[idx] = (z <= 0.2f) ? 0 : 1;
}

int launch_preprocess_fwd(const PreprocessParams& p, cudaStream_t stream) {
    if (p.P <= 0) return 0;
    const int blocks = (p.P + PRE_THREADS - 1) / PRE_THREADS;
    const bool sh16 = p.colors_precomp == nullptr && p.M == 16 &&
                      (reinterpret_cast<uintptr_t>(p.shs) % 16 == 0);
    if (sh16) {
        B200S_SET_FUNC_ATTR_ONCE(preprocess_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)PRE_SMEM_SH16);
        count_launch();
        preprocess_fwd_kernel<true><<<blocks, PRE_THREADS, PRE_SMEM_SH16, stream>>>(p);
    } else {
        count_launch();
        preprocess_fwd_kernel<false><<<blocks, PRE_THREADS, 0, stream>>>(p);
    }
    return (int)cudaGetLastError();
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present,
                        cudaStream_t stream) {
    if (P <= 0) return 0;
    count_launch();
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
    return (int)cudaGetLastError();
}

}  // namespace b200s
