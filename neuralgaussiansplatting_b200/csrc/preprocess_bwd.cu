// K8+K9 fused — backward of the preprocess: conic -> cov2D -> (cov3D, mean), projected-mean
// gradient, SH backward (incl. view direction -> mean), cov3D -> (scale, quaternion).
// Replaces computeCov2DCUDA + preprocessCUDA<3> (backward) + computeColorFromSH (backward) +
// computeCov3D (backward) (reference backward.cu:20-139,144-274,278-341,346-396).
//
// One kernel instead of two: dL/dcov3D never round-trips through HBM, and the kernel writes
// EVERY element of every output (zeros for culled Gaussians), so the nine zero-filled
// tensors the reference allocates per step (rasterize_points.cu:151-159) become plain
// uninitialised allocations. SH rows are staged through shared memory with coalesced 16-B
// transfers in both directions (read via cp.async, dL/dsh written back as full 128-B lines).

#include <cstdlib>
#include "kernels.cuh"
#include "sh_grad.cuh"

namespace b200s {


constexpr int PB_THREADS = 256;
constexpr int PB_ROW_FLOATS = 48;
constexpr int PB_ROW_PAD = 52;
constexpr size_t PB_SMEM_SH16 = (size_t)PB_THREADS * PB_ROW_PAD * sizeof(float);

#define bSH_C0 gSH_C0
#define bSH_C1 gSH_C1
#define bSH_C2 gSH_C2
#define bSH_C3 gSH_C3

__device__ __forceinline__ void pb_cp_async16(void* smem_dst, const void* gmem_src) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}

__device__ __forceinline__ float3 f3(float a, float b, float c) { return make_float3(a, b, c); }
__device__ __forceinline__ float3 operator*(float s, float3 v) { return f3(s * v.x, s * v.y, s * v.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// SH backward for one Gaussian (backward.cu:20-139). get(k) returns coefficient k (RGB), put(k, v)
// stores dL/dsh[k]. Returns dL/d(direction) (before the normalisation Jacobian).
template <typename Get, typename Put>
__device__ __forceinline__ float3 sh_backward(int deg, float x, float y, float z, float3 dRGB, Get get, Put put) {
    float3 dx = f3(0, 0, 0), dy = f3(0, 0, 0), dz = f3(0, 0, 0);
    // the coefficients are consumed first (get) and dL/dsh produced last (put), so that put may overwrite the
    // storage get reads from and the 48 coefficients never have to be live in registers together with the basis
    if (deg > 0) {
        const float3 s1 = get(1), s2 = get(2), s3 = get(3);
        dx = (-bSH_C1) * s3;
        dy = (-bSH_C1) * s1;
        dz = bSH_C1 * s2;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const float3 s4 = get(4), s5 = get(5), s6 = get(6), s7 = get(7), s8 = get(8);
            dx = dx + (bSH_C2[0] * y) * s4 + (bSH_C2[2] * 2.f * -x) * s6 + (bSH_C2[3] * z) * s7 + (bSH_C2[4] * 2.f * x) * s8;
            dy = dy + (bSH_C2[0] * x) * s4 + (bSH_C2[1] * z) * s5 + (bSH_C2[2] * 2.f * -y) * s6 + (bSH_C2[4] * 2.f * -y) * s8;
            dz = dz + (bSH_C2[1] * y) * s5 + (bSH_C2[2] * 2.f * 2.f * z) * s6 + (bSH_C2[3] * x) * s7;
            if (deg > 2) {
                const float3 s9 = get(9), s10 = get(10), s11 = get(11), s12 = get(12), s13 = get(13),
                             s14 = get(14), s15 = get(15);
                dx = dx + (bSH_C3[0] * 3.f * 2.f * xy) * s9 + (bSH_C3[1] * yz) * s10 + (bSH_C3[2] * -2.f * xy) * s11 +
                     (bSH_C3[3] * -3.f * 2.f * xz) * s12 + (bSH_C3[4] * (-3.f * xx + 4.f * zz - yy)) * s13 +
                     (bSH_C3[5] * 2.f * xz) * s14 + (bSH_C3[6] * 3.f * (xx - yy)) * s15;
                dy = dy + (bSH_C3[0] * 3.f * (xx - yy)) * s9 + (bSH_C3[1] * xz) * s10 +
                     (bSH_C3[2] * (-3.f * yy + 4.f * zz - xx)) * s11 + (bSH_C3[3] * -3.f * 2.f * yz) * s12 +
                     (bSH_C3[4] * -2.f * xy) * s13 + (bSH_C3[5] * -2.f * yz) * s14 + (bSH_C3[6] * -3.f * 2.f * xy) * s15;
                dz = dz + (bSH_C3[1] * xy) * s10 + (bSH_C3[2] * 4.f * 2.f * yz) * s11 +
                     (bSH_C3[3] * 3.f * (2.f * zz - xx - yy)) * s12 + (bSH_C3[4] * 4.f * 2.f * xz) * s13 +
                     (bSH_C3[5] * (xx - yy)) * s14;
            }
        }
    }
    // dL/dsh[k] = basis_k * dL/dRGB: single-rounding products of the shared basis (sh_grad.cuh), which the
    // factored-gradient reconstruction reproduces bit for bit; coefficients above the active degree get zeros
    float bas[16];
    sh_basis(deg, x, y, z, bas);
    const int ncoef = (deg + 1) * (deg + 1);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        put(k, k < ncoef ? f3(fmul(bas[k], dRGB.x), fmul(bas[k], dRGB.y), fmul(bas[k], dRGB.z)) : f3(0, 0, 0));
    return f3(dot(dx, dRGB), dot(dy, dRGB), dot(dz, dRGB));
}

template <bool SH16, int MINB>
__global__ void __launch_bounds__(PB_THREADS, MINB)
preprocess_bwd_kernel(const __grid_constant__ PreprocessBwdParams p) {
    extern __shared__ __align__(16) float sh_tile[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t idx = (int64_t)blockIdx.x * PB_THREADS + tid;
    const bool in_range = idx < p.P;
    const bool visible = in_range && p.radii[idx] > 0;
    const uint32_t vis_mask = __ballot_sync(FULL_MASK, visible);
    const bool use_sh = p.shs != nullptr && p.colors_precomp == nullptr;

    float* tile = nullptr;
    if (SH16) {
        tile = sh_tile + (size_t)warp * 32 * PB_ROW_PAD;
        const float* gsrc = p.shs + ((int64_t)blockIdx.x * PB_THREADS + warp * 32) * PB_ROW_FLOATS;
        if (vis_mask != 0) {
#pragma unroll
            for (int it = 0; it < 12; ++it) {
                const int c = it * 32 + lane;
                const int row = c / 12;
                const int ch = c - row * 12;
                if ((vis_mask >> row) & 1u) pb_cp_async16(tile + row * PB_ROW_PAD + ch * 4, gsrc + (size_t)c * 4);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
    }

    float3 g_mean = f3(0, 0, 0), g_m2d = f3(0, 0, 0), g_scale = f3(0, 0, 0), g_col = f3(0, 0, 0);
    float4 g_rot = make_float4(0, 0, 0, 0);
    float g_op = 0.f;
    float g_cov[6] = {0, 0, 0, 0, 0, 0};
    float g_con[3] = {0, 0, 0};
    float3 dir = f3(0, 0, 0), dir_orig = f3(0, 0, 0);
    uint32_t clamp_bits = 0;

    if (visible) {
        // gather the blend's raw sums
        float a9[9];
        if (p.partial != nullptr) {
            // deterministic: sum this Gaussian's instance slots in tile order
#pragma unroll
            for (int i = 0; i < 9; ++i) a9[i] = 0.f;
            const uint32_t first = idx == 0 ? 0u : p.point_offsets[idx - 1];
            const uint32_t end = p.point_offsets[idx];
            for (uint32_t u = first; u < end; ++u) {
                const float4* s = reinterpret_cast<const float4*>(p.partial + (size_t)u * ACC_STRIDE);
                const float4 v0 = s[0], v1 = s[1], v2 = s[2];
                a9[0] += v0.x; a9[1] += v0.y; a9[2] += v0.z; a9[3] += v0.w;
                a9[4] += v1.x; a9[5] += v1.y; a9[6] += v1.z; a9[7] += v1.w; a9[8] += v2.x;
            }
        } else {
            const float4* s = reinterpret_cast<const float4*>(p.accum + (size_t)idx * ACC_STRIDE);
            const float4 v0 = s[0], v1 = s[1], v2 = s[2];
            a9[0] = v0.x; a9[1] = v0.y; a9[2] = v0.z; a9[3] = v0.w;
            a9[4] = v1.x; a9[5] = v1.y; a9[6] = v1.z; a9[7] = v1.w; a9[8] = v2.x;
        }
        const SplatRec* r = p.rec + idx;
        const float4 q0 = r->q0;
        const float con_c = r->q1.x;
        const float con_a = q0.z, con_b = q0.w;
        g_col = f3(a9[0], a9[1], a9[2]);
        g_op = a9[8];
        // dL/dmean2D (backward.cu:545-546): ddelx_dx = 0.5*W, ddely_dy = 0.5*H
        g_m2d.x = -(con_a * a9[3] + con_b * a9[4]) * (0.5f * (float)p.W);
        g_m2d.y = -(con_c * a9[4] + con_b * a9[3]) * (0.5f * (float)p.H);
        // dL/dconic (backward.cu:549-551)
        g_con[0] = -0.5f * a9[5]; g_con[1] = -0.5f * a9[6]; g_con[2] = -0.5f * a9[7];

        const float mx = p.means3D[3 * idx], my = p.means3D[3 * idx + 1], mz = p.means3D[3 * idx + 2];
        float cov3D[6];
        float qr = 0, qx = 0, qy = 0, qz = 0, sx = 0, sy = 0, sz = 0;
        if (p.cov3D_precomp != nullptr) {
#pragma unroll
            for (int i = 0; i < 6; ++i) cov3D[i] = p.cov3D_precomp[6 * idx + i];
        } else {
            const float4 q = *reinterpret_cast<const float4*>(p.rotations + 4 * idx);
            qr = q.x; qx = q.y; qy = q.z; qz = q.w;
            sx = p.scales[3 * idx]; sy = p.scales[3 * idx + 1]; sz = p.scales[3 * idx + 2];
            cov3d_ref(sx, sy, sz, p.scale_modifier, qr, qx, qy, qz, cov3D);
        }
        const float* view = p.viewmatrix;
        const Cov2D c2 = cov2d_ref(mx, my, mz, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov3D, view);

        // ---- computeCov2DCUDA (backward.cu:144-274)
        const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
        const float x_grad_mul = (c2.txtz < -limx || c2.txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (c2.tytz < -limy || c2.tytz > limy) ? 0.f : 1.f;
        const float a = c2.a, b = c2.b, c = c2.c;
        const float denom = a * c - b * b;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float T00 = c2.T00, T01 = c2.T01, T02 = c2.T02, T10 = c2.T10, T11 = c2.T11, T12 = c2.T12;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * g_con[0] + 2 * b * c * g_con[1] + (denom - a * c) * g_con[2]);
            dL_dc = denom2inv * (-a * a * g_con[2] + 2 * a * b * g_con[1] + (denom - a * c) * g_con[0]);
            dL_db = denom2inv * 2 * (b * c * g_con[0] - (denom + 2 * b * b) * g_con[1] + a * b * g_con[2]);
            g_cov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            g_cov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            g_cov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            g_cov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
            g_cov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
            g_cov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
        }
        // V = Vrk (symmetric)
        const float V00 = cov3D[0], V01 = cov3D[1], V02 = cov3D[2], V11 = cov3D[3], V12 = cov3D[4], V22 = cov3D[5];
        const float r0x = T00 * V00 + T01 * V01 + T02 * V02;   // row 0 of T against V columns
        const float r0y = T00 * V01 + T01 * V11 + T02 * V12;
        const float r0z = T00 * V02 + T01 * V12 + T02 * V22;
        const float r1x = T10 * V00 + T11 * V01 + T12 * V02;
        const float r1y = T10 * V01 + T11 * V11 + T12 * V12;
        const float r1z = T10 * V02 + T11 * V12 + T12 * V22;
        const float dL_dT00 = 2 * r0x * dL_da + r1x * dL_db;
        const float dL_dT01 = 2 * r0y * dL_da + r1y * dL_db;
        const float dL_dT02 = 2 * r0z * dL_da + r1z * dL_db;
        const float dL_dT10 = 2 * r1x * dL_dc + r0x * dL_db;
        const float dL_dT11 = 2 * r1y * dL_dc + r0y * dL_db;
        const float dL_dT12 = 2 * r1z * dL_dc + r0z * dL_db;
        // W columns: W[0] = (v0,v4,v8), W[1] = (v1,v5,v9), W[2] = (v2,v6,v10)
        const float dL_dJ00 = view[0] * dL_dT00 + view[4] * dL_dT01 + view[8] * dL_dT02;
        const float dL_dJ02 = view[2] * dL_dT00 + view[6] * dL_dT01 + view[10] * dL_dT02;
        const float dL_dJ11 = view[1] * dL_dT10 + view[5] * dL_dT11 + view[9] * dL_dT12;
        const float dL_dJ12 = view[2] * dL_dT10 + view[6] * dL_dT11 + view[10] * dL_dT12;
        const float tz = 1.f / c2.tz, tz2 = tz * tz, tz3 = tz2 * tz;
        const float h_x = p.focal_x, h_y = p.focal_y;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * c2.tx) * tz3 * dL_dJ02 +
                             (2 * h_y * c2.ty) * tz3 * dL_dJ12;
        g_mean.x = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
        g_mean.y = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
        g_mean.z = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;

        // ---- projected mean (backward.cu:370-387)
        const float* proj = p.projmatrix;
        const float m_w = 1.0f / (proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15] + 0.0000001f);
        const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
        g_mean.x += (proj[0] * m_w - proj[3] * mul1) * g_m2d.x + (proj[1] * m_w - proj[3] * mul2) * g_m2d.y;
        g_mean.y += (proj[4] * m_w - proj[7] * mul1) * g_m2d.x + (proj[5] * m_w - proj[7] * mul2) * g_m2d.y;
        g_mean.z += (proj[8] * m_w - proj[11] * mul1) * g_m2d.x + (proj[9] * m_w - proj[11] * mul2) * g_m2d.y;

        // ---- cov3D -> scale / rotation (backward.cu:278-341)
        if (p.cov3D_precomp == nullptr) {
            const float s0 = p.scale_modifier * sx, s1 = p.scale_modifier * sy, s2 = p.scale_modifier * sz;
            // rotation entries R[c][r] (GLM column-major)
            const float R00 = 1.f - 2.f * (qy * qy + qz * qz), R01 = 2.f * (qx * qy - qr * qz), R02 = 2.f * (qx * qz + qr * qy);
            const float R10 = 2.f * (qx * qy + qr * qz), R11 = 1.f - 2.f * (qx * qx + qz * qz), R12 = 2.f * (qy * qz - qr * qx);
            const float R20 = 2.f * (qx * qz - qr * qy), R21 = 2.f * (qy * qz + qr * qx), R22 = 1.f - 2.f * (qx * qx + qy * qy);
            // M = S*R: M[c][r] = s_r * R[c][r]
            const float M00 = s0 * R00, M01 = s1 * R01, M02 = s2 * R02;
            const float M10 = s0 * R10, M11 = s1 * R11, M12 = s2 * R12;
            const float M20 = s0 * R20, M21 = s1 * R21, M22 = s2 * R22;
            // dL_dSigma (symmetric, off-diagonals halved), D[c][r]
            const float D00 = g_cov[0], D01 = 0.5f * g_cov[1], D02 = 0.5f * g_cov[2];
            const float D11 = g_cov[3], D12 = 0.5f * g_cov[4], D22 = g_cov[5];
            // dL_dM = 2 * M * dL_dSigma  (GLM: (A*B)[c][r] = sum_k A[k][r] * B[c][k])
            const float dM00 = 2.f * (M00 * D00 + M10 * D01 + M20 * D02);
            const float dM01 = 2.f * (M01 * D00 + M11 * D01 + M21 * D02);
            const float dM02 = 2.f * (M02 * D00 + M12 * D01 + M22 * D02);
            const float dM10 = 2.f * (M00 * D01 + M10 * D11 + M20 * D12);
            const float dM11 = 2.f * (M01 * D01 + M11 * D11 + M21 * D12);
            const float dM12 = 2.f * (M02 * D01 + M12 * D11 + M22 * D12);
            const float dM20 = 2.f * (M00 * D02 + M10 * D12 + M20 * D22);
            const float dM21 = 2.f * (M01 * D02 + M11 * D12 + M21 * D22);
            const float dM22 = 2.f * (M02 * D02 + M12 * D12 + M22 * D22);
            // Rt[i] = (R[0][i], R[1][i], R[2][i]); dL_dMt[i] = (dM[0][i], dM[1][i], dM[2][i])
            g_scale.x = R00 * dM00 + R10 * dM10 + R20 * dM20;
            g_scale.y = R01 * dM01 + R11 * dM11 + R21 * dM21;
            g_scale.z = R02 * dM02 + R12 * dM12 + R22 * dM22;
            // dL_dMt[i] *= s_i ; t[i][j] = dL_dMt[i][j] = dM[j][i] * s_i
            const float t00 = dM00 * s0, t01 = dM10 * s0, t02 = dM20 * s0;
            const float t10 = dM01 * s1, t11 = dM11 * s1, t12 = dM21 * s1;
            const float t20 = dM02 * s2, t21 = dM12 * s2, t22 = dM22 * s2;
            g_rot.x = 2 * qz * (t01 - t10) + 2 * qy * (t20 - t02) + 2 * qx * (t12 - t21);
            g_rot.y = 2 * qy * (t10 + t01) + 2 * qz * (t20 + t02) + 2 * qr * (t12 - t21) - 4 * qx * (t22 + t11);
            g_rot.z = 2 * qx * (t10 + t01) + 2 * qr * (t20 - t02) + 2 * qz * (t12 + t21) - 4 * qy * (t22 + t00);
            g_rot.w = 2 * qr * (t01 - t10) + 2 * qx * (t20 + t02) + 2 * qy * (t12 + t21) - 4 * qz * (t11 + t00);
        }
        if (use_sh) {
            sh_view_dir(mx, my, mz, p.campos[0], p.campos[1], p.campos[2], dir_orig, dir);
            clamp_bits = p.clamped[idx];
        }
    }

    // ---- SH backward + dL/dsh output (all rows of the block are written)
    const bool write_rows = p.dL_dsh != nullptr;
    if (use_sh && (write_rows || p.dL_dsh_factors != nullptr)) {
        float3 dRGB = g_col;
        if (clamp_bits & 1u) dRGB.x = 0.f;
        if (clamp_bits & 2u) dRGB.y = 0.f;
        if (clamp_bits & 4u) dRGB.z = 0.f;
        if (p.dL_dsh_factors != nullptr) {
            // factored SH gradient: the 3-float factor of this view's rank-1 dL/dsh row (+ the camera centre)
            if (in_range) {
                p.dL_dsh_factors[3 * idx] = dRGB.x; p.dL_dsh_factors[3 * idx + 1] = dRGB.y; p.dL_dsh_factors[3 * idx + 2] = dRGB.z;
            }
            if (blockIdx.x == 0 && tid < 4) p.dL_dsh_factors[3 * (int64_t)p.P + tid] = tid < 3 ? p.campos[tid] : 0.f;
        }
        if (SH16) {
            if (vis_mask != 0) asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            float* row = tile + lane * PB_ROW_PAD;
            if (!visible && write_rows) {
#pragma unroll
                for (int k = 0; k < 12; ++k) reinterpret_cast<float4*>(row)[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (visible) {
                // the lane's row is read with 16-byte loads (conflict-free at the 52-float pitch), four coefficients
                // at a time; its shared-memory slot is then reused for dL/dsh
                float o[PB_ROW_FLOATS];
                const float3 dd = sh_backward(
                    p.D, dir.x, dir.y, dir.z, dRGB,
                    [&](int k) {
                        const float4 t0 = reinterpret_cast<const float4*>(row)[(3 * k) >> 2];
                        const float4 t1 = reinterpret_cast<const float4*>(row)[(3 * k + 2) >> 2];
                        const int r = (3 * k) & 3;   // offset of the coefficient's first float inside t0
                        return r == 0 ? f3(t0.x, t0.y, t0.z) : r == 1 ? f3(t0.y, t0.z, t0.w)
                             : r == 2 ? f3(t0.z, t0.w, t1.x) : f3(t0.w, t1.x, t1.y);
                    },
                    [&](int k, float3 g) { o[3 * k] = g.x; o[3 * k + 1] = g.y; o[3 * k + 2] = g.z; });
                if (write_rows) {
#pragma unroll
                    for (int k = 0; k < 12; ++k)
                        reinterpret_cast<float4*>(row)[k] = make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
                }
                // dnormvdv (auxiliary.h:107-117)
                const float3 vv = dir_orig;
                const float sum2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
                const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
                g_mean.x += ((sum2 - vv.x * vv.x) * dd.x - vv.y * vv.x * dd.y - vv.z * vv.x * dd.z) * inv;
                g_mean.y += (-vv.x * vv.y * dd.x + (sum2 - vv.y * vv.y) * dd.y - vv.z * vv.y * dd.z) * inv;
                g_mean.z += (-vv.x * vv.z * dd.x - vv.y * vv.z * dd.y + (sum2 - vv.z * vv.z) * dd.z) * inv;
            }
            __syncwarp();
            if (write_rows) {
                // coalesced write-back of the warp's 32 rows (6 KB contiguous)
                const int64_t gbase = ((int64_t)blockIdx.x * PB_THREADS + warp * 32);
                float* gdst = p.dL_dsh + gbase * PB_ROW_FLOATS;
#pragma unroll
                for (int it = 0; it < 12; ++it) {
                    const int c = it * 32 + lane;
                    const int r = c / 12;
                    const int ch = c - r * 12;
                    if (gbase + r < p.P)
                        reinterpret_cast<float4*>(gdst)[c] = *reinterpret_cast<const float4*>(tile + r * PB_ROW_PAD + ch * 4);
                }
            }
        } else if (in_range) {
            float* out = write_rows ? p.dL_dsh + (int64_t)idx * p.M * 3 : nullptr;
            if (write_rows)
                for (int k = 0; k < p.M * 3; ++k) out[k] = 0.f;
            if (visible) {
                const float* sh = p.shs + (int64_t)idx * p.M * 3;
                const float3 dd = sh_backward(
                    p.D, dir.x, dir.y, dir.z, dRGB, [&](int k) { return f3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]); },
                    // sh_backward offers all 16 coefficients; a row of this tensor holds only M of them (M = 1 / 4 / 9 for
                    // models trained at --sh_degree 0 / 1 / 2) and has been zeroed above
                    [&](int k, float3 g) { if (write_rows && k < p.M) { out[3 * k] = g.x; out[3 * k + 1] = g.y; out[3 * k + 2] = g.z; } });
                const float3 vv = dir_orig;
                const float sum2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
                const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
                g_mean.x += ((sum2 - vv.x * vv.x) * dd.x - vv.y * vv.x * dd.y - vv.z * vv.x * dd.z) * inv;
                g_mean.y += (-vv.x * vv.y * dd.x + (sum2 - vv.y * vv.y) * dd.y - vv.z * vv.y * dd.z) * inv;
                g_mean.z += (-vv.x * vv.z * dd.x - vv.y * vv.z * dd.y + (sum2 - vv.z * vv.z) * dd.z) * inv;
            }
        }
    } else if (SH16 && vis_mask != 0) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }

    if (in_range) {
        p.dL_dmeans3D[3 * idx] = g_mean.x; p.dL_dmeans3D[3 * idx + 1] = g_mean.y; p.dL_dmeans3D[3 * idx + 2] = g_mean.z;
        p.dL_dmeans2D[3 * idx] = g_m2d.x; p.dL_dmeans2D[3 * idx + 1] = g_m2d.y; p.dL_dmeans2D[3 * idx + 2] = 0.f;
        p.dL_dopacity[idx] = g_op;
        if (p.dL_dscales) { p.dL_dscales[3 * idx] = g_scale.x; p.dL_dscales[3 * idx + 1] = g_scale.y; p.dL_dscales[3 * idx + 2] = g_scale.z; }
        if (p.dL_drotations) *reinterpret_cast<float4*>(p.dL_drotations + 4 * idx) = g_rot;
        if (p.dL_dcolors) { p.dL_dcolors[3 * idx] = g_col.x; p.dL_dcolors[3 * idx + 1] = g_col.y; p.dL_dcolors[3 * idx + 2] = g_col.z; }
        if (p.dL_dcov3D) {
#pragma unroll
            for (int i = 0; i < 6; ++i) p.dL_dcov3D[6 * idx + i] = g_cov[i];
        }
        if (p.dL_dconic) *reinterpret_cast<float4*>(p.dL_dconic + 4 * idx) = make_float4(g_con[0], g_con[1], 0.f, g_con[2]);
    }
}

int launch_preprocess_bwd(const PreprocessBwdParams& p, cudaStream_t stream) {
    if (p.P <= 0) return 0;
    const int blocks = (p.P + PB_THREADS - 1) / PB_THREADS;
    const bool sh16 = p.shs != nullptr && p.colors_precomp == nullptr &&
                      (p.dL_dsh != nullptr || p.dL_dsh_factors != nullptr) && p.M == 16 &&
                      (reinterpret_cast<uintptr_t>(p.shs) % 16 == 0) &&
                      (reinterpret_cast<uintptr_t>(p.dL_dsh) % 16 == 0);
    if (sh16) {
        static int minb = [] { const char* v = getenv("B200S_PB_MINB"); return v && *v ? atoi(v) : 3; }();
        auto kern = minb >= 4 ? preprocess_bwd_kernel<true, 4> : minb == 3 ? preprocess_bwd_kernel<true, 3>
                                                                            : preprocess_bwd_kernel<true, 2>;
        B200S_SET_FUNC_ATTR_ONCE(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PB_SMEM_SH16);
        count_launch();
        kern<<<blocks, PB_THREADS, PB_SMEM_SH16, stream>>>(p);
    } else {
        count_launch();
        preprocess_bwd_kernel<false, 2><<<blocks, PB_THREADS, 0, stream>>>(p);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200s
