// SH basis and view direction shared by preprocess_bwd (which emits dL/dsh[k] = basis_k * dL/dRGB per view,
// reference backward.cu:20-139) and the factored-gradient reconstruction (sh_reconstruct.cu), which must
// reproduce those products bit for bit from the 3-float factor: every operation is an explicit single-rounding
// intrinsic, so both kernels compile to the same arithmetic regardless of FMA contraction choices.
#pragma once
#include "common.cuh"

namespace b200s {

__device__ const float gSH_C0 = 0.28209479177387814f;
__device__ const float gSH_C1 = 0.4886025119029199f;
__device__ const float gSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                    -1.0925484305920792f, 0.5462742152960396f};
__device__ const float gSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                    0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                    -0.5900435899266435f};

// dir_orig = mean - campos, dir = dir_orig / |dir_orig| (forward.cu:27-29 as the backward recomputes it,
// backward.cu:28-29).
__device__ __forceinline__ void sh_view_dir(float mx, float my, float mz, float cx, float cy, float cz,
                                            float3& dir_orig, float3& dir) {
    dir_orig = make_float3(fsub(mx, cx), fsub(my, cy), fsub(mz, cz));
    const float len2 = ffma(dir_orig.z, dir_orig.z, ffma(dir_orig.y, dir_orig.y, fmul(dir_orig.x, dir_orig.x)));
    const float len = __fsqrt_rn(len2);
    dir = make_float3(__fdiv_rn(dir_orig.x, len), __fdiv_rn(dir_orig.y, len), __fdiv_rn(dir_orig.z, len));
}

// b[k] = d RGB / d sh[k] for the active degree (backward.cu:41-110); entries above the degree are left untouched.
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float (&b)[16]) {
    b[0] = gSH_C0;
    if (deg > 0) {
        b[1] = fmul(-gSH_C1, y);
        b[2] = fmul(gSH_C1, z);
        b[3] = fmul(-gSH_C1, x);
        if (deg > 1) {
            const float xx = fmul(x, x), yy = fmul(y, y), zz = fmul(z, z);
            const float xy = fmul(x, y), yz = fmul(y, z), xz = fmul(x, z);
            b[4] = fmul(gSH_C2[0], xy);
            b[5] = fmul(gSH_C2[1], yz);
            b[6] = fmul(gSH_C2[2], fsub(ffma(2.0f, zz, -xx), yy));
            b[7] = fmul(gSH_C2[3], xz);
            b[8] = fmul(gSH_C2[4], fsub(xx, yy));
            if (deg > 2) {
                b[9] = fmul(fmul(gSH_C3[0], y), ffma(3.0f, xx, -yy));
                b[10] = fmul(fmul(gSH_C3[1], xy), z);
                b[11] = fmul(fmul(gSH_C3[2], y), fsub(ffma(4.0f, zz, -xx), yy));
                b[12] = fmul(fmul(gSH_C3[3], z), ffma(-3.0f, yy, ffma(-3.0f, xx, fmul(2.0f, zz))));
                b[13] = fmul(fmul(gSH_C3[4], x), fsub(ffma(4.0f, zz, -xx), yy));
                b[14] = fmul(fmul(gSH_C3[5], z), fsub(xx, yy));
                b[15] = fmul(fmul(gSH_C3[6], x), ffma(-3.0f, yy, xx));
            }
        }
    }
}

constexpr int SH_ROW_FLOATS = 48;   // 16 coefficients x RGB
constexpr int SH_ROW_PAD = 52;      // shared-memory row pitch: 16-B aligned, conflict-free float4 access

}  // namespace b200s
