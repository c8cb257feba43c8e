// Fused activation-backward + Adam + activation-forward for the replicated Gaussian parameters
// (the step right after the rasterizer's backward / gradient all-reduce).
//
// Reference: the rasterizer's gradients are w.r.t. ACTIVATED values (get_xyz / get_features /
// get_opacity / get_scaling / get_rotation, scene/gaussian_model.py:113-141); autograd chains them
// through sigmoid / exp / normalize to the raw parameters, torch.optim.Adam(eps=1e-15) with one
// param group (= learning rate) per tensor updates those (gaussian_model.py:252-279, train.py:127),
// and the next iteration re-applies the activations and torch.cat's f_dc/f_rest (a 576 MB copy at
// 3 M Gaussians). That is ~25 elementwise passes over 59 floats per Gaussian.
//
// Here: ONE launch, one pass. Per float4 of a tensor: read grad (straight from the peer-reduced
// gradient buffer), raw parameter, both moments; chain the gradient through the activation; Adam
// with torch's exact operation order (lerp / mul+addcmul / sqrt-div-add / addcdiv); write the
// parameter, the moments and the ACTIVATED value the next forward consumes. 8 x 4 bytes per
// float: 1,888 B per Gaussian — HBM-bound by construction.
//
// Tensors are described by B200sAdamTensor; `mode` selects the activation:
//   IDENTITY  activated == param (xyz); `activated` may be NULL
//   SH        identity, row of 48 floats: coefficient 0 (3 floats) uses neg_step, the other 15 use
//             neg_step_rest (f_dc / f_rest learning rates on the concatenated [P,16,3] layout,
//             so get_features' torch.cat disappears)
//   EXP       scaling      a = exp(p),          dL/dp = dL/da * a
//   SIGMOID   opacity      a = 1/(1+exp(-p)),   dL/dp = dL/da * a * (1 - a)
//   NORMALIZE rotation     a = q / max(|q|, 1e-12), dL/dq = (g - a (a.g)) / |q|   (one float4 per row)

#include "kernels.cuh"

namespace b200s {

constexpr int ADAM_THREADS = 256;

struct AdamParams {
    B200sAdamTensor t[B200S_ADAM_MAX_TENSORS];
    long long block_start[B200S_ADAM_MAX_TENSORS + 1];
    int count;
    float w1;        // 1 - beta1
    float beta2;
    float w2;        // 1 - beta2
    float eps;
    float bc2_sqrt;  // sqrt(1 - beta2^step)
};

__device__ __forceinline__ float adam_update(float& p, float& m, float& v, float g, float neg_step,
                                             const AdamParams& a) {
    m = m + a.w1 * (g - m);                        // exp_avg.lerp_(grad, 1 - beta1)
    v = v * a.beta2;                               // exp_avg_sq.mul_(beta2)
    v = v + (a.w2 * g) * g;                        //   .addcmul_(grad, grad, value = 1 - beta2)
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p + neg_step * (m / denom);                // param.addcdiv_(exp_avg, denom, value = -step_size)
    return p;
}

__global__ void __launch_bounds__(ADAM_THREADS)
adam_step_kernel(const __grid_constant__ AdamParams a) {
    int s = 0;
#pragma unroll
    for (int i = 1; i < B200S_ADAM_MAX_TENSORS; ++i)
        if (i < a.count && (long long)blockIdx.x >= a.block_start[i]) s = i;
    const B200sAdamTensor& t = a.t[s];
    const long long i4 = ((long long)blockIdx.x - a.block_start[s]) * ADAM_THREADS + threadIdx.x;
    const long long f0 = i4 * 4;
    if (f0 >= t.n) return;
    const int lanes = (int)min(4LL, t.n - f0);

    float g[4], p[4], m[4], v[4], act[4];
    if (lanes == 4) {
        const float4 G = __ldcs(reinterpret_cast<const float4*>(t.grad) + i4);
        const float4 P = __ldcs(reinterpret_cast<const float4*>(t.param) + i4);
        const float4 M = __ldcs(reinterpret_cast<const float4*>(t.exp_avg) + i4);
        const float4 V = __ldcs(reinterpret_cast<const float4*>(t.exp_avg_sq) + i4);
        g[0] = G.x; g[1] = G.y; g[2] = G.z; g[3] = G.w;
        p[0] = P.x; p[1] = P.y; p[2] = P.z; p[3] = P.w;
        m[0] = M.x; m[1] = M.y; m[2] = M.z; m[3] = M.w;
        v[0] = V.x; v[1] = V.y; v[2] = V.z; v[3] = V.w;
    } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const bool ok = c < lanes;
            g[c] = ok ? t.grad[f0 + c] : 0.0f;
            p[c] = ok ? t.param[f0 + c] : 0.0f;
            m[c] = ok ? t.exp_avg[f0 + c] : 0.0f;
            v[c] = ok ? t.exp_avg_sq[f0 + c] : 0.0f;
        }
    }

    float step[4] = {t.neg_step, t.neg_step, t.neg_step, t.neg_step};
    switch (t.mode) {
        case B200S_ADAM_SH: {
            const int r = (int)(f0 % 48);          // 48 floats per row, rows start on multiples of 4
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (r + c >= 3) step[c] = t.neg_step_rest;
            break;
        }
        case B200S_ADAM_EXP:
#pragma unroll
            for (int c = 0; c < 4; ++c) g[c] = g[c] * expf(p[c]);
            break;
        case B200S_ADAM_SIGMOID:
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float y = 1.0f / (1.0f + expf(-p[c]));
                g[c] = g[c] * ((1.0f - y) * y);
            }
            break;
        case B200S_ADAM_NORMALIZE: {
            const float n = fmaxf(sqrtf(p[0] * p[0] + p[1] * p[1] + p[2] * p[2] + p[3] * p[3]), 1e-12f);
            const float inv = 1.0f / n;
            const float y0 = p[0] * inv, y1 = p[1] * inv, y2 = p[2] * inv, y3 = p[3] * inv;
            const float d = y0 * g[0] + y1 * g[1] + y2 * g[2] + y3 * g[3];
            g[0] = (g[0] - y0 * d) * inv; g[1] = (g[1] - y1 * d) * inv;
            g[2] = (g[2] - y2 * d) * inv; g[3] = (g[3] - y3 * d) * inv;
            break;
        }
        default: break;
    }

#pragma unroll
    for (int c = 0; c < 4; ++c) adam_update(p[c], m[c], v[c], g[c], step[c], a);

    switch (t.mode) {
        case B200S_ADAM_EXP:
#pragma unroll
            for (int c = 0; c < 4; ++c) act[c] = expf(p[c]);
            break;
        case B200S_ADAM_SIGMOID:
#pragma unroll
            for (int c = 0; c < 4; ++c) act[c] = 1.0f / (1.0f + expf(-p[c]));
            break;
        case B200S_ADAM_NORMALIZE: {
            const float n = fmaxf(sqrtf(p[0] * p[0] + p[1] * p[1] + p[2] * p[2] + p[3] * p[3]), 1e-12f);
#pragma unroll
            for (int c = 0; c < 4; ++c) act[c] = p[c] / n;
            break;
        }
        default:
#pragma unroll
            for (int c = 0; c < 4; ++c) act[c] = p[c];
            break;
    }

    const bool write_act = t.activated != nullptr && t.activated != t.param;
    if (lanes == 4) {
        __stcs(reinterpret_cast<float4*>(t.param) + i4, make_float4(p[0], p[1], p[2], p[3]));
        __stcs(reinterpret_cast<float4*>(t.exp_avg) + i4, make_float4(m[0], m[1], m[2], m[3]));
        __stcs(reinterpret_cast<float4*>(t.exp_avg_sq) + i4, make_float4(v[0], v[1], v[2], v[3]));
        if (write_act) __stcs(reinterpret_cast<float4*>(t.activated) + i4, make_float4(act[0], act[1], act[2], act[3]));
    } else {
        for (int c = 0; c < lanes; ++c) {
            t.param[f0 + c] = p[c]; t.exp_avg[f0 + c] = m[c]; t.exp_avg_sq[f0 + c] = v[c];
            if (write_act) t.activated[f0 + c] = act[c];
        }
    }
}

int launch_adam_step(const B200sAdamArgs& in, cudaStream_t stream) {
    AdamParams a{};
    a.count = in.count;
    long long blocks = 0;
    for (int i = 0; i < in.count; ++i) {
        a.t[i] = in.t[i];
        a.block_start[i] = blocks;
        const long long n4 = (in.t[i].n + 3) / 4;
        blocks += (n4 + ADAM_THREADS - 1) / ADAM_THREADS;
    }
    a.block_start[in.count] = blocks;
    a.w1 = in.one_minus_beta1;
    a.beta2 = in.beta2;
    a.w2 = in.one_minus_beta2;
    a.eps = in.eps;
    a.bc2_sqrt = in.bias_correction2_sqrt;
    if (blocks == 0) return 0;
    if (blocks > 0x7fffffffLL) return B200S_E_BADARG;
    count_launch();
    adam_step_kernel<<<(unsigned)blocks, ADAM_THREADS, 0, stream>>>(a);
    return (int)cudaGetLastError();
}

}  // namespace b200s
