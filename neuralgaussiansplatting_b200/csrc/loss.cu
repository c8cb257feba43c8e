// Fused photometric loss: (1 - lambda) * L1 + lambda * (1 - SSIM), value and dL/dimage
// (the step between the rasterizer's forward and backward).
//
// Reference (utils/loss_utils.py:17-63, train.py:91-93): l1_loss = |x - y|.mean(); ssim = five
// depthwise 11x11 conv2d (zero padding 5) of x, y, x*x, y*y, x*y with the outer product of a
// normalised sigma-1.5 Gaussian, ten elementwise maps, a mean — and autograd runs all of it
// backwards: ~40 kernels and ~60 full-image passes per iteration.
//
// Here: two kernels. Forward, per 32x16 pixel tile and channel: stage x and y with a 5-pixel halo
// in shared memory, run the window separably (rows, then columns) for the five moments, evaluate
// the SSIM map and its partial derivatives w.r.t. mu1, E[x^2], E[xy] (three floats per pixel kept
// for the backward), and reduce sum(SSIM), sum|x-y| per block; a one-block finalize adds the block
// partials in index order (deterministic) into {loss, l1, ssim}. Backward: the same separable
// window over the three derivative maps (the window is symmetric, so correlation = convolution),
//   dL/dx = s * [ (1-lambda)/N * sign(x-y) - lambda/N * (W*dm1 + 2x * W*de11 + y * W*de12) ].

#include "kernels.cuh"

namespace b200s {

constexpr int LS_TX = 32, LS_TY = 16, LS_R = 5, LS_K = 2 * LS_R + 1;
constexpr int LS_HX = LS_TX + 2 * LS_R, LS_HY = LS_TY + 2 * LS_R;   // 42 x 26 staged pixels
constexpr int LS_THREADS = 256;

struct LossParams {
    int C, H, W;
    float lambda, inv_n, grad_scale;
    const float* x;
    const float* y;
    float* dm1;      // [C,H,W] dS/dmu1 (total, through sigma1_sq and sigma12 as well)
    float* de11;     // dS/dE[x^2]
    float* de12;     // dS/dE[xy]
    float* partial;  // [blocks][2]
    float* out;      // [3] loss, l1, ssim
    float* grad;     // [C,H,W]
    int blocks;
    float w[LS_K];
};

__global__ void __launch_bounds__(LS_THREADS)
loss_fwd_kernel(const __grid_constant__ LossParams p) {
    __shared__ float s_x[LS_HY][LS_HX + 1];
    __shared__ float s_y[LS_HY][LS_HX + 1];
    __shared__ float s_h[5][LS_HY][LS_TX + 1];
    __shared__ float s_red[2][LS_THREADS / 32];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LS_TX, y0 = blockIdx.y * LS_TY, c = blockIdx.z;
    const float* X = p.x + (size_t)c * p.H * p.W;
    const float* Y = p.y + (size_t)c * p.H * p.W;

    for (int i = tid; i < LS_HY * LS_HX; i += LS_THREADS) {
        const int r = i / LS_HX, q = i % LS_HX;
        const int gy = y0 + r - LS_R, gx = x0 + q - LS_R;
        const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
        s_x[r][q] = in ? X[(size_t)gy * p.W + gx] : 0.0f;     // conv2d zero padding
        s_y[r][q] = in ? Y[(size_t)gy * p.W + gx] : 0.0f;
    }
    __syncthreads();
    for (int i = tid; i < LS_HY * LS_TX; i += LS_THREADS) {
        const int r = i / LS_TX, q = i % LS_TX;
        float a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
#pragma unroll
        for (int k = 0; k < LS_K; ++k) {
            const float xv = s_x[r][q + k], yv = s_y[r][q + k], w = p.w[k];
            a0 += w * xv; a1 += w * yv; a2 += w * (xv * xv); a3 += w * (yv * yv); a4 += w * (xv * yv);
        }
        s_h[0][r][q] = a0; s_h[1][r][q] = a1; s_h[2][r][q] = a2; s_h[3][r][q] = a3; s_h[4][r][q] = a4;
    }
    __syncthreads();

    float sum_s = 0.0f, sum_l1 = 0.0f;
    const int tx = tid % LS_TX;
#pragma unroll
    for (int rep = 0; rep < LS_TY * LS_TX / LS_THREADS; ++rep) {
        const int ty = tid / LS_TX + rep * (LS_THREADS / LS_TX);
        const int gx = x0 + tx, gy = y0 + ty;
        if (gx >= p.W || gy >= p.H) continue;
        float m1 = 0, m2 = 0, e11 = 0, e22 = 0, e12 = 0;
#pragma unroll
        for (int k = 0; k < LS_K; ++k) {
            const float w = p.w[k];
            m1 += w * s_h[0][ty + k][tx]; m2 += w * s_h[1][ty + k][tx];
            e11 += w * s_h[2][ty + k][tx]; e22 += w * s_h[3][ty + k][tx]; e12 += w * s_h[4][ty + k][tx];
        }
        const float C1 = 0.0001f, C2 = 0.0009f;
        const float m11 = m1 * m1, m22 = m2 * m2, m12 = m1 * m2;
        const float s1 = e11 - m11, s2 = e22 - m22, s12 = e12 - m12;
        const float a = 2.0f * m12 + C1, b = 2.0f * s12 + C2, cc = m11 + m22 + C1, d = s1 + s2 + C2;
        const float inv_cd = 1.0f / (cc * d);
        const float S = a * b * inv_cd;
        sum_s += S;
        sum_l1 += fabsf(s_x[ty + LS_R][tx + LS_R] - s_y[ty + LS_R][tx + LS_R]);
        if (p.dm1) {
            const size_t o = ((size_t)c * p.H + gy) * p.W + gx;
            // S = a b / (c d); a, c depend on mu1 directly, b and d through sigma12 = e12 - mu1 mu2, sigma1 = e11 - mu1^2
            p.dm1[o] = 2.0f * m2 * (b - a) * inv_cd - S * 2.0f * m1 * (1.0f / cc - 1.0f / d);
            p.de11[o] = -S / d;
            p.de12[o] = 2.0f * a * inv_cd;
        }
    }
    // block reduction in a fixed order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum_s += __shfl_xor_sync(FULL_MASK, sum_s, o);
        sum_l1 += __shfl_xor_sync(FULL_MASK, sum_l1, o);
    }
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = sum_s; s_red[1][tid >> 5] = sum_l1; }
    __syncthreads();
    if (tid == 0) {
        float a = 0, b = 0;
        for (int w = 0; w < LS_THREADS / 32; ++w) { a += s_red[0][w]; b += s_red[1][w]; }
        const int blk = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.partial[2 * blk] = a;
        p.partial[2 * blk + 1] = b;
    }
}

// One block: sums the per-block partials in index order (double accumulators) -> loss, l1, ssim.
__global__ void __launch_bounds__(LS_THREADS)
loss_finalize_kernel(const __grid_constant__ LossParams p) {
    __shared__ double s_a[LS_THREADS], s_b[LS_THREADS];
    double a = 0.0, b = 0.0;
    const int per = (p.blocks + LS_THREADS - 1) / LS_THREADS;
    const int lo = threadIdx.x * per, hi = min(p.blocks, lo + per);
    for (int i = lo; i < hi; ++i) { a += (double)p.partial[2 * i]; b += (double)p.partial[2 * i + 1]; }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (int i = 0; i < LS_THREADS; ++i) { ta += s_a[i]; tb += s_b[i]; }
        const float ssim = (float)(ta * (double)p.inv_n), l1 = (float)(tb * (double)p.inv_n);
        p.out[0] = (1.0f - p.lambda) * l1 + p.lambda * (1.0f - ssim);
        p.out[1] = l1;
        p.out[2] = ssim;
    }
}

__global__ void __launch_bounds__(LS_THREADS)
loss_bwd_kernel(const __grid_constant__ LossParams p) {
    __shared__ float s_m[3][LS_HY][LS_HX + 1];
    __shared__ float s_h[3][LS_HY][LS_TX + 1];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LS_TX, y0 = blockIdx.y * LS_TY, c = blockIdx.z;
    const size_t plane = (size_t)c * p.H * p.W;
    for (int i = tid; i < LS_HY * LS_HX; i += LS_THREADS) {
        const int r = i / LS_HX, q = i % LS_HX;
        const int gy = y0 + r - LS_R, gx = x0 + q - LS_R;
        const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
        const size_t o = plane + (size_t)gy * p.W + gx;
        s_m[0][r][q] = in ? p.dm1[o] : 0.0f;
        s_m[1][r][q] = in ? p.de11[o] : 0.0f;
        s_m[2][r][q] = in ? p.de12[o] : 0.0f;
    }
    __syncthreads();
    for (int i = tid; i < LS_HY * LS_TX; i += LS_THREADS) {
        const int r = i / LS_TX, q = i % LS_TX;
        float a0 = 0, a1 = 0, a2 = 0;
#pragma unroll
        for (int k = 0; k < LS_K; ++k) {
            const float w = p.w[k];
            a0 += w * s_m[0][r][q + k]; a1 += w * s_m[1][r][q + k]; a2 += w * s_m[2][r][q + k];
        }
        s_h[0][r][q] = a0; s_h[1][r][q] = a1; s_h[2][r][q] = a2;
    }
    __syncthreads();
    const int tx = tid % LS_TX;
    const float k_l1 = (1.0f - p.lambda) * p.inv_n * p.grad_scale;
    const float k_ss = -p.lambda * p.inv_n * p.grad_scale;
#pragma unroll
    for (int rep = 0; rep < LS_TY * LS_TX / LS_THREADS; ++rep) {
        const int ty = tid / LS_TX + rep * (LS_THREADS / LS_TX);
        const int gx = x0 + tx, gy = y0 + ty;
        if (gx >= p.W || gy >= p.H) continue;
        float c0 = 0, c1 = 0, c2 = 0;
#pragma unroll
        for (int k = 0; k < LS_K; ++k) {
            const float w = p.w[k];
            c0 += w * s_h[0][ty + k][tx]; c1 += w * s_h[1][ty + k][tx]; c2 += w * s_h[2][ty + k][tx];
        }
        const size_t o = plane + (size_t)gy * p.W + gx;
        const float xv = p.x[o], yv = p.y[o];
        const float df = xv - yv;
        const float sgn = df > 0.0f ? 1.0f : (df < 0.0f ? -1.0f : 0.0f);
        p.grad[o] = k_l1 * sgn + k_ss * (c0 + 2.0f * xv * c1 + yv * c2);
    }
}

static inline int loss_blocks(int C, int H, int W) {
    return ((W + LS_TX - 1) / LS_TX) * ((H + LS_TY - 1) / LS_TY) * C;
}

size_t loss_workspace_size(int C, int H, int W) {
    const size_t px = (size_t)C * H * W;
    return align_up(3 * px * sizeof(float), 256) + align_up((size_t)loss_blocks(C, H, W) * 2 * sizeof(float), 256);
}

int launch_l1_ssim(const B200sLossArgs& a, cudaStream_t stream) {
    LossParams p{};
    p.C = a.C; p.H = a.H; p.W = a.W;
    p.lambda = a.lambda_dssim;
    p.inv_n = (float)(1.0 / ((double)a.C * a.H * a.W));
    p.grad_scale = a.grad_scale;
    p.x = a.image; p.y = a.gt;
    const size_t px = (size_t)a.C * a.H * a.W;
    float* ws = (float*)a.ws;
    p.dm1 = a.dL_dimage ? ws : nullptr;
    p.de11 = ws + px; p.de12 = ws + 2 * px;
    p.partial = (float*)((char*)a.ws + align_up(3 * px * sizeof(float), 256));
    p.out = a.loss_out; p.grad = a.dL_dimage;
    p.blocks = loss_blocks(a.C, a.H, a.W);
    // the reference's window: exp(-(i-5)^2 / (2 * 1.5^2)), normalised, rounded to float (loss_utils.py:23-25)
    double g[LS_K], sum = 0.0;
    for (int i = 0; i < LS_K; ++i) { g[i] = exp(-(double)((i - LS_R) * (i - LS_R)) / (2.0 * 1.5 * 1.5)); }
    float gf[LS_K]; float fsum = 0.0f;
    for (int i = 0; i < LS_K; ++i) { gf[i] = (float)g[i]; fsum += gf[i]; }   // torch.Tensor([...]) is float32; .sum() in float
    (void)sum;
    for (int i = 0; i < LS_K; ++i) p.w[i] = gf[i] / fsum;
    const dim3 grid((a.W + LS_TX - 1) / LS_TX, (a.H + LS_TY - 1) / LS_TY, a.C);
    count_launch(2);
    loss_fwd_kernel<<<grid, LS_THREADS, 0, stream>>>(p);
    loss_finalize_kernel<<<1, LS_THREADS, 0, stream>>>(p);
    if (a.dL_dimage) {
        count_launch();
        loss_bwd_kernel<<<grid, LS_THREADS, 0, stream>>>(p);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200s
