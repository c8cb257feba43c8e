// SM-side peaks of this B200 for the blend-kernel rooflines: FP32 FMA, MUFU.EX2,
// shared-memory LDS.128 (broadcast, as the blend kernels read a staged Gaussian), warp shuffles, and the
// instruction-issue rate. Register-resident chains with 8 independent accumulators per thread, 148 x 8 CTAs of
// 256 threads, timed with CUDA events after a warm-up; prints one JSON object.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/sm_peaks tools/sm_peaks.cu && tools/_bin/sm_peaks
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int THREADS = 256;

__global__ void k_ffma(float* out, float a, float b) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = fmaf(x[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}

__global__ void k_ex2(float* out, float a) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = -(threadIdx.x * 1e-3f + i) * a;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}

__global__ void k_lds(float* out, int stride) {
    __shared__ float4 sm[256];
    sm[threadIdx.x] = make_float4(threadIdx.x, 1, 2, 3);
    __syncthreads();
    float4 acc = make_float4(0, 0, 0, 0);
    int idx = (threadIdx.x >> 5) * stride & 255;          // warp-uniform address: one broadcast wavefront
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float4 v;
            unsigned addr = (unsigned)__cvta_generic_to_shared(&sm[(idx + i) & 255]);
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        idx = (idx + stride) & 255;
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) out[0] = acc.x;
}

__global__ void k_shfl(float* out) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = __shfl_xor_sync(0xffffffffu, x[i], 1 + (i & 3));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}

// integer ALU + FMA mix: the issue-slot bound (one warp instruction per scheduler per clock)
__global__ void k_issue(float* out, float a, int b) {
    float x[4];
    int y[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { x[i] = threadIdx.x + i; y[i] = threadIdx.x * 7 + i; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { x[i] = fmaf(x[i], a, 1.0f); y[i] = (y[i] ^ b) + it; }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s += x[i] + y[i];
    if (s == 123.456f) out[0] = s;
}

template <typename F>
static double time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) launch();
    cudaDeviceSynchronize();
    double best = 1e30;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { printf("{\"error\": \"no CUDA device\"}\n"); return 1; }
    const int sms = prop.multiProcessorCount;
    const int blocks = sms * 8;
    float* out;
    cudaMalloc(&out, 64);
    const double threads = (double)blocks * THREADS, warps = threads / 32.0;
    const double ms_ffma = time_ms([&] { k_ffma<<<blocks, THREADS>>>(out, 1.0001f, 0.5f); });
    const double ms_ex2 = time_ms([&] { k_ex2<<<blocks, THREADS>>>(out, 0.37f); });
    const double ms_lds = time_ms([&] { k_lds<<<blocks, THREADS>>>(out, 3); });
    const double ms_shfl = time_ms([&] { k_shfl<<<blocks, THREADS>>>(out); });
    const double ms_issue = time_ms([&] { k_issue<<<blocks, THREADS>>>(out, 1.0001f, 12345); });
    const double n = (double)ITERS * 8.0;
    const double ffma_per_s = threads * n / (ms_ffma * 1e-3);
    const double ex2_per_s = threads * n / (ms_ex2 * 1e-3);
    const double lds_winst_per_s = warps * n / (ms_lds * 1e-3);
    const double shfl_winst_per_s = warps * n / (ms_shfl * 1e-3);
    const double issue_winst_per_s = warps * ((double)ITERS * 12.0) / (ms_issue * 1e-3);   // 4 FFMA + 4 LOP3 + 4 IADD per iteration
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"fp32_tflops\": %.3f, \"ffma_implied_sm_mhz\": %.1f, "
           "\"mufu_ex2_gops\": %.1f, \"mufu_ex2_per_clk_per_sm\": %.2f, "
           "\"lds128_gwinst_per_s\": %.2f, \"lds128_broadcast_gbs\": %.1f, \"shfl_gwinst_per_s\": %.2f, "
           "\"issue_ginst_per_s\": %.2f, \"issue_winst_per_clk_per_sm\": %.2f, "
           "\"how\": \"tools/sm_peaks.cu: %d CTAs x %d threads, %d iterations x 8 independent chains, best of 5, CUDA events\"}\n",
           prop.name, sms, 2.0 * ffma_per_s / 1e12, ffma_per_s / (sms * 128.0) / 1e6,
           ex2_per_s / 1e9, ex2_per_s / (ffma_per_s / (sms * 128.0)) / sms,
           lds_winst_per_s / 1e9, lds_winst_per_s * 512.0 / 1e9, shfl_winst_per_s / 1e9,
           issue_winst_per_s / 1e9, issue_winst_per_s / (ffma_per_s / (sms * 128.0)) / sms,
           blocks, THREADS, ITERS);
    cudaFree(out);
    return 0;
}
