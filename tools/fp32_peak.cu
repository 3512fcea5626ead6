// FP32 FFMA peak of this GPU (SURVEY.md section 6 "first GPU task"): every thread runs 16 independent
// FFMA chains; prints TFLOP/s (FMA = 2 flop) for a few occupancies.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu && tools/fp32_peak
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void __launch_bounds__(256) ffma_kernel(float *out, int iters, float b, float c) {
    float a[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) a[k] = threadIdx.x * 1e-6f + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) a[k] = fmaf(a[k], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    int dev = 0, sms = 0, mhz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, dev);
    float *out;
    cudaMalloc(&out, sizeof(float) * sms * 16 * 256);
    const int iters = 1 << 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    printf("{\"sms\": %d, \"clock_khz\": %d, \"theoretical_tflops_at_clock\": %.2f, \"runs\": [", sms, mhz,
           sms * 128.0 * 2.0 * mhz * 1e3 / 1e12);
    bool first = true;
    for (int per_sm = 2; per_sm <= 8; per_sm *= 2) {
        const int blocks = sms * per_sm;
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0);
            ffma_kernel<16><<<blocks, 256>>>(out, iters, 1.000001f, 1e-7f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double flop = 2.0 * 16 * (double)iters * blocks * 256;
        printf("%s{\"ctas_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}", first ? "" : ", ", per_sm, best,
               flop / (best * 1e-3) / 1e12);
        first = false;
    }
    printf("]}\n");
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
