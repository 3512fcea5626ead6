// Measures what a WRITE-ONLY stream reaches on this GPU (context for the IoU
// kernel's roofline: MEASURED_PEAKS.json's hbm_gbs is a read+write copy).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/hbm_write_probe tools/hbm_write_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void fill_gridstride(float4 *p, size_t n4) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n4; i += step) p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// the IoU kernel's pattern: CTA = 64 rows x 128 cols tile, warp stores 512 B rows
__global__ void fill_tiles(float *out, int64_t ld, unsigned tiles_x) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row_base = (int64_t)(blockIdx.x / tiles_x) * 64;
    const int64_t col_base = (int64_t)(blockIdx.x % tiles_x) * 128;
    float *dst = out + (row_base + warp) * ld + col_base + 4 * lane;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        dst += 8 * ld;
    }
}

// same tiles, but a whole tile row-block of 8 consecutive rows per warp
__global__ void fill_tiles_rowmajor(float *out, int64_t ld, unsigned tiles_x) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row_base = (int64_t)(blockIdx.x / tiles_x) * 64;
    const int64_t col_base = (int64_t)(blockIdx.x % tiles_x) * 128;
    float *dst = out + (row_base + 8 * warp) * ld + col_base + 4 * lane;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        dst += ld;
    }
}

// CTA owns 8 full rows x 2048 columns (contiguous 8 KB runs)
__global__ void fill_wide(float *out, int64_t ld, unsigned chunks_x) {
    const int64_t row = (int64_t)(blockIdx.x / chunks_x) * 8 + (threadIdx.x >> 5);
    const int64_t col = (int64_t)(blockIdx.x % chunks_x) * 2048 + 4 * (threadIdx.x & 31);
    float *dst = out + row * ld + col;
#pragma unroll
    for (int it = 0; it < 16; ++it) *reinterpret_cast<float4 *>(dst + it * 128) = make_float4(0.f, 0.f, 0.f, 0.f);
}

// persistent variant: CTA c fills tiles c, c+G, ...
__global__ void fill_tiles_persistent(float *out, int64_t ld, unsigned tiles_x, unsigned tiles) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (unsigned t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int64_t row_base = (int64_t)(t / tiles_x) * 64;
        const int64_t col_base = (int64_t)(t % tiles_x) * 128;
        float *dst = out + (row_base + warp) * ld + col_base + 4 * lane;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
            dst += 8 * ld;
        }
    }
}
// persistent, contiguous chunk of tiles per CTA
__global__ void fill_tiles_persistent_chunk(float *out, int64_t ld, unsigned tiles_x, unsigned tiles) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned per = (tiles + gridDim.x - 1) / gridDim.x;
    const unsigned t0 = blockIdx.x * per, t1 = min(tiles, t0 + per);
    for (unsigned t = t0; t < t1; ++t) {
        const int64_t row_base = (int64_t)(t / tiles_x) * 64;
        const int64_t col_base = (int64_t)(t % tiles_x) * 128;
        float *dst = out + (row_base + warp) * ld + col_base + 4 * lane;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
            dst += 8 * ld;
        }
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// TMA bulk stores: one thread per CTA pushes 64 rows x 512 B from a zeroed smem row
__global__ void fill_tiles_tma(float *out, int64_t ld, unsigned tiles_x) {
    __shared__ __align__(128) float zrow[128];
    if (threadIdx.x < 128) zrow[threadIdx.x] = 0.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int64_t row_base = (int64_t)(blockIdx.x / tiles_x) * 64;
    const int64_t col_base = (int64_t)(blockIdx.x % tiles_x) * 128;
    if (threadIdx.x < 64) {
        float *dst = out + (row_base + threadIdx.x) * ld + col_base;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 512;" ::"l"(dst), "r"(smem_u32(zrow)) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

template <typename F>
float time_it(F fn, void *flush, size_t flush_bytes, int reps, float *best) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<float> ts;
    for (int i = 0; i < reps + 3; ++i) {
        cudaMemsetAsync(flush, 1, flush_bytes, 0);
        cudaEventRecord(e0, 0);
        fn();
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 3) ts.push_back(ms);
    }
    std::sort(ts.begin(), ts.end());
    *best = ts[0];
    return ts[ts.size() / 2];
}

int main() {
    const int64_t N = 8192, M = 8192;
    const size_t bytes = (size_t)N * M * 4, flush_bytes = 512u << 20;
    float *out; void *flush;
    CK(cudaMalloc(&out, bytes));
    CK(cudaMalloc(&flush, flush_bytes));
    float best, med;
#define REPORT(name) printf("%-34s median %.1f us (%.0f GB/s)  best %.1f us (%.0f GB/s)\n", name, med * 1e3, bytes / med / 1e6, best * 1e3, bytes / best / 1e6)
    med = time_it([&] { cudaMemsetAsync(out, 0, bytes, 0); }, flush, flush_bytes, 20, &best); REPORT("cudaMemsetAsync");
    for (int ctas : {148 * 4, 148 * 8, 148 * 16, 148 * 32}) {
        med = time_it([&] { fill_gridstride<<<ctas, 256>>>((float4 *)out, bytes / 16); }, flush, flush_bytes, 20, &best);
        char nm[64]; snprintf(nm, 64, "gridstride float4 %d CTAs", ctas); REPORT(nm);
    }
    med = time_it([&] { fill_gridstride<<<(unsigned)(bytes / 16 / 256), 256>>>((float4 *)out, bytes / 16); }, flush, flush_bytes, 20, &best); REPORT("one float4 per thread");
    med = time_it([&] { fill_tiles<<<8192, 256>>>(out, M, 64); }, flush, flush_bytes, 20, &best); REPORT("64x128 tiles, rows w+8k");
    med = time_it([&] { fill_tiles_rowmajor<<<8192, 256>>>(out, M, 64); }, flush, flush_bytes, 20, &best); REPORT("64x128 tiles, rows 8w+k");
    med = time_it([&] { fill_wide<<<8192 / 8 * 4, 256>>>(out, M, 4); }, flush, flush_bytes, 20, &best); REPORT("8 rows x 2048 cols per CTA");
    med = time_it([&] { fill_tiles_tma<<<8192, 128>>>(out, M, 64); }, flush, flush_bytes, 20, &best); REPORT("64x128 tiles, TMA bulk stores");
    for (int k : {1, 2, 4, 5, 8}) {
        med = time_it([&] { fill_tiles_persistent<<<148 * k, 256>>>(out, M, 64, 8192); }, flush, flush_bytes, 20, &best);
        char nm[64]; snprintf(nm, 64, "persistent strided, %d CTA/SM", k); REPORT(nm);
        med = time_it([&] { fill_tiles_persistent_chunk<<<148 * k, 256>>>(out, M, 64, 8192); }, flush, flush_bytes, 20, &best);
        snprintf(nm, 64, "persistent chunked, %d CTA/SM", k); REPORT(nm);
    }
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    printf("done\n");
    return 0;
}
