// Pairwise rotated IoU / IoF matrix on packed boxes.  sm_100a.
//
// One CTA (256 threads) owns a 64-row x 128-column tile of the output:
//   0. the tile's row and column boxes (two float4 planes each, 6 KB) are staged
//      in shared memory by four TMA bulk copies signalled on one mbarrier; while
//      they are in flight the CTA streams the tile's zeros to HBM (coalesced
//      float4 stores, 512 B per warp instruction) — most of the matrix is zero;
//   1. cull level 1: lane l of warp w tests rows {w, w+8, ...} against its four
//      columns 4l..4l+3 with the modular bin masks (two ANDs a pair) and
//      remembers survivors in a 32-bit mask (8 rows x 4 columns);
//   2. survivors are compacted into a shared-memory queue (one warp scan and one
//      shared atomic per warp);
//   3. the whole CTA walks the queue: circumscribed-circle test (cull level 2),
//      then the exact clipper, and scatters the non-zero values over the zeros.
// Disjoint pairs therefore cost ~4 instructions and 4 B of HBM write; the
// kernel is bound by the HBM write stream (DESIGN.md "Kernels").
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTileRows = 64;
constexpr int kTileCols = 128;
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kRowsPerWarp = kTileRows / kWarps;  // 8

struct __align__(16) IouSmem {
    float4 row0[kTileRows];
    float4 row1[kTileRows];
    float4 col0[kTileCols];
    float4 col1[kTileCols];
    uint16_t queue[kTileRows * kTileCols];
    uint64_t bar;
    int qcount;
};

template <int MODE, bool VEC>
__global__ void __launch_bounds__(kThreads)
iou_tile_kernel(const float4 *__restrict__ a0, const float4 *__restrict__ a1, int64_t n1,
                const float4 *__restrict__ b0, const float4 *__restrict__ b1, int64_t n2,
                float *__restrict__ out, int64_t ld_out, unsigned tiles_x) {
    __shared__ IouSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // 1-D grid, column tiles fastest: neighbouring CTAs share their row boxes
    const int64_t row_base = (int64_t)(blockIdx.x / tiles_x) * kTileRows;
    const int64_t col_base = (int64_t)(blockIdx.x % tiles_x) * kTileCols;

    if (tid == 0) {
        sm.qcount = 0;
        mbar_init(&sm.bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(&sm.bar, (2 * kTileRows + 2 * kTileCols) * sizeof(float4));
        tma_load_1d(sm.row0, a0 + row_base, kTileRows * sizeof(float4), &sm.bar);
        tma_load_1d(sm.row1, a1 + row_base, kTileRows * sizeof(float4), &sm.bar);
        tma_load_1d(sm.col0, b0 + col_base, kTileCols * sizeof(float4), &sm.bar);
        tma_load_1d(sm.col1, b1 + col_base, kTileCols * sizeof(float4), &sm.bar);
    }

    // ---- 0. the tile's zeros, while the TMA copies are in flight ---------------
    {
        const int64_t gcol = col_base + 4 * lane;
        const int cols_here = (int)min((int64_t)4, n2 - gcol);  // <= 0: lane is past the edge
        const int rows_here = (int)min((int64_t)kTileRows, n1 - row_base);
        float *dst = out + (row_base + warp) * ld_out + gcol;
        const int64_t step = (int64_t)kWarps * ld_out;
        if (VEC) {
            if (rows_here == kTileRows && cols_here == 4) {   // interior tile: 8 plain stores
#pragma unroll
                for (int it = 0; it < kRowsPerWarp; ++it) {
                    *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
                    dst += step;
                }
            } else if (cols_here > 0) {
                for (int r = warp; r < rows_here; r += kWarps) {
                    *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
                    dst += step;
                }
            }
        } else {
            for (int it = 0; it < kRowsPerWarp; ++it)
                if (warp + kWarps * it < rows_here)
                    for (int k = 0; k < cols_here; ++k) dst[it * step + k] = 0.f;
        }
    }
    mbar_wait(&sm.bar, 0);

    // ---- 1. cull level 1: bin masks (padding rows / columns carry empty masks) --
    uint32_t cxm[4], cym[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 c = sm.col0[4 * lane + k];
        cxm[k] = __float_as_uint(c.z);
        cym[k] = __float_as_uint(c.w);
    }
    uint32_t hits = 0;
#pragma unroll
    for (int it = 0; it < kRowsPerWarp; ++it) {
        const float2 rm = *reinterpret_cast<const float2 *>(&sm.row0[warp + kWarps * it].z);
        const uint32_t rxm = __float_as_uint(rm.x), rym = __float_as_uint(rm.y);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool hit = ((rxm & cxm[k]) != 0u) & ((rym & cym[k]) != 0u);
            hits |= (hit ? 1u : 0u) << (4 * it + k);
        }
    }

    // ---- 2. compact survivors into the queue ---------------------------------
    if (__any_sync(0xffffffffu, hits != 0)) {
        const int cnt = __popc(hits);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        int base = 0;
        if (lane == 31) base = atomicAdd(&sm.qcount, incl);
        base = __shfl_sync(0xffffffffu, base, 31);
        int pos = base + incl - cnt;
        uint32_t m = hits;
        while (m) {
            const int bit = __ffs(m) - 1;
            m &= m - 1;
            const int r = warp + kWarps * (bit >> 2);
            const int c = 4 * lane + (bit & 3);
            sm.queue[pos++] = (uint16_t)((r << 7) | c);
        }
    }
    __syncthreads();   // also orders the zero stores before the value stores below

    // ---- 3. cull level 2 + exact pass over the queue ---------------------------
    const int total = sm.qcount;
    for (int e = tid; e < total; e += kThreads) {
        const int q = sm.queue[e];
        const int r = q >> 7, c = q & 127;
        const float v = pair_value<MODE>(sm.row0[r], sm.row1[r], sm.col0[c], sm.col1[c]);
        if (v != 0.f) out[(row_base + r) * ld_out + (col_base + c)] = v;
    }
}

int launch_iou_pairwise(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                        float *out, int64_t ld_out, cudaStream_t stream) {
    if (n1 == 0 || n2 == 0) return 0;
    const PackedView A = packed_view(packed1, padded_count(n1));
    const PackedView B = packed_view(packed2, padded_count(n2));
    const int64_t tiles_x = (n2 + kTileCols - 1) / kTileCols;
    const int64_t tiles = tiles_x * ((n1 + kTileRows - 1) / kTileRows);
    if (tiles > 0x7fffffffLL) return -1;
    const unsigned grid = (unsigned)tiles;
    const bool vec = (ld_out % 4 == 0) && (n2 % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
#define ROTGEO_LAUNCH(M, V)                                                                     \
    iou_tile_kernel<M, V><<<grid, kThreads, 0, stream>>>(A.p0, A.p1, n1, B.p0, B.p1, n2, out, ld_out, \
                                                         (unsigned)tiles_x)
    if (mode == 0) {
        if (vec) ROTGEO_LAUNCH(0, true); else ROTGEO_LAUNCH(0, false);
    } else {
        if (vec) ROTGEO_LAUNCH(1, true); else ROTGEO_LAUNCH(1, false);
    }
#undef ROTGEO_LAUNCH
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

// ---- aligned variant: out[i] = overlap(boxes1[i], boxes2[i]) -------------------
__device__ __forceinline__ void load_box(const float *b, float min_wh, float4 &p0, float4 &p1) {
    float w = b[2], h = b[3];
    if (min_wh > 0.f) {
        w = fmaxf(w, min_wh);
        h = fmaxf(h, min_wh);
    }
    double sd, cd;
    sincos((double)b[4], &sd, &cd);
    p0 = make_float4(b[0], b[1], 0.f, 0.f);
    p1 = make_float4((float)cd, (float)sd, 0.5f * w, 0.5f * h);
}

template <int MODE>
__global__ void __launch_bounds__(128)
iou_aligned_kernel(const float *__restrict__ b1, const float *__restrict__ b2, int64_t n, int s1,
                   int s2, float min_wh, float *__restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 A0, A1, B0, B1;
    load_box(b1 + i * s1, min_wh, A0, A1);
    load_box(b2 + i * s2, min_wh, B0, B1);
    out[i] = rect_overlap<MODE>(A0, A1, B0, B1);
}

int launch_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1,
                       int stride2, int mode, float min_wh, float *out, cudaStream_t stream) {
    if (n == 0) return 0;
    const int threads = 128;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (mode == 0)
        iou_aligned_kernel<0><<<blocks, threads, 0, stream>>>(boxes1, boxes2, n, stride1, stride2, min_wh, out);
    else
        iou_aligned_kernel<1><<<blocks, threads, 0, stream>>>(boxes1, boxes2, n, stride1, stride2, min_wh, out);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
