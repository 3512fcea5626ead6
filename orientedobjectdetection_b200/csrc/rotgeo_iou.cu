// Pairwise rotated IoU / IoF matrix on packed boxes.  sm_100a.
//
// One CTA (256 threads) owns a 64-row x 128-column tile of the output:
//   0. the tile's row and column boxes (two float4 planes each, 6 KB) are staged
//      in shared memory by four TMA bulk copies signalled on one mbarrier;
//   1. cull pass: lane l of warp w tests rows {w, w+8, ...} against its four
//      columns 4l..4l+3 with the circumscribed-circle test (7 FP32 ops a pair),
//      stores the row's zeros as one coalesced float4, and remembers survivors
//      in a 32-bit mask (8 rows x 4 columns);
//   2. survivors are compacted into a shared-memory queue (one warp scan and one
//      shared atomic per warp);
//   3. exact pass: the whole CTA walks the queue with every lane busy on a pair
//      that really needs the clipper, and scatters the values over the zeros.
// Disjoint pairs therefore cost a handful of instructions and 4 B of HBM
// write, overlapping pairs cost the ~170-instruction clipper at full lane
// occupancy regardless of how sparse they are in the matrix.
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
    mbar_wait(&sm.bar, 0);

    // ---- 1. cull pass -------------------------------------------------------
    float cxs[4], cys[4], crs[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 c = sm.col0[4 * lane + k];
        cxs[k] = c.x; cys[k] = c.y; crs[k] = c.z;
    }
    const int64_t gcol = col_base + 4 * lane;
    const int cols_here = (int)min((int64_t)4, n2 - gcol);  // <= 0: lane is past the edge
    uint32_t hits = 0;
#pragma unroll
    for (int it = 0; it < kRowsPerWarp; ++it) {
        const int r = warp + kWarps * it;
        const int64_t grow = row_base + r;
        const float4 R = sm.row0[r];
        const bool row_ok = grow < n1;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool hit = row_ok && (k < cols_here) &&
                             !circles_disjoint(R.x, R.y, R.z, cxs[k], cys[k], crs[k]);
            hits |= (hit ? 1u : 0u) << (4 * it + k);
        }
        if (row_ok) {
            float *dst = out + grow * ld_out + gcol;
            if (VEC) {
                if (cols_here > 0) *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (k < cols_here) dst[k] = 0.f;
            }
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
    __syncthreads();

    // ---- 3. exact pass over the queue ----------------------------------------
    const int total = sm.qcount;
    for (int e = tid; e < total; e += kThreads) {
        const int q = sm.queue[e];
        const int r = q >> 7, c = q & 127;
        const float v = rect_overlap<MODE>(sm.row0[r], sm.row1[r], sm.col0[c], sm.col1[c]);
        out[(row_base + r) * ld_out + (col_base + c)] = v;
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
    p0 = make_float4(b[0], b[1], 0.f, w * h);
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
