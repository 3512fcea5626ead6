// Segmented greedy rotated NMS, entirely on the device.  sm_100a.
//
// Input: boxes packed in SEGMENT-MAJOR order (segment = class / FPN level /
// image x level ...), score-descending inside each segment.  Three kernels:
//
//   nms_mask_kernel  one CTA per 128x128 tile of the UPPER TRIANGLE of each
//                    segment's pair matrix (block-diagonal: pairs from different
//                    segments are never touched).  Boxes are staged by TMA bulk
//                    copies; cull level 1 (modular bin masks, two ANDs a pair)
//                    compacts surviving pairs into a shared queue, the queue pass
//                    applies the circle test and the exact clipper and ORs
//                    `IoU >= thr` bits into a shared 128x2-word tile that is
//                    written out once.
//   nms_scan_kernel  one CTA per segment walks its 64-box chunks in score order.
//                    Inside a chunk the kept set is found by peeling (warp
//                    redux.or rounds, usually 1-3) instead of a 64-step serial
//                    loop; kept rows are then OR-ed into the `removed` bitmap
//                    (shared memory) by the whole CTA.  The mask never leaves the GPU.
//   nms_emit_kernel  (several segments only) compacts the kept flags in GLOBAL
//                    score order into keep_out — the order mmcv.ops.nms_rotated
//                    returns and every reference caller slices with [:max_num].
//
// Replaces mmcv's nms_rotated (64x64 bitmask over the full square + D2H copy of
// the mask + serial host scan; reference call sites
// mmrotate/core/post_processing/bbox_nms_rotated.py:80,122,
// mmrotate/core/patch/merge_results.py:116-118, mmrotate/datasets/dota.py:379).
#include <math.h>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTile = 128;      // tile edge in boxes (two 64-bit words per row)
constexpr int kHalf = 64;       // rows per cull/exact pass
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

struct __align__(16) MaskSmem {
    float4 row0[kTile];
    float4 row1[kTile];
    float4 col0[kTile];
    float4 col1[kTile];
    uint2 rowm[kTile];
    uint2 colm[kTile];
    unsigned long long bits[kTile][2];
    uint16_t queue[kHalf * kTile];
    uint64_t bar;
    int qcount;
};

__device__ __forceinline__ int find_segment(const NmsSegment *__restrict__ segs, int nseg,
                                            int64_t tile) {
    int lo = 0, hi = nseg - 1;  // last segment with tile_base <= tile
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (segs[mid].tile_base <= tile) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(kThreads)
nms_mask_kernel(const float4 *__restrict__ p0, const float4 *__restrict__ p1,
                const uint2 *__restrict__ pm, const NmsSegment *__restrict__ segs, int nseg, float thr,
                unsigned long long *__restrict__ mask) {
    __shared__ MaskSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const int si = find_segment(segs, nseg, blockIdx.x);
    const NmsSegment seg = segs[si];
    const int64_t t = (int64_t)blockIdx.x - seg.tile_base;
    const int C = seg.words >> 1;  // 128-box column blocks
    // tiles before block-row i: i*C - i*(i-1)/2
    int I = (int)(((2.0 * C + 1.0) - sqrt((2.0 * C + 1.0) * (2.0 * C + 1.0) - 8.0 * (double)t)) * 0.5);
    I = max(0, min(I, C - 1));
    while (I + 1 < C && (int64_t)(I + 1) * C - (int64_t)(I + 1) * I / 2 <= t) ++I;
    while (I > 0 && (int64_t)I * C - (int64_t)I * (I - 1) / 2 > t) --I;
    const int J = I + (int)(t - ((int64_t)I * C - (int64_t)I * (I - 1) / 2));
    const int row_base = I * kTile, col_base = J * kTile;  // positions inside the segment
    const int L = (int)seg.len;

    if (tid == 0) {
        mbar_init(&sm.bar, 1);
        fence_mbar_init();
    }
    sm.bits[tid >> 1][tid & 1] = 0ull;
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(&sm.bar, 4 * kTile * sizeof(float4));
        tma_load_1d(sm.row0, p0 + seg.start + row_base, kTile * sizeof(float4), &sm.bar);
        tma_load_1d(sm.row1, p1 + seg.start + row_base, kTile * sizeof(float4), &sm.bar);
        tma_load_1d(sm.col0, p0 + seg.start + col_base, kTile * sizeof(float4), &sm.bar);
        tma_load_1d(sm.col1, p1 + seg.start + col_base, kTile * sizeof(float4), &sm.bar);
    }
    // bin masks: plain coalesced loads (a segment may start at an odd box, and a
    // bulk copy needs 16-byte alignment; the float4 geometry planes always have it)
    if (tid < kTile) sm.rowm[tid] = pm[seg.start + row_base + tid];
    else sm.colm[tid - kTile] = pm[seg.start + col_base + tid - kTile];
    __syncthreads();
    mbar_wait(&sm.bar, 0);

    // cull level 1 operands: bin masks of this lane's four columns; positions past
    // the segment's end belong to the next segment (or are padding) -> empty masks
    uint32_t cxm[4], cym[4];
    const int cpos = col_base + 4 * lane;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint2 c = sm.colm[4 * lane + k];
        const bool live = cpos + k < L;
        cxm[k] = live ? c.x : 0u;
        cym[k] = live ? c.y : 0u;
    }
    const bool diagonal = I == J;   // only there can a column precede its row

    for (int half = 0; half < 2; ++half) {
        const int hrow = half * kHalf;
        if (row_base + hrow >= L) break;                 // uniform
        if (col_base + kTile - 1 <= row_base + hrow) continue;  // tile half entirely below the diagonal
        if (tid == 0) sm.qcount = 0;
        __syncthreads();
        uint32_t hits = 0;
#pragma unroll
        for (int it = 0; it < kHalf / kWarps; ++it) {
            const int r = hrow + warp + kWarps * it;
            const int rpos = row_base + r;
            const uint2 rm = sm.rowm[r];
            const uint32_t rxm = rpos < L ? rm.x : 0u;
            const uint32_t rym = rm.y;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                bool hit = ((rxm & cxm[k]) != 0u) & ((rym & cym[k]) != 0u);
                if (diagonal) hit = hit && (cpos + k > rpos);
                hits |= (hit ? 1u : 0u) << (4 * it + k);
            }
        }
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
                const int r = warp + kWarps * (bit >> 2);   // row inside this half
                const int c = 4 * lane + (bit & 3);
                sm.queue[pos++] = (uint16_t)((r << 7) | c);
            }
        }
        __syncthreads();
        const int total = sm.qcount;
        for (int e = tid; e < total; e += kThreads) {
            const int q = sm.queue[e];
            const int r = hrow + (q >> 7), c = q & 127;
            const float v = pair_value<0>(sm.row0[r], sm.row1[r], sm.col0[c], sm.col1[c]);
            if (v >= thr) atomicOr(&sm.bits[r][c >> 6], 1ull << (c & 63));
        }
        __syncthreads();
    }
    __syncthreads();
    {
        const int r = tid >> 1, w = tid & 1;
        const int rpos = row_base + r;
        if (rpos < L)
            mask[seg.mask_base + (int64_t)rpos * seg.words + 2 * J + w] = sm.bits[r][w];
    }
}

// ---- greedy scan ---------------------------------------------------------------
constexpr int kScanThreads = 512;

__device__ __forceinline__ unsigned long long warp_or64(unsigned long long v) {
    const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
    const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

template <bool EMIT>
__global__ void __launch_bounds__(kScanThreads)
nms_scan_kernel(const NmsSegment *__restrict__ segs, const unsigned long long *__restrict__ mask,
                const int64_t *__restrict__ seg_order, uint8_t *__restrict__ flags,
                int64_t *__restrict__ keep_out, int64_t *__restrict__ num_keep_out) {
    extern __shared__ unsigned long long removed[];  // seg.words entries
    __shared__ unsigned long long s_kept;
    __shared__ int s_rows[64];
    __shared__ int s_nk;

    const NmsSegment seg = segs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = seg.words;
    const int L = (int)seg.len;
    const unsigned long long *__restrict__ M = mask + seg.mask_base;

    for (int w = tid; w < W; w += kScanThreads) {
        // positions >= L do not exist: mark them removed
        const int first = w * 64;
        unsigned long long v = 0ull;
        if (first >= L) v = ~0ull;
        else if (first + 64 > L) v = ~0ull << (L - first);
        removed[w] = v;
    }
    __syncthreads();

    int64_t nkeep = 0;
    const int chunks = seg.row_blocks;
    for (int c = 0; c < chunks; ++c) {
        if (warp == 0) {
            // lane holds diagonal words of rows c*64+lane and c*64+32+lane
            const int r0 = c * 64 + lane, r1 = r0 + 32;
            const unsigned long long d0 = r0 < L ? M[(int64_t)r0 * W + c] : 0ull;
            const unsigned long long d1 = r1 < L ? M[(int64_t)r1 * W + c] : 0ull;
            unsigned long long cand = ~removed[c];
            unsigned long long kept = 0ull;
            while (cand) {
                const bool in0 = (cand >> lane) & 1ull, in1 = (cand >> (lane + 32)) & 1ull;
                // everything some undecided box could still suppress
                const unsigned long long threat = warp_or64((in0 ? d0 : 0ull) | (in1 ? d1 : 0ull));
                const unsigned long long safe = cand & ~threat;   // nobody left can suppress these
                kept |= safe;
                const bool s0 = (safe >> lane) & 1ull, s1 = (safe >> (lane + 32)) & 1ull;
                const unsigned long long killed = warp_or64((s0 ? d0 : 0ull) | (s1 ? d1 : 0ull));
                cand &= ~(safe | killed);
            }
            if (lane == 0) {
                s_kept = kept;
                s_nk = __popcll(kept);
            }
            const bool k0 = (kept >> lane) & 1ull, k1 = (kept >> (lane + 32)) & 1ull;
            const int rk0 = __popcll(kept & ((1ull << lane) - 1ull));
            const int rk1 = __popcll(kept & ((1ull << (lane + 32)) - 1ull));
            if (k0) s_rows[rk0] = r0;
            if (k1) s_rows[rk1] = r1;
            if (EMIT) {
                if (k0) keep_out[nkeep + rk0] = seg_order[seg.start + r0];
                if (k1) keep_out[nkeep + rk1] = seg_order[seg.start + r1];
            } else {
                if (r0 < L) flags[seg_order[seg.start + r0]] = k0 ? 1 : 0;
                if (r1 < L) flags[seg_order[seg.start + r1]] = k1 ? 1 : 0;
            }
        }
        __syncthreads();
        const int nk = s_nk;
        nkeep += nk;
        const int wrem = W - (c + 1);
        if (wrem > 0 && nk > 0) {
            const int items = nk * wrem;
            for (int it = tid; it < items; it += kScanThreads) {
                const int k = it / wrem;
                const int w = c + 1 + (it - k * wrem);
                const unsigned long long v = M[(int64_t)s_rows[k] * W + w];
                if (v) atomicOr(&removed[w], v);
            }
        }
        __syncthreads();
    }
    if (EMIT && tid == 0) *num_keep_out = nkeep;
}

// ---- ordered compaction of the kept flags (several segments) ---------------------
__global__ void __launch_bounds__(1024)
nms_emit_kernel(const int64_t *__restrict__ order, const uint8_t *__restrict__ flags, int64_t n,
                int64_t *__restrict__ keep_out, int64_t *__restrict__ num_keep_out) {
    __shared__ int warp_cnt[32];
    __shared__ int64_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int64_t start = 0; start < n; start += 1024) {
        const int64_t i = start + tid;
        int64_t idx = -1;
        bool f = false;
        if (i < n) {
            idx = order[i];
            f = flags[idx] != 0;
        }
        const unsigned b = __ballot_sync(0xffffffffu, f);
        if (lane == 0) warp_cnt[warp] = __popc(b);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 32; ++w) {
            const int cnt = warp_cnt[w];
            before += (w < warp) ? cnt : 0;
            total += cnt;
        }
        const int64_t base = s_base;
        if (f) keep_out[base + before + __popc(b & ((1u << lane) - 1u))] = idx;
        __syncthreads();
        if (tid == 0) s_base = base + total;
        __syncthreads();
    }
    if (tid == 0) *num_keep_out = s_base;
}

// ---- host-side plan + launch -------------------------------------------------------
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

void nms_plan(int64_t n, const int64_t *seg_offsets_host, int64_t num_segments, NmsPlan *plan,
              NmsSegment *segments_out) {
    plan->n = n;
    plan->n_pad = padded_count(n) + kTile;  // slack: a tile may start at any box
    int64_t one[2] = {0, n};
    if (seg_offsets_host == nullptr || num_segments <= 0) {
        seg_offsets_host = one;
        num_segments = 1;
    }
    plan->num_segments = num_segments;
    int64_t tiles = 0, words = 0;
    for (int64_t s = 0; s < num_segments; ++s) {
        const int64_t start = seg_offsets_host[s], len = seg_offsets_host[s + 1] - start;
        const int64_t C = (len + kTile - 1) / kTile;
        if (segments_out) {
            NmsSegment &g = segments_out[s];
            g.start = start;
            g.len = len;
            g.mask_base = words;
            g.tile_base = tiles;
            g.words = (int32_t)(2 * C);
            g.row_blocks = (int32_t)((len + 63) / 64);
        }
        tiles += C * (C + 1) / 2;
        words += len * 2 * C;
    }
    plan->total_tiles = tiles;
    plan->mask_words = words;
    size_t off = 0;
    plan->off_packed = off;   off = align_up(off + (size_t)plan->n_pad * kPackedBytesPerBox, 256);
    plan->off_segments = off; off = align_up(off + (size_t)num_segments * sizeof(NmsSegment), 256);
    plan->off_mask = off;     off = align_up(off + (size_t)words * sizeof(uint64_t), 256);
    plan->off_flags = off;    off = align_up(off + (size_t)n, 256);
    plan->off_seg_keep = off;
    plan->total_bytes = off;
}

int launch_nms(const NmsPlan &plan, const void *packed, const NmsSegment *segments_dev,
               int max_words, const int64_t *order, const int64_t *seg_order, float thr,
               uint64_t *mask, uint8_t *flags, int64_t *keep_out, int64_t *num_keep_out,
               cudaStream_t stream) {
    const PackedView P = packed_view(packed, plan.n_pad);
    if (plan.total_tiles > 0x7fffffffLL) return -1;
    if (plan.total_tiles > 0)
        nms_mask_kernel<<<(unsigned)plan.total_tiles, kThreads, 0, stream>>>(
            P.p0, P.p1, P.pm, segments_dev, (int)plan.num_segments, thr,
            reinterpret_cast<unsigned long long *>(mask));
    if (cudaGetLastError() != cudaSuccess) return -3;
    const size_t dyn = (size_t)max_words * sizeof(unsigned long long);
    const bool single = plan.num_segments == 1;
    if (single) {
        if (dyn > 48 * 1024 &&
            cudaFuncSetAttribute(nms_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)dyn) != cudaSuccess)
            return -3;
        nms_scan_kernel<true><<<1, kScanThreads, dyn, stream>>>(
            segments_dev, reinterpret_cast<const unsigned long long *>(mask), seg_order, flags,
            keep_out, num_keep_out);
    } else {
        if (dyn > 48 * 1024 &&
            cudaFuncSetAttribute(nms_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)dyn) != cudaSuccess)
            return -3;
        nms_scan_kernel<false><<<(unsigned)plan.num_segments, kScanThreads, dyn, stream>>>(
            segments_dev, reinterpret_cast<const unsigned long long *>(mask), seg_order, flags,
            keep_out, num_keep_out);
        if (cudaGetLastError() != cudaSuccess) return -3;
        nms_emit_kernel<<<1, 1024, 0, stream>>>(order, flags, plan.n, keep_out, num_keep_out);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
