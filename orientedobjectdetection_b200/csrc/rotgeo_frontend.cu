// Fused front-end of the class-batched rotated NMS.  sm_100a.
//
// Replaces the python body of the reference's multiclass_nms_rotated
// (mmrotate/core/post_processing/bbox_nms_rotated.py:33-92): expand boxes to
// (n*C, 5), threshold the scores, nonzero() (a host sync), three gathers, two
// global max reductions for the class offset, the sort inside mmcv.ops.nms_rotated
// and the final gathers/cat — a dozen small launches and several device->host
// round trips.  Here the whole call is a fixed sequence of launches and NO
// device->host read:
//
//   mc_keys_kernel   one thread per (row, class): candidate test, 32-bit key (image of
//                    the score, descending) + flat index, per-class counts,
//                    max(cx,cy) / max(w,h) of the candidates
//   cub radix sort   stable pairs sort on the 32 key bits (4 passes): global score order
//   mc_class_kernel  class of every entry of that order (1 byte); thread 0 also makes the
//                    segment table and tile count from the class counts (the plan the
//                    host makes in rotgeo_nms, made on the device)
//   cub radix sort   one stable 8-bit pass on the class: segment-major order, score
//                    order preserved inside a class
//   pack / mask / resolve kernels of rotgeo_nms, launched over host-side upper bounds;
//                    the resolve kernel's compaction runs in global score order and writes
//                    the (box, score) rows and labels directly
//
// Sizes that depend on the data (candidates, tiles, mask words) never reach the host:
// grids and workspace are sized for the worst case (every class keeps all n rows)
// and the kernels read the actual counts from the device-side header.
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_run_length_encode.cuh>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTile = 128;
constexpr uint32_t kInvalidKey = 0xffffffffu;   // behind every candidate (their keys are < 2^31)

struct McHeader {
    long long n_valid;        // candidates (score > score_thr)
    long long total_tiles;    // mask tiles over all segments
    unsigned int max_xy, max_wh;   // order-preserving integer images of the two maxima
    float mc_max[2];          // the same as floats (written by mc_plan_kernel)
    int counts[kMcMaxClasses];
};

// float -> unsigned whose integer order is the float order
__device__ __forceinline__ unsigned int ordered_bits(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_float(unsigned int o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__global__ void __launch_bounds__(256)
mc_keys_kernel(const float *__restrict__ boxes, int shared, const float *__restrict__ scores,
               const float *__restrict__ factors, int64_t rows, int classes, float score_thr,
               uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, McHeader *__restrict__ hdr) {
    __shared__ int s_cnt[kMcMaxClasses];
    __shared__ unsigned int s_max[2];
    const int tid = threadIdx.x;
    if (tid < kMcMaxClasses) s_cnt[tid] = 0;
    if (tid < 2) s_max[tid] = 0u;
    __syncthreads();
    const int64_t total = rows * classes;
    const int64_t f = blockIdx.x * (int64_t)blockDim.x + tid;
    if (f < total) {
        const int64_t row = f / classes;
        const int cls = (int)(f - row * classes);
        const float raw = scores[row * (classes + 1) + cls];
        uint32_t key = kInvalidKey;
        if (raw > score_thr) {                       // tested BEFORE the factors, like the reference
            float sc = raw;
            if (factors) sc *= factors[row];
            // descending score: complement of the order-preserving image (+0.0f: -0 -> +0);
            // equal scores keep their flat-index order because the sort is stable
            key = 0xffffffffu - ordered_bits(sc + 0.0f);
            if (key == kInvalidKey) key = kInvalidKey - 1u;   // -NaN payloads only
            atomicAdd(&s_cnt[cls], 1);
            const float *b = boxes + (shared ? row : f) * 5;
            atomicMax(&s_max[0], ordered_bits(fmaxf(b[0], b[1])));
            atomicMax(&s_max[1], ordered_bits(fmaxf(b[2], b[3])));
        }
        keys[f] = key;
        vals[f] = (uint32_t)f;
    }
    __syncthreads();
    if (tid < classes && s_cnt[tid]) atomicAdd(&hdr->counts[tid], s_cnt[tid]);
    if (tid == 0 && s_max[0]) atomicMax(&hdr->max_xy, s_max[0]);
    if (tid == 1 && s_max[1]) atomicMax(&hdr->max_wh, s_max[1]);
}

// class byte of every entry of the global score order (255 behind the candidates); thread 0
// of the grid also makes the plan nms_plan() makes on the host (<= 64 segments)
__global__ void __launch_bounds__(256)
mc_class_kernel(const uint32_t *__restrict__ keys_sorted, const uint32_t *__restrict__ vals_sorted,
                int64_t total, int classes, uint8_t *__restrict__ cls_out, McHeader *hdr,
                NmsSegment *segs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < total)
        cls_out[i] = keys_sorted[i] == kInvalidKey ? (uint8_t)255 : (uint8_t)(vals_sorted[i] % (uint32_t)classes);
    if (i != 0) return;
    long long start = 0, slots = 0, words = 0, tiles = 0;
    for (int c = 0; c < classes; ++c) {
        const long long len = hdr->counts[c];
        const long long C = (len + kTile - 1) / kTile;
        NmsSegment &g = segs[c];
        g.start = start;
        g.pstart = slots;
        g.len = len;
        g.mask_base = words;
        g.tile_base = tiles;
        g.words = (int32_t)(2 * C);
        g.row_blocks = (int32_t)((len + 63) / 64);
        start += len;
        slots += C * kTile;
        tiles += C * (C + 1) / 2;
        words += len * 2 * C;
    }
    hdr->n_valid = start;
    hdr->total_tiles = tiles;
    hdr->mc_max[0] = ordered_float(hdr->max_xy);
    hdr->mc_max[1] = ordered_float(hdr->max_wh);
}

// ---- host side ---------------------------------------------------------------------
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct McLayout {
    int64_t total;           // rows * classes
    int64_t blocks_per_cls;  // ceil(rows / 128)
    int64_t slots_bound, tiles_bound, words_bound;
    size_t cub_bytes;
    size_t off_hdr, off_segs, off_keys, off_keys2, off_vals, off_vals2, off_cls, off_cls2, off_segvals,
        off_cub, off_packed, off_mask, off_flags, off_kr, off_summary, off_sync, off_counts, total_bytes;
    int sum_words;
};

static void mc_layout(int64_t rows, int classes, McLayout *l) {
    l->total = rows * classes;
    l->blocks_per_cls = (rows + kTile - 1) / kTile;
    // worst case: every class keeps all rows
    l->slots_bound = classes * l->blocks_per_cls * kTile;
    l->tiles_bound = classes * (l->blocks_per_cls * (l->blocks_per_cls + 1) / 2);
    l->words_bound = (int64_t)classes * rows * 2 * l->blocks_per_cls;
    size_t cub1 = 0, cub2 = 0;
    cub::DoubleBuffer<uint32_t> dk(nullptr, nullptr), dv(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, cub1, dk, dv, (int)l->total, 0, 32);
    cub::DeviceRadixSort::SortPairs(nullptr, cub2, (const uint8_t *)nullptr, (uint8_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)l->total, 0, 8);
    l->cub_bytes = cub1 > cub2 ? cub1 : cub2;
    size_t off = 0;
    l->off_hdr = off;        off = align_up(off + sizeof(McHeader), 256);
    l->off_segs = off;       off = align_up(off + (size_t)classes * sizeof(NmsSegment), 256);
    l->off_keys = off;       off = align_up(off + (size_t)l->total * 4, 256);
    l->off_keys2 = off;      off = align_up(off + (size_t)l->total * 4, 256);
    l->off_vals = off;       off = align_up(off + (size_t)l->total * 4, 256);
    l->off_vals2 = off;      off = align_up(off + (size_t)l->total * 4, 256);
    l->off_cls = off;        off = align_up(off + (size_t)l->total, 256);
    l->off_cls2 = off;       off = align_up(off + (size_t)l->total, 256);
    l->off_segvals = off;    off = align_up(off + (size_t)l->total * 4, 256);
    l->off_cub = off;        off = align_up(off + l->cub_bytes, 256);
    l->off_packed = off;     off = align_up(off + (size_t)l->slots_bound * kPackedBytesPerBox, 256);
    l->off_mask = off;       off = align_up(off + (size_t)l->words_bound * 8, 256);
    l->off_flags = off;      off = align_up(off + (size_t)l->total, 256);
    l->sum_words = nms_sum_words(l->blocks_per_cls);
    l->off_kr = off;         off = align_up(off + (size_t)(l->slots_bound / 64) * 16, 256);
    l->off_summary = off;    off = align_up(off + (size_t)l->slots_bound * l->sum_words * 8, 256);
    l->off_sync = off;       off = align_up(off + kNmsSyncWords * sizeof(unsigned int), 256);
    l->off_counts = off;     off = align_up(off + kNmsMaxBlocks * sizeof(int), 256);
    l->total_bytes = off;
}

size_t multiclass_nms_workspace_bytes(int64_t rows, int classes) {
    if (rows <= 0 || classes <= 0) return 0;
    McLayout l;
    mc_layout(rows, classes, &l);
    return l.total_bytes;
}

int launch_multiclass_nms(const float *multi_bboxes, int shared, const float *multi_scores,
                          const float *score_factors, int64_t rows, int classes, float score_thr,
                          float iou_threshold, int64_t cap, float *dets_out, int64_t *labels_out,
                          int64_t *keep_flat_out, int64_t *num_out, void *workspace,
                          size_t workspace_bytes, cudaStream_t stream) {
    McLayout l;
    mc_layout(rows, classes, &l);
    if (workspace_bytes < l.total_bytes) return -2;
    char *ws = reinterpret_cast<char *>(workspace);
    McHeader *hdr = reinterpret_cast<McHeader *>(ws + l.off_hdr);
    NmsSegment *segs = reinterpret_cast<NmsSegment *>(ws + l.off_segs);
    uint32_t *keys = reinterpret_cast<uint32_t *>(ws + l.off_keys);
    uint32_t *keys2 = reinterpret_cast<uint32_t *>(ws + l.off_keys2);
    uint32_t *vals = reinterpret_cast<uint32_t *>(ws + l.off_vals);
    uint32_t *vals2 = reinterpret_cast<uint32_t *>(ws + l.off_vals2);
    uint8_t *cls = reinterpret_cast<uint8_t *>(ws + l.off_cls);
    uint8_t *cls2 = reinterpret_cast<uint8_t *>(ws + l.off_cls2);
    uint32_t *segvals = reinterpret_cast<uint32_t *>(ws + l.off_segvals);
    const unsigned grid = (unsigned)((l.total + 255) / 256);

    // zero bits: no candidates, no tiles, maxima below the image of every float
    if (cudaMemsetAsync(hdr, 0, sizeof(McHeader), stream) != cudaSuccess) return -3;
    mc_keys_kernel<<<grid, 256, 0, stream>>>(multi_bboxes, shared, multi_scores, score_factors, rows,
                                             classes, score_thr, keys, vals, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    cub::DoubleBuffer<uint32_t> dk(keys, keys2), dv(vals, vals2);
    size_t cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, dk, dv, (int)l.total, 0, 32,
                                        stream) != cudaSuccess)
        return -3;
    const uint32_t *order32 = dv.Current();
    mc_class_kernel<<<grid, 256, 0, stream>>>(dk.Current(), order32, l.total, classes, cls, hdr, segs);
    if (cudaGetLastError() != cudaSuccess) return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint8_t *)cls, cls2, order32,
                                        segvals, (int)l.total, 0, 8, stream) != cudaSuccess)
        return -3;

    // pack: slots follow the device-made segment table; the gather list is the segment-major order
    PackSet s1 = make_pack_set(multi_bboxes, l.total, 5, nullptr, nullptr, ws + l.off_packed, l.slots_bound);
    s1.segs = segs;
    s1.nseg = classes;
    s1.n_dev = reinterpret_cast<const int64_t *>(&hdr->n_valid);
    s1.gather32 = segvals;
    s1.mc_classes = classes;
    s1.mc_shared = shared;
    s1.mc_max = hdr->mc_max;
    NmsScratch scratch;
    scratch.kr = reinterpret_cast<unsigned long long *>(ws + l.off_kr);
    scratch.summary = reinterpret_cast<unsigned long long *>(ws + l.off_summary);
    scratch.sum_words = l.sum_words;
    scratch.sync = reinterpret_cast<unsigned int *>(ws + l.off_sync);
    scratch.block_counts = reinterpret_cast<int *>(ws + l.off_counts);
    s1.z_kr = scratch.kr;
    s1.z_summary = scratch.summary;
    s1.z_sum_words = scratch.sum_words;
    s1.z_sync = scratch.sync;
    const PackSet none = make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (launch_pack(s1, none, 0.f, 0.f, nms_area_axis_ratio(iou_threshold), stream)) return -3;

    NmsDevLaunch a;
    a.packed = ws + l.off_packed;
    a.n_pad_bound = l.slots_bound;
    a.segs_dev = segs;
    a.nseg = classes;
    a.tiles_bound = l.tiles_bound;
    a.total_tiles_dev = reinterpret_cast<const int64_t *>(&hdr->total_tiles);
    a.scratch = scratch;
    a.max_chunks_bound = 2 * l.blocks_per_cls;
    a.order32 = order32;
    a.seg_order32 = segvals;
    a.n_bound = l.total;
    a.n_dev = reinterpret_cast<const int64_t *>(&hdr->n_valid);
    a.thr = iou_threshold;
    a.mask = reinterpret_cast<uint64_t *>(ws + l.off_mask);
    a.flags = reinterpret_cast<uint8_t *>(ws + l.off_flags);
    a.keep_out = keep_flat_out;
    a.num_keep_out = num_out;
    a.mc.boxes = multi_bboxes;
    a.mc.scores = multi_scores;
    a.mc.factors = score_factors;
    a.mc.classes = classes;
    a.mc.score_stride = classes + 1;
    a.mc.shared = shared;
    a.mc.dets_out = dets_out;
    a.mc.labels_out = labels_out;
    a.mc.cap = cap;
    return launch_nms_device_planned(a, stream);
}


// ============================================================================================
// Generic front-end: score order and segment order of (scores[, ids]) for rotgeo_nms_prepared.
// Replaces what the reference does around mmcv.ops.nms_rotated / batched_nms
// (mmcv/ops/nms.py: scores.sort(descending=True), the per-id loop or the offset trick) and
// what this package used to do with a dozen torch launches: key transform, one 64-bit sort,
// unique() for the segment sizes.  The ONE device->host read of the call sequence — the
// sizes of the segments, which the host needs to lay out the mask — happens here.
//
//   prep_keys_kernel    32-bit key (descending image of the score) + index; extremes of the
//                       ids; max(cx,cy) / max(w,h) for the id-derived centre offsets
//   cub radix sort      stable pairs sort on the 32 key bits: global score order
//   prep_segkey_kernel  16-bit segment key (id - min id) of every entry of that order
//   cub radix sort      stable 16-bit pairs sort: segment-major order
//   cub run-length      sizes of the segments present, ascending id
// ============================================================================================
struct PrepHeader {
    unsigned long long id_max_u, id_min_inv_u;   // images of the extreme ids (atomicMax from zero)
    unsigned int max_xy, max_wh;
    float mc_max[2];
    int range_bad;                               // ids span 2^16 or more
    int n_runs;
};

struct PrepLayout {
    size_t off_hdr, off_keys, off_keys2, off_vals, off_order, off_segkey, off_segkey2, off_seg_order,
        off_runs_unique, off_runs_count, off_cub, total_bytes;
    size_t cub_bytes;
};

static void prep_layout(int64_t n, PrepLayout *l) {
    size_t c1 = 0, c2 = 0, c3 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, c1, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 32);
    cub::DeviceRadixSort::SortPairs(nullptr, c2, (const uint16_t *)nullptr, (uint16_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 16);
    cub::DeviceRunLengthEncode::Encode(nullptr, c3, (const uint16_t *)nullptr, (uint16_t *)nullptr,
                                       (int *)nullptr, (int *)nullptr, (int)n);
    l->cub_bytes = c1 > c2 ? c1 : c2;
    if (c3 > l->cub_bytes) l->cub_bytes = c3;
    size_t off = 0;
    l->off_hdr = off;         off = align_up(off + sizeof(PrepHeader), 256);
    l->off_keys = off;        off = align_up(off + (size_t)n * 4, 256);
    l->off_keys2 = off;       off = align_up(off + (size_t)n * 4, 256);
    l->off_vals = off;        off = align_up(off + (size_t)n * 4, 256);
    l->off_order = off;       off = align_up(off + (size_t)n * 4, 256);
    l->off_segkey = off;      off = align_up(off + (size_t)n * 2, 256);
    l->off_segkey2 = off;     off = align_up(off + (size_t)n * 2, 256);
    l->off_seg_order = off;   off = align_up(off + (size_t)n * 4, 256);
    l->off_runs_unique = off; off = align_up(off + (size_t)n * 2, 256);
    l->off_runs_count = off;  off = align_up(off + (size_t)n * 4, 256);
    l->off_cub = off;         off = align_up(off + l->cub_bytes, 256);
    l->total_bytes = off;
}

size_t nms_prepare_bytes(int64_t n) {
    if (n <= 0) return 0;
    PrepLayout l;
    prep_layout(n, &l);
    return l.total_bytes;
}

__device__ __forceinline__ unsigned long long id_image(int64_t id) {   // order-preserving
    return (unsigned long long)id ^ 0x8000000000000000ull;
}

__global__ void __launch_bounds__(256)
prep_keys_kernel(const float *__restrict__ boxes, int stride, const float *__restrict__ scores,
                 const int64_t *__restrict__ ids, int64_t n, uint32_t *__restrict__ keys,
                 uint32_t *__restrict__ vals, PrepHeader *__restrict__ hdr) {
    __shared__ unsigned long long s_id[2];
    __shared__ unsigned int s_max[2];
    const int tid = threadIdx.x;
    if (tid < 2) { s_id[tid] = 0ull; s_max[tid] = 0u; }
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + tid;
    if (i < n) {
        // descending score: complement of the order-preserving image (+0.0f: -0 -> +0); equal
        // scores keep their index order because the sort is stable
        keys[i] = 0xffffffffu - ordered_bits(scores[i] + 0.0f);
        vals[i] = (uint32_t)i;
        if (ids) {
            const unsigned long long u = id_image(ids[i]);
            atomicMax(&s_id[0], u);
            atomicMax(&s_id[1], ~u);
            const float *b = boxes + i * stride;
            atomicMax(&s_max[0], ordered_bits(fmaxf(b[0], b[1])));
            atomicMax(&s_max[1], ordered_bits(fmaxf(b[2], b[3])));
        }
    }
    __syncthreads();
    if (ids && tid == 0) {
        atomicMax(&hdr->id_max_u, s_id[0]);
        atomicMax(&hdr->id_min_inv_u, s_id[1]);
        atomicMax(&hdr->max_xy, s_max[0]);
        atomicMax(&hdr->max_wh, s_max[1]);
    }
}

__global__ void __launch_bounds__(256)
prep_segkey_kernel(const int64_t *__restrict__ ids, const uint32_t *__restrict__ order, int64_t n,
                   uint16_t *__restrict__ segkey, PrepHeader *hdr) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const unsigned long long lo = ~hdr->id_min_inv_u;
    if (i == 0) {
        if (hdr->id_max_u - lo >= 65536ull) hdr->range_bad = 1;
        hdr->mc_max[0] = ordered_float(hdr->max_xy);
        hdr->mc_max[1] = ordered_float(hdr->max_wh);
    }
    if (i < n) segkey[i] = (uint16_t)(id_image(ids[order[i]]) - lo);
}

// returns 0, -3 (CUDA), -4 (ids span >= 2^16 values), -5 (more segments than seg_cap)
int nms_prepare(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids,
                void *prep, int64_t *seg_counts_host, int64_t seg_cap, int64_t *num_segments_host,
                cudaStream_t stream) {
    PrepLayout l;
    prep_layout(n, &l);
    char *ws = reinterpret_cast<char *>(prep);
    PrepHeader *hdr = reinterpret_cast<PrepHeader *>(ws + l.off_hdr);
    uint32_t *keys = reinterpret_cast<uint32_t *>(ws + l.off_keys);
    uint32_t *keys2 = reinterpret_cast<uint32_t *>(ws + l.off_keys2);
    uint32_t *vals = reinterpret_cast<uint32_t *>(ws + l.off_vals);
    uint32_t *order = reinterpret_cast<uint32_t *>(ws + l.off_order);
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (cudaMemsetAsync(hdr, 0, sizeof(PrepHeader), stream) != cudaSuccess) return -3;
    prep_keys_kernel<<<grid, 256, 0, stream>>>(boxes, stride, scores, ids, n, keys, vals, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    size_t cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint32_t *)keys, keys2,
                                        (const uint32_t *)vals, order, (int)n, 0, 32, stream) != cudaSuccess)
        return -3;
    if (!ids) {
        seg_counts_host[0] = n;
        *num_segments_host = 1;
        return 0;
    }
    uint16_t *segkey = reinterpret_cast<uint16_t *>(ws + l.off_segkey);
    uint16_t *segkey2 = reinterpret_cast<uint16_t *>(ws + l.off_segkey2);
    uint32_t *seg_order = reinterpret_cast<uint32_t *>(ws + l.off_seg_order);
    uint16_t *runs_unique = reinterpret_cast<uint16_t *>(ws + l.off_runs_unique);
    int *runs_count = reinterpret_cast<int *>(ws + l.off_runs_count);
    prep_segkey_kernel<<<grid, 256, 0, stream>>>(ids, order, n, segkey, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint16_t *)segkey, segkey2,
                                        (const uint32_t *)order, seg_order, (int)n, 0, 16, stream) != cudaSuccess)
        return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRunLengthEncode::Encode(ws + l.off_cub, cub_bytes, (const uint16_t *)segkey2, runs_unique,
                                           runs_count, &hdr->n_runs, (int)n, stream) != cudaSuccess)
        return -3;
    // the one device->host read: header + the first seg_cap run lengths
    PrepHeader h;
    const int64_t ncopy = n < seg_cap ? n : seg_cap;
    std::vector<int> counts((size_t)ncopy);
    if (cudaMemcpyAsync(&h, hdr, sizeof(PrepHeader), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaMemcpyAsync(counts.data(), runs_count, (size_t)ncopy * sizeof(int), cudaMemcpyDeviceToHost,
                        stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess)
        return -3;
    if (h.range_bad) return -4;
    if (h.n_runs > seg_cap) return -5;
    for (int r = 0; r < h.n_runs; ++r) seg_counts_host[r] = counts[(size_t)r];
    *num_segments_host = h.n_runs;
    return 0;
}


// ============================================================================================
// Batch of images front-end: what the reference does once per image around batched_nms
// (mmrotate/models/dense_heads/oriented_rpn_head.py:275, rotated_rpn_head.py:414-426,515: a python
// loop over the images of the batch, each with its own boxes.max(), offsets, sort and NMS) for
// ALL images in one pass.  The images lie back to back in (boxes, scores, ids); a segment is one
// (image, id); the centre offsets are mmcv.batched_nms's id * (max(cx,cy) + max(w,h) + 1) with the
// maxima of the box's OWN image; the order the kept boxes are emitted in is image-major (score
// order inside an image), so one more pass turns them into per-image lists:
//
//   batch_keys_kernel    32-bit score key + index, image of every box (16 bits), extremes of the
//                        ids, per-image maxima (block-local reduction, then one atomic per block)
//   cub radix sort       stable 32-bit pairs sort: global score order
//   batch_segkey_kernel  16-bit segment key image * id_range + (id - min id) and the image key of
//                        every entry of that order; per-box centre offsets
//   cub radix sort x2    stable 16-bit sorts of the score order: segment-major and image-major
//   cub run-length       sizes of the segments present
//   batch_finish_kernel  (after the NMS) global kept index -> index inside its image + kept count
//                        per image
// ============================================================================================
struct BatchExtra {
    size_t off_starts, off_mxy, off_mwh, total;
};
static void batch_extra_layout(int64_t nimg, BatchExtra *e) {
    size_t off = 0;
    e->off_starts = off; off = align_up(off + (size_t)(nimg + 1) * 8, 256);
    e->off_mxy = off;    off = align_up(off + (size_t)nimg * 4, 256);
    e->off_mwh = off;    off = align_up(off + (size_t)nimg * 4, 256);
    e->total = off;
}

size_t nms_batch_prepare_bytes(int64_t n, int64_t nimg) {
    if (n <= 0 || nimg <= 0) return 0;
    PrepLayout l;
    prep_layout(n, &l);
    BatchExtra e;
    batch_extra_layout(nimg, &e);
    return l.total_bytes + e.total;
}

// last image whose first row is <= i (images are contiguous row ranges, empty ones allowed)
__device__ __forceinline__ int image_of_row(const int64_t *__restrict__ starts, int nimg, int64_t i) {
    int lo = 0, hi = nimg - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (starts[mid] <= i) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(256)
batch_keys_kernel(const float *__restrict__ boxes, int stride, const float *__restrict__ scores,
                  const int64_t *__restrict__ ids, int64_t n, const int64_t *__restrict__ starts, int nimg,
                  uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, uint16_t *__restrict__ imgkey,
                  unsigned int *__restrict__ mxy, unsigned int *__restrict__ mwh, PrepHeader *__restrict__ hdr) {
    __shared__ unsigned long long s_id[2];
    __shared__ unsigned int s_max[2][2];   // [image - first image of the block][xy, wh]
    __shared__ int s_img0;
    const int tid = threadIdx.x;
    if (tid < 2) { s_id[tid] = 0ull; s_max[tid][0] = 0u; s_max[tid][1] = 0u; }
    if (tid == 0) s_img0 = image_of_row(starts, nimg, blockIdx.x * (int64_t)blockDim.x);
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + tid;
    if (i < n) {
        keys[i] = 0xffffffffu - ordered_bits(scores[i] + 0.0f);
        vals[i] = (uint32_t)i;
        int img = s_img0;
        if (i >= starts[img + 1]) img = image_of_row(starts, nimg, i);
        imgkey[i] = (uint16_t)img;
        const unsigned long long u = id_image(ids[i]);
        atomicMax(&s_id[0], u);
        atomicMax(&s_id[1], ~u);
        const float *b = boxes + i * stride;
        const unsigned int xy = ordered_bits(fmaxf(b[0], b[1])), wh = ordered_bits(fmaxf(b[2], b[3]));
        const int d = img - s_img0;
        if (d < 2) {
            atomicMax(&s_max[d][0], xy);
            atomicMax(&s_max[d][1], wh);
        } else {
            atomicMax(mxy + img, xy);
            atomicMax(mwh + img, wh);
        }
    }
    __syncthreads();
    if (tid == 0) {
        atomicMax(&hdr->id_max_u, s_id[0]);
        atomicMax(&hdr->id_min_inv_u, s_id[1]);
    }
    if (tid < 2 && s_max[tid][0] != 0u) {       // 0 = no box of that image in this block
        atomicMax(mxy + s_img0 + tid, s_max[tid][0]);
        atomicMax(mwh + s_img0 + tid, s_max[tid][1]);
    }
}

__global__ void __launch_bounds__(256)
batch_segkey_kernel(const int64_t *__restrict__ ids, const uint32_t *__restrict__ order,
                    const uint16_t *__restrict__ imgkey, int64_t n, int nimg,
                    const unsigned int *__restrict__ mxy, const unsigned int *__restrict__ mwh,
                    uint16_t *__restrict__ segkey, uint16_t *__restrict__ imgkey_sorted,
                    float *__restrict__ offsets_out, PrepHeader *hdr) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const unsigned long long lo = ~hdr->id_min_inv_u;
    const unsigned long long range = hdr->id_max_u - lo + 1ull;
    if (p == 0 && (range > 65536ull || range * (unsigned long long)nimg > 65536ull)) hdr->range_bad = 1;
    if (p < n) {
        // entry p of the score order
        const uint32_t i = order[p];
        const unsigned img = imgkey[i];
        segkey[p] = (uint16_t)(img * range + (id_image(ids[i]) - lo));
        imgkey_sorted[p] = (uint16_t)img;
        // box p: id * (max(cx,cy) + max(w,h) + 1) with the maxima of its image, the float
        // arithmetic of mmcv.ops.batched_nms (separate roundings, no fused multiply-add)
        const unsigned im = imgkey[p];
        const float m = __fadd_rn(ordered_float(mxy[im]), ordered_float(mwh[im]));
        offsets_out[p] = __fmul_rn((float)ids[p], __fadd_rn(m, 1.0f));
    }
}

// returns 0, -3 (CUDA), -4 (images x id range >= 2^16), -5 (more segments than seg_cap)
int nms_batch_prepare(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids,
                      const int64_t *img_sizes_host, int64_t nimg, void *prep, float *offsets_out,
                      int64_t *seg_counts_host, int64_t seg_cap, int64_t *num_segments_host,
                      cudaStream_t stream) {
    PrepLayout l;
    prep_layout(n, &l);
    BatchExtra x;
    batch_extra_layout(nimg, &x);
    char *ws = reinterpret_cast<char *>(prep);
    char *xs = ws + l.total_bytes;
    PrepHeader *hdr = reinterpret_cast<PrepHeader *>(ws + l.off_hdr);
    uint32_t *keys = reinterpret_cast<uint32_t *>(ws + l.off_keys);
    uint32_t *keys2 = reinterpret_cast<uint32_t *>(ws + l.off_keys2);
    uint32_t *vals = reinterpret_cast<uint32_t *>(ws + l.off_vals);
    uint32_t *order = reinterpret_cast<uint32_t *>(ws + l.off_order);           // image-major score order
    uint32_t *score_order = reinterpret_cast<uint32_t *>(ws + l.off_runs_count);   // free until the run-length pass
    uint16_t *imgkey = reinterpret_cast<uint16_t *>(ws + l.off_runs_unique);       // free until the run-length pass
    uint16_t *imgkey_sorted = reinterpret_cast<uint16_t *>(ws + l.off_keys);       // keys are dead after the first sort
    uint16_t *imgkey_sorted2 = reinterpret_cast<uint16_t *>(ws + l.off_keys2);     // so are the sorted keys
    uint16_t *segkey = reinterpret_cast<uint16_t *>(ws + l.off_segkey);
    uint16_t *segkey2 = reinterpret_cast<uint16_t *>(ws + l.off_segkey2);
    uint32_t *seg_order = reinterpret_cast<uint32_t *>(ws + l.off_seg_order);
    uint16_t *runs_unique = reinterpret_cast<uint16_t *>(ws + l.off_runs_unique);
    int *runs_count = reinterpret_cast<int *>(ws + l.off_runs_count);
    int64_t *starts = reinterpret_cast<int64_t *>(xs + x.off_starts);
    unsigned int *mxy = reinterpret_cast<unsigned int *>(xs + x.off_mxy);
    unsigned int *mwh = reinterpret_cast<unsigned int *>(xs + x.off_mwh);

    std::vector<int64_t> st((size_t)nimg + 1, 0);
    for (int64_t g = 0; g < nimg; ++g) st[(size_t)g + 1] = st[(size_t)g] + img_sizes_host[g];
    if (st[(size_t)nimg] != n) return -1;
    // pageable source: staged by the runtime before the call returns
    if (cudaMemcpyAsync(starts, st.data(), st.size() * sizeof(int64_t), cudaMemcpyHostToDevice, stream) != cudaSuccess ||
        cudaMemsetAsync(hdr, 0, sizeof(PrepHeader), stream) != cudaSuccess ||
        cudaMemsetAsync(mxy, 0, x.total - x.off_mxy, stream) != cudaSuccess)
        return -3;
    const unsigned grid = (unsigned)((n + 255) / 256);
    batch_keys_kernel<<<grid, 256, 0, stream>>>(boxes, stride, scores, ids, n, starts, (int)nimg, keys, vals,
                                                imgkey, mxy, mwh, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    size_t cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint32_t *)keys, keys2,
                                        (const uint32_t *)vals, score_order, (int)n, 0, 32, stream) != cudaSuccess)
        return -3;
    batch_segkey_kernel<<<grid, 256, 0, stream>>>(ids, score_order, imgkey, n, (int)nimg, mxy, mwh, segkey,
                                                  imgkey_sorted, offsets_out, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint16_t *)segkey, segkey2,
                                        (const uint32_t *)score_order, seg_order, (int)n, 0, 16, stream) != cudaSuccess)
        return -3;
    int img_bits = 1;
    while (img_bits < 16 && (1ll << img_bits) < nimg) ++img_bits;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortPairs(ws + l.off_cub, cub_bytes, (const uint16_t *)imgkey_sorted, imgkey_sorted2,
                                        (const uint32_t *)score_order, order, (int)n, 0, img_bits, stream) != cudaSuccess)
        return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRunLengthEncode::Encode(ws + l.off_cub, cub_bytes, (const uint16_t *)segkey2, runs_unique,
                                           runs_count, &hdr->n_runs, (int)n, stream) != cudaSuccess)
        return -3;
    PrepHeader h;
    const int64_t ncopy = n < seg_cap ? n : seg_cap;
    std::vector<int> counts((size_t)ncopy);
    if (cudaMemcpyAsync(&h, hdr, sizeof(PrepHeader), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaMemcpyAsync(counts.data(), runs_count, (size_t)ncopy * sizeof(int), cudaMemcpyDeviceToHost,
                        stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess)
        return -3;
    if (h.range_bad) return -4;
    if (h.n_runs > seg_cap) return -5;
    for (int r = 0; r < h.n_runs; ++r) seg_counts_host[r] = counts[(size_t)r];
    *num_segments_host = h.n_runs;
    return 0;
}

__global__ void __launch_bounds__(256)
batch_finish_kernel(int64_t *__restrict__ keep, const int64_t *__restrict__ num_keep, const int64_t *__restrict__ starts,
                    int nimg, unsigned long long *__restrict__ counts) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= *num_keep) return;
    const int64_t gidx = keep[i];
    const int img = image_of_row(starts, nimg, gidx);
    keep[i] = gidx - starts[img];
    // kept boxes are image-major: most of a warp counts for the same image
    const unsigned peers = __match_any_sync(__activemask(), img);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(counts + img, (unsigned long long)__popc(peers));
}

int nms_batch_finish(int64_t *keep, const int64_t *num_keep_dev, int64_t n, const void *prep, int64_t nimg,
                     int64_t *counts_out_dev, cudaStream_t stream) {
    PrepLayout l;
    prep_layout(n, &l);
    BatchExtra x;
    batch_extra_layout(nimg, &x);
    const char *xs = reinterpret_cast<const char *>(prep) + l.total_bytes;
    const int64_t *starts = reinterpret_cast<const int64_t *>(xs + x.off_starts);
    if (cudaMemsetAsync(counts_out_dev, 0, (size_t)nimg * sizeof(int64_t), stream) != cudaSuccess) return -3;
    batch_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(
        keep, num_keep_dev, starts, (int)nimg, reinterpret_cast<unsigned long long *>(counts_out_dev));
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int launch_nms_prepared(const float *boxes, int64_t n, int stride, const float *offsets,
                        const int64_t *offset_ids, bool segmented, const void *prep, const NmsPlan &plan,
                        const NmsSegment *segs_dev, float thr, void *packed, uint64_t *mask,
                        uint8_t *flags, const NmsScratch &scratch, int64_t *keep_out,
                        int64_t *num_keep_out, cudaStream_t stream) {
    PrepLayout l;
    prep_layout(n, &l);
    const char *ws = reinterpret_cast<const char *>(prep);
    const PrepHeader *hdr = reinterpret_cast<const PrepHeader *>(ws + l.off_hdr);
    const uint32_t *order = reinterpret_cast<const uint32_t *>(ws + l.off_order);
    // prepared without ids: one segment, the segment-major order is the score order
    const uint32_t *seg_order = segmented ? reinterpret_cast<const uint32_t *>(ws + l.off_seg_order) : order;
    PackSet s1 = make_pack_set(boxes, n, stride, nullptr, offsets, packed, plan.n_pad);
    s1.gather32 = seg_order;
    s1.segs = segs_dev;
    s1.nseg = (int)plan.num_segments;
    if (offset_ids) {
        s1.offset_ids = offset_ids;
        s1.mc_max = hdr->mc_max;
    }
    s1.z_kr = scratch.kr;
    s1.z_summary = scratch.summary;
    s1.z_sum_words = scratch.sum_words;
    s1.z_sync = scratch.sync;
    const PackSet none = make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (launch_pack(s1, none, 0.f, 0.f, nms_area_axis_ratio(thr), stream)) return -3;
    return launch_nms(plan, packed, segs_dev, nullptr, nullptr, order, seg_order, thr, mask, flags,
                      scratch, keep_out, num_keep_out, stream);
}

}  // namespace rotgeo
