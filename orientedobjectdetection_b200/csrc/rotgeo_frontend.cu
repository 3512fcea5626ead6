// Fused front-end of the class-batched rotated NMS.  sm_100a.
//
// Replaces the python body of the reference's multiclass_nms_rotated
// (mmrotate/core/post_processing/bbox_nms_rotated.py:33-92): expand boxes to
// (n*C, 5), threshold the scores, nonzero() (a host sync), three gathers, two
// global max reductions for the class offset, the sort inside mmcv.ops.nms_rotated
// and the final gathers/cat — a dozen small launches and several device->host
// round trips.  Here the whole call is eight launches and NO device->host read:
//
//   mc_keys_kernel   one thread per (row, class): candidate test, 64-bit sort keys
//                    (class | descending score | flat index), per-class counts,
//                    max(cx,cy) / max(w,h) of the candidates
//   cub radix sort   twice, keys only: segment-major order and global score order
//   mc_plan_kernel   segment table and tile count from the class counts (the plan
//                    the host makes in rotgeo_nms, made on the device)
//   pack / mask / scan kernels of rotgeo_nms, launched over host-side upper bounds
//   nms_emit_kernel  global score order, writes (box, score) rows and labels directly
//
// Sizes that depend on the data (candidates, tiles, mask words) never reach the host:
// grids and workspace are sized for the worst case (every class keeps all n rows)
// and the kernels read the actual counts from the device-side header.
#include <cub/device/device_radix_sort.cuh>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kIdxBits = 26;                       // flat index row * C + class
constexpr unsigned long long kIdxMask = (1ull << kIdxBits) - 1ull;
constexpr int kTile = 128;

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

__global__ void mc_init_kernel(McHeader *h) {
    const int t = threadIdx.x;
    if (t == 0) {
        h->n_valid = 0;
        h->total_tiles = 0;
        h->max_xy = 0u;     // below the image of every float
        h->max_wh = 0u;
    }
    if (t < kMcMaxClasses) h->counts[t] = 0;
}

__global__ void __launch_bounds__(256)
mc_keys_kernel(const float *__restrict__ boxes, int shared, const float *__restrict__ scores,
               const float *__restrict__ factors, int64_t rows, int classes, float score_thr,
               unsigned long long *__restrict__ keys_seg, unsigned long long *__restrict__ keys_glob,
               McHeader *__restrict__ hdr) {
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
        unsigned long long ks = ~0ull, kg = ~0ull;
        if (raw > score_thr) {                       // tested BEFORE the factors, like the reference
            float sc = raw;
            if (factors) sc *= factors[row];
            // descending score: complement of the order-preserving image (+0.0f: -0 -> +0)
            const unsigned long long desc = 0xffffffffu - ordered_bits(sc + 0.0f);
            kg = (desc << kIdxBits) | (unsigned long long)f;
            ks = ((unsigned long long)cls << (32 + kIdxBits)) | kg;
            atomicAdd(&s_cnt[cls], 1);
            const float *b = boxes + (shared ? row : f) * 5;
            atomicMax(&s_max[0], ordered_bits(fmaxf(b[0], b[1])));
            atomicMax(&s_max[1], ordered_bits(fmaxf(b[2], b[3])));
        }
        keys_seg[f] = ks;
        keys_glob[f] = kg;
    }
    __syncthreads();
    if (tid < classes && s_cnt[tid]) atomicAdd(&hdr->counts[tid], s_cnt[tid]);
    if (tid == 0 && s_max[0]) atomicMax(&hdr->max_xy, s_max[0]);
    if (tid == 1 && s_max[1]) atomicMax(&hdr->max_wh, s_max[1]);
}

// the plan nms_plan() makes on the host, on the device (one thread: <= 64 segments)
__global__ void mc_plan_kernel(McHeader *hdr, NmsSegment *segs, int classes) {
    if (threadIdx.x != 0) return;
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
    size_t off_hdr, off_segs, off_keys_seg, off_keys_seg2, off_keys_glob, off_keys_glob2, off_cub,
        off_packed, off_mask, off_flags, total_bytes;
};

static int sort_end_bit(int classes) {
    int cbits = 0;
    while ((1 << cbits) < classes) ++cbits;
    return 32 + kIdxBits + cbits;
}

static void mc_layout(int64_t rows, int classes, McLayout *l) {
    l->total = rows * classes;
    l->blocks_per_cls = (rows + kTile - 1) / kTile;
    // worst case: every class keeps all rows
    l->slots_bound = classes * l->blocks_per_cls * kTile;
    l->tiles_bound = classes * (l->blocks_per_cls * (l->blocks_per_cls + 1) / 2);
    l->words_bound = (int64_t)classes * rows * 2 * l->blocks_per_cls;
    size_t cub = 0;
    cub::DoubleBuffer<unsigned long long> db(nullptr, nullptr);
    cub::DeviceRadixSort::SortKeys(nullptr, cub, db, (int)l->total, 0, sort_end_bit(classes));
    l->cub_bytes = cub;
    size_t off = 0;
    l->off_hdr = off;        off = align_up(off + sizeof(McHeader), 256);
    l->off_segs = off;       off = align_up(off + (size_t)classes * sizeof(NmsSegment), 256);
    l->off_keys_seg = off;   off = align_up(off + (size_t)l->total * 8, 256);
    l->off_keys_seg2 = off;  off = align_up(off + (size_t)l->total * 8, 256);
    l->off_keys_glob = off;  off = align_up(off + (size_t)l->total * 8, 256);
    l->off_keys_glob2 = off; off = align_up(off + (size_t)l->total * 8, 256);
    l->off_cub = off;        off = align_up(off + cub, 256);
    l->off_packed = off;     off = align_up(off + (size_t)l->slots_bound * kPackedBytesPerBox, 256);
    l->off_mask = off;       off = align_up(off + (size_t)l->words_bound * 8, 256);
    l->off_flags = off;      off = align_up(off + (size_t)l->total, 256);
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
    unsigned long long *ks = reinterpret_cast<unsigned long long *>(ws + l.off_keys_seg);
    unsigned long long *ks2 = reinterpret_cast<unsigned long long *>(ws + l.off_keys_seg2);
    unsigned long long *kg = reinterpret_cast<unsigned long long *>(ws + l.off_keys_glob);
    unsigned long long *kg2 = reinterpret_cast<unsigned long long *>(ws + l.off_keys_glob2);

    mc_init_kernel<<<1, 64, 0, stream>>>(hdr);
    mc_keys_kernel<<<(unsigned)((l.total + 255) / 256), 256, 0, stream>>>(
        multi_bboxes, shared, multi_scores, score_factors, rows, classes, score_thr, ks, kg, hdr);
    if (cudaGetLastError() != cudaSuccess) return -3;
    // invalid entries are all ones and sort behind every candidate
    cub::DoubleBuffer<unsigned long long> dbs(ks, ks2), dbg(kg, kg2);
    size_t cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortKeys(ws + l.off_cub, cub_bytes, dbs, (int)l.total, 0,
                                       sort_end_bit(classes), stream) != cudaSuccess)
        return -3;
    cub_bytes = l.cub_bytes;
    if (cub::DeviceRadixSort::SortKeys(ws + l.off_cub, cub_bytes, dbg, (int)l.total, 0,
                                       32 + kIdxBits, stream) != cudaSuccess)
        return -3;
    mc_plan_kernel<<<1, 32, 0, stream>>>(hdr, segs, classes);
    if (cudaGetLastError() != cudaSuccess) return -3;

    // pack: slots follow the device-made segment table; gather[] = sorted keys (index in the low bits)
    PackSet s1 = make_pack_set(multi_bboxes, l.total, 5, reinterpret_cast<const int64_t *>(dbs.Current()),
                               nullptr, ws + l.off_packed, l.slots_bound);
    s1.segs = segs;
    s1.nseg = classes;
    s1.n_dev = reinterpret_cast<const int64_t *>(&hdr->n_valid);
    s1.gather_mask = (int64_t)kIdxMask;
    s1.mc_classes = classes;
    s1.mc_shared = shared;
    s1.mc_max = hdr->mc_max;
    const PackSet none = make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (launch_pack(s1, none, 0.f, 0.f, stream)) return -3;

    NmsDevLaunch a;
    a.packed = ws + l.off_packed;
    a.n_pad_bound = l.slots_bound;
    a.segs_dev = segs;
    a.nseg = classes;
    a.tiles_bound = l.tiles_bound;
    a.total_tiles_dev = reinterpret_cast<const int64_t *>(&hdr->total_tiles);
    a.max_words_bound = (int)(2 * l.blocks_per_cls);
    a.order = reinterpret_cast<const int64_t *>(dbg.Current());
    a.seg_order = reinterpret_cast<const int64_t *>(dbs.Current());
    a.idx_mask = (int64_t)kIdxMask;
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

}  // namespace rotgeo
