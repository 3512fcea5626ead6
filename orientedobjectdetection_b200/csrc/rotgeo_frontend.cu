// Front-ends of the rotated NMS: what runs between the caller's (boxes, scores[, ids]) and the
// pack / mask / resolve kernels of rotgeo_nms.cu.  sm_100a.
//
// Every front-end is the SAME launch sequence and contains no device->host read:
//
//   sort (rotgeo_sort.cu)  one cluster per group: candidate filter, score order, segment-major
//                          order, centre offsets, segment table + launch totals in a device header
//   pack                   gathers the boxes in segment-major order into the packed planes
//   mask, resolve          over host-side UPPER BOUNDS of the slot / tile / mask-word counts; the
//                          kernels read the actual counts from the device header
//
//   * launch_multiclass_nms   the whole body of the reference's multiclass_nms_rotated
//                             (mmrotate/core/post_processing/bbox_nms_rotated.py:33-92): candidates
//                             are the (row, class) pairs with score > score_thr, one group, one
//                             segment per class; the resolve kernel's compaction writes the
//                             (box, score) rows and labels directly.
//   * nms_sort + nms_run      mmcv.ops.nms_rotated / mmcv.ops.batched_nms (one group, one segment
//                             per id) and the RPN heads' per-image python loop around batched_nms
//                             (rotated_rpn_head.py:414-426,515, oriented_rpn_head.py:275) or the
//                             per-class loop of merge_results (merge_results.py:90-119) as ONE
//                             pass: one group per image / class.
//
// When the worst-case workspace of a job is too large to be worth allocating (hundreds of
// thousands of boxes in one call), the caller reads the 64-byte header back once (nms_read_plan)
// and runs the same kernels over the exact counts.  Round 1 ran 5-7 cub::DeviceRadixSort passes
// and a run-length encode here, and always synchronised to hand the segment sizes to the host.
#include <vector>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTile = 128;

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- workspace of pack / mask / resolve for given (bounds of the) counts ---------------------------
void nms_run_layout(int64_t n, const NmsSizes &z, bool spatial, NmsRunLayout *l) {
    size_t off = 0;
    l->sum_words = nms_sum_words(z.max_blocks);
    l->off_packed = off;   off = align_up(off + (size_t)z.n_pad * kPackedBytesPerBox, 256);
    l->off_mask = off;     off = align_up(off + (size_t)z.mask_words * 8, 256);
    l->off_flags = off;    off = align_up(off + (size_t)n, 256);
    l->off_kr = off;       off = align_up(off + (size_t)(z.n_pad / 64) * 16, 256);
    l->off_summary = off;  off = align_up(off + (size_t)z.n_pad * l->sum_words * 8, 256);
    l->off_sync = off;     off = align_up(off + kNmsSyncWords * sizeof(unsigned int), 256);
    l->off_counts = off;   off = align_up(off + kNmsMaxBlocks * sizeof(int), 256);
    l->off_rank = off;     off = align_up(off + (spatial ? (size_t)z.n_pad * 4 : 0), 256);
    l->off_aabb = off;     off = align_up(off + (spatial ? (size_t)(z.n_pad / kTile) * 16 : 0), 256);
    l->off_tiles = off;    off = align_up(off + (spatial ? (size_t)z.tiles * 8 : 0), 256);
    l->total_bytes = off;
}

// worst case of the counts the device plan can produce, from what the host knows: the group sizes
// and the cap on the segments of a group
void nms_bounds(const int64_t *group_sizes_host, int64_t ngroups, int64_t n, int64_t seg_cap, NmsSizes *z) {
    int64_t blocks = 0, tiles = 0, words = 0, maxc = 0, nseg = 0;
    for (int64_t g = 0; g < ngroups; ++g) {
        const int64_t ng = group_sizes_host ? group_sizes_host[g] : n;
        if (ng <= 0) continue;
        const int64_t C = (ng + kTile - 1) / kTile;
        const int64_t S = seg_cap < ng ? seg_cap : ng;
        const int64_t B = C + S - 1;          // every further segment adds at most one partial block
        blocks += B;
        // tiles = sum of C_i (C_i + 1) / 2 over the segments, with sum C_i <= C + k - 1 for k <= S segments and every
        // C_i >= 1: convex, so the worst case is one segment with C blocks and k - 1 single-block ones
        // (B (B + 1) / 2, all blocks in one segment, was 16 times the possible for one C4 image)
        tiles += C * (C + 1) / 2 + (S - 1);
        words += ng * 2 * C;                  // a row of a segment has at most 2 C words
        maxc = C > maxc ? C : maxc;
        nseg += S;
    }
    z->n_pad = blocks > 0 ? blocks * kTile : kTile;
    z->tiles = tiles;
    z->mask_words = words;
    z->max_blocks = maxc > 0 ? maxc : 1;
    z->nseg = nseg < kSortSegsMax ? nseg : kSortSegsMax;
    z->n_emit = n;
}

// ---- kept global rows -> rows inside their group + kept boxes per group -----------------------------
struct GroupTable {
    int ngroups;
    const int64_t *gstart;                           // device (ngroups + 1), or null: inline
    int64_t gstart_inline[kSortGroupsInline + 1];
};
__device__ __forceinline__ int64_t gt_start(const GroupTable &t, int g) {
    return t.gstart ? t.gstart[g] : t.gstart_inline[g];
}

__global__ void __launch_bounds__(256)
group_finish_kernel(int64_t *__restrict__ keep, const int64_t *__restrict__ num_keep, const GroupTable t,
                    unsigned long long *__restrict__ counts) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= *num_keep) return;
    const int64_t row = keep[i];
    int lo = 0, hi = t.ngroups - 1;     // last group whose first row is <= row (empty groups allowed)
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (gt_start(t, mid) <= row) lo = mid; else hi = mid - 1;
    }
    keep[i] = row - gt_start(t, lo);
    // kept boxes are group-major: most of a warp counts for the same group
    const unsigned peers = __match_any_sync(__activemask(), lo);
    if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(counts + lo, (unsigned long long)__popc(peers));
}

// ---- generic front-end -----------------------------------------------------------------------------
int nms_sort(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids, int id_offsets,
             const int64_t *group_sizes_host, int64_t ngroups, int64_t seg_cap, void *sortws, cudaStream_t stream) {
    int64_t max_group = 0;
    for (int64_t g = 0; g < ngroups; ++g) {
        const int64_t ng = group_sizes_host ? group_sizes_host[g] : n;
        max_group = ng > max_group ? ng : max_group;
    }
    SortLayout l;
    sort_layout(n, ngroups, seg_cap, max_group, &l);
    SortArgs a = {};
    a.mode = 0;
    a.boxes = boxes;
    a.stride = stride;
    a.scores = scores;
    a.ids = ids;
    a.id_offsets = ids ? id_offsets : 0;
    a.ngroups = (int)ngroups;
    a.seg_cap = (int)seg_cap;
    a.spatial = nms_spatial(max_group) ? 1 : 0;
    return launch_sort(a, group_sizes_host, n, sortws, l, stream);
}

size_t nms_sort_bytes(int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group) {
    SortLayout l;
    sort_layout(n, ngroups, seg_cap, max_group, &l);
    return l.total_bytes;
}

// synchronising read of the device plan (the ONE device->host read of the large-job path)
int nms_read_plan(const void *sortws, int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group, NmsSizes *z,
                  unsigned int *error, cudaStream_t stream) {
    SortLayout l;
    sort_layout(n, ngroups, seg_cap, max_group, &l);
    SortHeader h;
    if (cudaMemcpyAsync(&h, reinterpret_cast<const char *>(sortws) + l.off_hdr, sizeof(SortHeader),
                        cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess)
        return -3;
    z->n_pad = h.n_pad;
    z->tiles = h.total_tiles;
    z->mask_words = h.mask_words;
    z->max_blocks = h.max_blocks > 0 ? h.max_blocks : 1;
    z->nseg = h.nseg;
    z->n_emit = h.n_valid;
    *error = h.error;
    return 0;
}

int nms_run(const float *boxes, int64_t n, int stride, const float *offsets, int sort_offsets,
            const int64_t *group_sizes_host, int64_t ngroups, int64_t seg_cap, const void *sortws, const NmsSizes &z,
            float thr, int64_t *keep_out, int64_t *num_keep_out, int64_t *group_counts_out, void *workspace,
            cudaStream_t stream, const float *det_scores, float *dets_out) {
    int64_t max_group = 0;
    for (int64_t g = 0; g < ngroups; ++g) {
        const int64_t ng = group_sizes_host ? group_sizes_host[g] : n;
        max_group = ng > max_group ? ng : max_group;
    }
    SortLayout sl;
    sort_layout(n, ngroups, seg_cap, max_group, &sl);
    const char *sw = reinterpret_cast<const char *>(sortws);
    const SortHeader *hdr = reinterpret_cast<const SortHeader *>(sw + sl.off_hdr);
    const NmsSegment *segs = reinterpret_cast<const NmsSegment *>(sw + sl.off_segs);
    const uint32_t *order = reinterpret_cast<const uint32_t *>(sw + sl.off_order);
    const uint32_t *seg_order = reinterpret_cast<const uint32_t *>(sw + sl.off_seg_order);
    const bool spatial = nms_spatial(max_group);
    NmsRunLayout l;
    nms_run_layout(n, z, spatial, &l);
    char *ws = reinterpret_cast<char *>(workspace);

    PackSet s1 = make_pack_set(boxes, n, stride, nullptr,
                               sort_offsets ? reinterpret_cast<const float *>(sw + sl.off_offsets) : offsets,
                               ws + l.off_packed, z.n_pad);
    s1.gather32 = seg_order;
    s1.segs = segs;
    s1.nseg = (int)z.nseg;
    s1.nseg_dev = &hdr->nseg;
    s1.n_pad_dev = &hdr->n_pad;
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
    if (spatial) {
        s1.rank_src = reinterpret_cast<const uint32_t *>(sw + sl.off_seg_rank);
        s1.rank_out = reinterpret_cast<uint32_t *>(ws + l.off_rank);
        s1.aabb_out = reinterpret_cast<float4 *>(ws + l.off_aabb);
    }
    const PackSet none = make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (launch_pack(s1, none, 0.f, 0.f, nms_area_axis_ratio(thr), stream)) return -3;

    NmsDevLaunch a = {};
    a.packed = ws + l.off_packed;
    a.n_pad_bound = z.n_pad;
    a.n_pad_dev = &hdr->n_pad;
    a.segs_dev = segs;
    a.nseg = (int)z.nseg;
    a.nseg_dev = &hdr->nseg;
    a.error_dev = &hdr->error;
    if (spatial) {
        a.rank = s1.rank_out;
        a.aabb = s1.aabb_out;
        a.tile_list = reinterpret_cast<unsigned long long *>(ws + l.off_tiles);
    }
    a.tiles_bound = z.tiles;
    a.total_tiles_dev = reinterpret_cast<const int64_t *>(&hdr->total_tiles);
    a.scratch = scratch;
    a.max_chunks_bound = 2 * z.max_blocks;
    a.order32 = order;
    a.seg_order32 = seg_order;
    a.n_bound = n;
    a.n_dev = nullptr;
    a.thr = thr;
    a.mask = reinterpret_cast<uint64_t *>(ws + l.off_mask);
    a.flags = reinterpret_cast<uint8_t *>(ws + l.off_flags);
    a.keep_out = keep_out;
    a.num_keep_out = num_keep_out;
    if (dets_out && det_scores) {
        a.mc.boxes = boxes;
        a.mc.box_stride = stride;
        a.mc.scores = det_scores;
        a.mc.classes = 0;
        a.mc.dets_out = dets_out;
        a.mc.cap = n;
    }
    int rc = launch_nms_device_planned(a, stream);
    if (rc) return rc;
    if (group_counts_out) {
        GroupTable t;
        t.ngroups = (int)ngroups;
        t.gstart = nullptr;
        if (ngroups <= kSortGroupsInline) {
            int64_t s = 0;
            for (int64_t g = 0; g < ngroups; ++g) {
                t.gstart_inline[g] = s;
                s += group_sizes_host ? group_sizes_host[g] : n;
            }
            t.gstart_inline[ngroups] = s;
        } else {
            t.gstart = reinterpret_cast<const int64_t *>(sw + sl.off_gstart);    // uploaded by launch_sort
        }
        if (cudaMemsetAsync(group_counts_out, 0, (size_t)ngroups * sizeof(int64_t), stream) != cudaSuccess) return -3;
        group_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(
            keep_out, num_keep_out, t, reinterpret_cast<unsigned long long *>(group_counts_out));
        if (cudaGetLastError() != cudaSuccess) return -3;
    }
    return 0;
}

// ---- multiclass front-end ---------------------------------------------------------------------------
struct McLayout {
    int64_t total;             // rows * classes
    NmsSizes z;                // worst case: every class keeps all rows
    SortLayout sort;
    NmsRunLayout run;
    size_t off_run, total_bytes;
};

static void mc_layout(int64_t rows, int classes, McLayout *l) {
    l->total = rows * classes;
    const int64_t bpc = (rows + kTile - 1) / kTile;
    l->z.n_pad = classes * bpc * kTile;
    l->z.tiles = classes * (bpc * (bpc + 1) / 2);
    l->z.mask_words = (int64_t)classes * rows * 2 * bpc;
    l->z.max_blocks = bpc;
    l->z.nseg = classes;
    l->z.n_emit = l->total;
    sort_layout(l->total, 1, classes, l->total, &l->sort);
    nms_run_layout(l->total, l->z, false, &l->run);
    l->off_run = align_up(l->sort.total_bytes, 256);
    l->total_bytes = l->off_run + l->run.total_bytes;
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
    char *sw = reinterpret_cast<char *>(workspace);
    char *ws = sw + l.off_run;

    SortArgs sa = {};
    sa.mode = 1;
    sa.boxes = multi_bboxes;
    sa.stride = 5;
    sa.scores = multi_scores;
    sa.factors = score_factors;
    sa.classes = classes;
    sa.shared = shared;
    sa.score_thr = score_thr;
    sa.ngroups = 1;
    sa.seg_cap = classes;
    int rc = launch_sort(sa, nullptr, l.total, sw, l.sort, stream);
    if (rc) return rc == -1 ? -1 : -3;
    const SortHeader *hdr = sa.hdr;

    // pack: slots follow the device-made segment table; the gather list is the segment-major order
    PackSet s1 = make_pack_set(multi_bboxes, l.total, 5, nullptr, nullptr, ws + l.run.off_packed, l.z.n_pad);
    s1.segs = sa.segs;
    s1.nseg = classes;
    s1.nseg_dev = &hdr->nseg;
    s1.n_pad_dev = &hdr->n_pad;
    s1.n_dev = reinterpret_cast<const int64_t *>(&hdr->n_valid);
    s1.gather32 = sa.seg_order;
    s1.mc_classes = classes;
    s1.mc_shared = shared;
    s1.mc_max = hdr->mc_max;
    NmsScratch scratch;
    scratch.kr = reinterpret_cast<unsigned long long *>(ws + l.run.off_kr);
    scratch.summary = reinterpret_cast<unsigned long long *>(ws + l.run.off_summary);
    scratch.sum_words = l.run.sum_words;
    scratch.sync = reinterpret_cast<unsigned int *>(ws + l.run.off_sync);
    scratch.block_counts = reinterpret_cast<int *>(ws + l.run.off_counts);
    s1.z_kr = scratch.kr;
    s1.z_summary = scratch.summary;
    s1.z_sum_words = scratch.sum_words;
    s1.z_sync = scratch.sync;
    const PackSet none = make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (launch_pack(s1, none, 0.f, 0.f, nms_area_axis_ratio(iou_threshold), stream)) return -3;

    NmsDevLaunch a = {};
    a.packed = ws + l.run.off_packed;
    a.n_pad_bound = l.z.n_pad;
    a.n_pad_dev = &hdr->n_pad;
    a.segs_dev = sa.segs;
    a.nseg = classes;
    a.nseg_dev = &hdr->nseg;
    a.error_dev = nullptr;
    a.tiles_bound = l.z.tiles;
    a.total_tiles_dev = reinterpret_cast<const int64_t *>(&hdr->total_tiles);
    a.scratch = scratch;
    a.max_chunks_bound = 2 * l.z.max_blocks;
    a.order32 = sa.order;
    a.seg_order32 = sa.seg_order;
    a.n_bound = l.total;
    a.n_dev = reinterpret_cast<const int64_t *>(&hdr->n_valid);
    a.thr = iou_threshold;
    a.mask = reinterpret_cast<uint64_t *>(ws + l.run.off_mask);
    a.flags = reinterpret_cast<uint8_t *>(ws + l.run.off_flags);
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
