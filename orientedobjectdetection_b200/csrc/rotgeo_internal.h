// Host-side launch helpers shared by the C-ABI translation unit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rotgeo {

// One segment of a segmented NMS.  `start` counts positions in seg_order (boxes in
// segment-major, score-descending order); `pstart` is the segment's first slot in the
// packed set, a multiple of 128, so every 128-slot table block lies inside one segment.
struct NmsSegment {
    int64_t start;      // first position in seg_order
    int64_t pstart;     // first packed slot (multiple of 128)
    int64_t len;        // boxes in the segment
    int64_t mask_base;  // first u64 word of this segment's suppression mask
    int64_t tile_base;  // first tile id of this segment in the bitmask grid
    int32_t words;      // u64 words per mask row (even)
    int32_t row_blocks; // ceil(len / 64)
};

// One box set to pack: AoS source rows (+ optional gather / centre offsets) and
// the two destination planes.  n_pad == 0 marks an absent second set.
struct PackSet {
    const float *boxes;
    int64_t n;
    int stride;
    const int64_t *gather;   // nullable: packed[i] = boxes[gather[i]]
    const float *offsets;    // nullable: added to cx, cy, indexed like the source rows
    const int64_t *offset_ids;  // nullable: centre offset = id * (mc_max[0] + mc_max[1] + 1) instead
    const NmsSegment *segs;  // nullable: segmented slot layout (see slot_source)
    int nseg;
    // device-planned sets (fused front-ends): the box count lives on the device, and gather32[]
    // holds FLAT candidate indices row * mc_classes + class of a (rows x classes) score matrix
    const int64_t *n_dev;    // nullable: overrides n
    const uint32_t *gather32;  // nullable: 32-bit gather list, used instead of gather
    int mc_classes;          // 0: the gather list holds source rows
    int mc_shared;           // 1: boxes (rows, 5) shared by all classes; 0: (rows * classes, 5)
    const float *mc_max;     // device float[2]: max(cx, cy), max(w, h) of the candidates; the class
                             // offset is class * (mc_max[0] + mc_max[1] + 1), bbox_nms_rotated.py:73-77
    // NMS only: scratch of the resolve stage, zeroed here so that no extra launch is needed
    unsigned long long *z_kr;       // nullable: n_pad / 64 (kept, removed) word pairs
    unsigned long long *z_summary;  // n_pad * z_sum_words row summaries
    int z_sum_words;
    unsigned int *z_sync;           // kNmsSyncWords words of barrier / round state
    float4 *p0, *p1;
    uint32_t *pr, *pt;
    int64_t n_pad;
};
constexpr int kNmsSyncWords = 16;

// The area axis of the inverted index pays off when the threshold is high enough for the
// admissible area range to be narrow (RPN: 0.8 -> a few of the 32 bins); at the usual 0.1 of
// the detection heads it would admit half of the bins and cost more than it removes.
static inline float nms_area_axis_ratio(float iou_threshold) {
    const float r = iou_threshold - 1e-4f;
    return (r >= 0.25f && r < 1.f) ? r : 0.f;
}
PackSet make_pack_set(const float *boxes, int64_t n, int stride, const int64_t *gather,
                      const float *offsets, void *packed, int64_t n_pad);

// cell > 0: bin width of the cull masks; cell <= 0: derived from the set(s) in the launch.
// area_ratio in (0, 1): NMS pack, also builds the area axis for pairs that need
// min(area) >= area_ratio * max(area); otherwise 0
int launch_pack(const PackSet &s1, const PackSet &s2, float min_wh, float cell, float area_ratio,
                cudaStream_t stream);

// after_pack: the pack kernel of these sets was the previous launch on `stream`
int launch_iou_pairwise(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                        float *out, int64_t ld_out, bool after_pack, cudaStream_t stream);

// reductions instead of the matrix (rotgeo_assign.cu): keys = float bits << 32 | ~index, 0 = none
int launch_iou_max(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                   unsigned long long *row_key, unsigned long long *col_key, bool after_pack,
                   cudaStream_t stream);
int launch_iou_row_max_hits(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                            const float *row_max, float min_pos_iou, int *col_lowq, cudaStream_t stream);

// fused assigner path (rotgeo_assign.cu)
size_t iou_max_workspace_bytes(int64_t n1, int64_t n2);
int launch_iou_rowcol_max(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2,
                          int stride2, int mode, float min_wh, float *row_max, int64_t *row_argmax,
                          float *col_max, int64_t *col_argmax, void *workspace, cudaStream_t stream);
int launch_max_iou_assign(const float *gts, int64_t num_gts, int stride_g, const float *boxes,
                          int64_t num_boxes, int stride_b, int mode, float min_wh, float pos_iou_thr,
                          float neg_lo, float neg_hi, float min_pos_iou, int match_low_quality,
                          int gt_max_assign_all, int64_t *assigned, float *max_overlaps, void *workspace,
                          cudaStream_t stream);

// mAP evaluation: TP / FP of every detection, all (class, image) segments at once (rotgeo_eval.cu)
size_t eval_tpfp_workspace_bytes(int64_t num_dets, int64_t num_gts);
int launch_eval_tpfp(const float *dets, int64_t num_dets, int det_stride, const int64_t *det_off,
                     const float *gts, int64_t num_gts, int gt_stride, const int64_t *gt_off,
                     const uint8_t *gt_ignore, int64_t num_segments, float iou_thr, uint8_t *tp,
                     uint8_t *fp, float *iou_max_out, void *workspace, cudaStream_t stream);

// differentiable aligned IoU (rotgeo_diff.cu): value + gradient w.r.t. both boxes, (n, 5) each, nullable
int launch_diff_iou(const float *boxes1, const float *boxes2, int64_t n, int stride1, int stride2,
                    float *iou_out, float *grad1, float *grad2, cudaStream_t stream);

int launch_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1,
                       int stride2, int mode, float min_wh, float *out, cudaStream_t stream);

struct NmsPlan {
    int64_t n;
    int64_t n_pad;       // packed slots: every segment padded to whole 128-slot blocks
    int64_t num_segments;
    int64_t total_tiles;
    int64_t mask_words;
    int64_t max_chunks;  // 64-box chunks of the largest segment (upper bound)
    int sum_words;       // u64 words of a row summary: ceil(column blocks of the largest segment / 64)
    // workspace layout (byte offsets)
    size_t off_packed, off_segments, off_mask, off_flags, off_kr, off_summary, off_sync, off_counts,
        total_bytes;
};

// scratch of the resolve stage (rotgeo_nms.cu); kr / summary / sync are zeroed by the pack kernel
struct NmsScratch {
    unsigned long long *kr;        // per 64-slot word: kept bits, removed bits
    unsigned long long *summary;   // per slot: bit J set = the row has a bit in column block J
    int sum_words;
    unsigned int *sync;
    int *block_counts;             // kNmsMaxBlocks entries
};
constexpr int kNmsMaxBlocks = 1024;
static inline int nms_sum_words(int64_t max_col_blocks) {
    return (int)((max_col_blocks + 63) / 64 > 0 ? (max_col_blocks + 63) / 64 : 1);
}

void nms_plan(int64_t n, const int64_t *seg_offsets_host, int64_t num_segments, NmsPlan *plan,
              NmsSegment *segments_out /* nullable, num_segments entries */);

constexpr int kMcMaxClasses = 64;   // classes the fused multiclass front-end supports

// fused output of the multiclass front-end: rows of (box, score) + labels of the kept candidates
struct McOut {
    const float *boxes;      // (rows, 5) if shared else (rows * classes, 5), contiguous
    const float *scores;     // (rows, score_stride) raw scores
    const float *factors;    // nullable (rows,)
    int classes, score_stride, shared;
    float *dets_out;         // (cap, 6)
    int64_t *labels_out;     // (cap,)
    int64_t cap;             // keep at most this many (max_num)
};

// orders: either the 64-bit pair or the 32-bit pair is given
int launch_nms(const NmsPlan &plan, const void *packed, const NmsSegment *segments_dev,
               const int64_t *order, const int64_t *seg_order, const uint32_t *order32,
               const uint32_t *seg_order32, float thr, uint64_t *mask, uint8_t *flags,
               const NmsScratch &scratch, int64_t *keep_out, int64_t *num_keep_out,
               cudaStream_t stream);

// generic front-end (rotgeo_frontend.cu)
size_t nms_prepare_bytes(int64_t n);
size_t nms_batch_prepare_bytes(int64_t n, int64_t nimg);
int nms_batch_prepare(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids,
                      const int64_t *img_sizes_host, int64_t nimg, void *prep, float *offsets_out,
                      int64_t *seg_counts_host, int64_t seg_cap, int64_t *num_segments_host,
                      cudaStream_t stream);
int nms_batch_finish(int64_t *keep, const int64_t *num_keep_dev, int64_t n, const void *prep, int64_t nimg,
                     int64_t *counts_out_dev, cudaStream_t stream);
int nms_prepare(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids,
                void *prep, int64_t *seg_counts_host, int64_t seg_cap, int64_t *num_segments_host,
                cudaStream_t stream);
int launch_nms_prepared(const float *boxes, int64_t n, int stride, const float *offsets,
                        const int64_t *offset_ids, bool segmented, const void *prep, const NmsPlan &plan,
                        const NmsSegment *segs_dev, float thr, void *packed, uint64_t *mask,
                        uint8_t *flags, const NmsScratch &scratch, int64_t *keep_out,
                        int64_t *num_keep_out, cudaStream_t stream);

// NMS over a plan made on the DEVICE (fused front-ends): grids and buffers are sized for
// host-side upper bounds, the kernels read the actual counts from device memory.
struct NmsDevLaunch {
    const void *packed;
    int64_t n_pad_bound;              // packed slots allocated
    const NmsSegment *segs_dev;
    int nseg;
    int64_t tiles_bound;              // mask grid
    const int64_t *total_tiles_dev;   // actual tile count
    NmsScratch scratch;
    int64_t max_chunks_bound;         // 64-box chunks of the largest possible segment
    const uint32_t *order32;          // global score order (flat candidate indices)
    const uint32_t *seg_order32;      // segment-major order
    int64_t n_bound;
    const int64_t *n_dev;             // actual candidate count
    float thr;
    uint64_t *mask;
    uint8_t *flags;
    int64_t *keep_out, *num_keep_out;
    McOut mc;
};
int launch_nms_device_planned(const NmsDevLaunch &a, cudaStream_t stream);

int launch_multiclass_nms(const float *multi_bboxes, int shared, const float *multi_scores,
                          const float *score_factors, int64_t rows, int classes, float score_thr,
                          float iou_threshold, int64_t cap, float *dets_out, int64_t *labels_out,
                          int64_t *keep_flat_out, int64_t *num_out, void *workspace,
                          size_t workspace_bytes, cudaStream_t stream);
size_t multiclass_nms_workspace_bytes(int64_t rows, int classes);

}  // namespace rotgeo
