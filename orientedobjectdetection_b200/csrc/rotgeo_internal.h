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
    const long long *nseg_dev;   // nullable: the segment count lives on the device (nseg is then a bound)
    const long long *n_pad_dev;  // nullable: slots actually used (n_pad is then a bound)
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
    // spatially packed NMS sets: the score rank of every slot and one bounding box per 128-slot block
    const uint32_t *rank_src;       // nullable: rank of every position of the gather list
    uint32_t *rank_out;             // per slot (0xffffffff: padding)
    float4 *aabb_out;               // per block: (x lo, y lo, x hi, y hi) of the circumscribed circles; empty: lo > hi
    float4 *p0, *p1;
    uint32_t *pr, *pt;
    int64_t n_pad;
};
constexpr int kNmsSyncWords = 16;     // [0..1] grid barrier of the resolve kernel, [8..9] 64-bit length of the tile list, [10..11] tiles handed out

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
    int classes, score_stride, shared;   // classes == 0: plain rows — kept index = box row, boxes have box_stride
    int box_stride;                      // floats between rows and scores is (n,), no labels

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

// counts that size the workspace and the grids of pack / mask / resolve: exact (read back from
// the device header) or upper bounds (made on the host from the group sizes)
struct NmsSizes {
    int64_t n_pad, tiles, mask_words, max_blocks, nseg, n_emit;
};
struct NmsRunLayout {
    size_t off_packed, off_mask, off_flags, off_kr, off_summary, off_sync, off_counts, total_bytes;
    size_t off_rank, off_aabb, off_tiles;     // spatially packed launches only
    int sum_words;
};
void nms_run_layout(int64_t n, const NmsSizes &z, bool spatial, NmsRunLayout *l);
void nms_bounds(const int64_t *group_sizes_host, int64_t ngroups, int64_t n, int64_t seg_cap, NmsSizes *z);

// generic front-end (rotgeo_frontend.cu): sort -> [read plan] -> run
size_t nms_sort_bytes(int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group);
int nms_sort(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *ids, int id_offsets,
             const int64_t *group_sizes_host, int64_t ngroups, int64_t seg_cap, void *sortws, cudaStream_t stream);
int nms_read_plan(const void *sortws, int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group, NmsSizes *z,
                  unsigned int *error, cudaStream_t stream);
// det_scores + dets_out (both nullable): the (kept, 6) rows cat(boxes[keep], scores[keep]) the reference's
// nms_rotated / batched_nms return, written by the resolve kernel's compaction
int nms_run(const float *boxes, int64_t n, int stride, const float *offsets, int sort_offsets,
            const int64_t *group_sizes_host, int64_t ngroups, int64_t seg_cap, const void *sortws, const NmsSizes &z,
            float thr, int64_t *keep_out, int64_t *num_keep_out, int64_t *group_counts_out, void *workspace,
            cudaStream_t stream, const float *det_scores = nullptr, float *dets_out = nullptr);

// NMS over a plan made on the DEVICE (fused front-ends): grids and buffers are sized for
// host-side upper bounds, the kernels read the actual counts from device memory.
struct NmsDevLaunch {
    const void *packed;
    int64_t n_pad_bound;              // packed slots allocated
    const long long *n_pad_dev;       // nullable: slots actually used
    const NmsSegment *segs_dev;
    int nseg;                         // segments (a bound when nseg_dev is set)
    const long long *nseg_dev;        // nullable: actual segment count
    const unsigned int *error_dev;    // nullable: front-end error bits, copied to num_keep_out[1]
    // spatially packed set (null otherwise): score rank per slot, bounding box per block, room for the tile list
    const uint32_t *rank;
    const float4 *aabb;
    unsigned long long *tile_list;
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

// ---- sort front-end (rotgeo_sort.cu) ------------------------------------------------------------
constexpr int kSortThreads = 1024;
constexpr int kSortItems = 8;
constexpr int kSortCap = kSortThreads * kSortItems;   // elements per CTA: 8192
constexpr int kSortMaxCluster = 8;                    // portable cluster size: 65 536 elements per group in shared memory
constexpr int kSortGroupsInline = 64;                 // group tables up to this size travel as kernel parameters
constexpr int64_t kSortSegsMax = 16384;               // segments one call handles
constexpr unsigned int kSortErrRange = 1u;            // the ids of a group span 2^16 values or more
constexpr int64_t kSpatialMinGroup = 16384;          // launches whose largest group has at least this many boxes pack spatially
static inline bool nms_spatial(int64_t max_group) { return max_group >= kSpatialMinGroup; }
constexpr unsigned int kSortErrSegments = 2u;         // more segments than seg_cap in a group / than kSortSegsMax in all

struct SortHeader {            // device-side result of the front-end (zeroed before the launch)
    unsigned int done;         // groups finished (ticket of the last-cluster plan)
    unsigned int error;        // kSortErr* bits
    long long n_valid;         // entries of `order` to emit (multiclass: the candidates)
    long long nseg, n_pad, total_tiles, mask_words;
    int max_blocks;            // 128-box blocks of the largest segment
    int pad;
    float mc_max[2];           // multiclass: max(cx, cy), max(w, h) of the candidates
};

struct SortArgs {
    int mode;                  // 0: scores (n,) [+ ids (n,)];  1: multiclass score matrix (rows, classes + 1)
    const float *boxes;
    int stride;
    const float *scores;
    const int64_t *ids;        // nullable (mode 0)
    const float *factors;      // nullable (mode 1)
    int classes, shared;       // mode 1
    float score_thr;           // mode 1
    int id_offsets;            // mode 0 with ids: offsets_out[row] = float(id) * (max(cx,cy) + max(w,h) + 1) of the row's group
    int ngroups;
    int seg_cap;               // segments per group
    int cta_cap;               // elements per CTA buffer of the cluster kernel (set by launch_sort)
    int spatial;               // pack order inside a segment: Morton order of the centres instead of score order
                               // (large sparse segments: only neighbouring 128-box blocks form mask tiles); mode 0 only
    const int64_t *gstart;     // device (ngroups + 1) row offsets, or null: gstart_inline
    int64_t gstart_inline[kSortGroupsInline + 1];
    // outputs (carved from the sort workspace by launch_sort)
    SortHeader *hdr;
    uint32_t *order;           // emission order: group-major, score-descending; global rows (mode 1: flat indices)
    uint32_t *seg_order;       // pack order: group-major, id-ascending, score-descending
    uint32_t *seg_rank;        // position inside the group's score order of every seg_order entry
    float *offsets_out;
    int *gseg_count, *gseg_off;
    uint32_t *gseg_start;      // (ngroups, seg_cap + 1)
    NmsSegment *segs;
    int segs_cap;
    // oversize groups
    uint32_t *big_buf;
    uint32_t big_cap;
    unsigned int *big_hist;
    void *big_red;
    unsigned int *big_count;
};

struct SortLayout {
    size_t off_hdr, off_order, off_seg_order, off_seg_rank, off_offsets, off_gseg_count, off_gseg_off,
        off_gseg_start, off_segs, off_gstart, off_big_buf, off_big_hist, off_big_red, off_big_count, total_bytes;
    int64_t segs_cap, big_cap;
};
void sort_layout(int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group, SortLayout *l);
// fills the output pointers of `a` from the workspace and launches; group_sizes_host null = one group of n.
// returns 0, -1 (sizes), -3 (CUDA), -6 (a group beyond what one cooperative grid sorts)
int launch_sort(SortArgs &a, const int64_t *group_sizes_host, int64_t n, void *sortws, const SortLayout &l,
                cudaStream_t stream);

// ---- multi-GPU exchange over peer memory (rotgeo_dist.cu) -----------------------------------------
constexpr int kPeerMax = 16;                          // ranks of one window
constexpr size_t kPeerPadBytes = 4096;                // signal pad at the start of every window
constexpr int kPeerErrorSlot = 64;                    // u64 slot of the pad raised when a barrier times out
constexpr unsigned long long kPeerTimeoutNs = 2000000000ull;   // 2 s
struct PeerTable {
    void *base[kPeerMax];      // window of every rank, mapped into this process
    int world, rank;
};
int launch_peer_barrier(const PeerTable &t, unsigned long long epoch, cudaStream_t stream);
int launch_dist_iou_allgather(const PeerTable &t, unsigned long long epoch, size_t matrix_off,
                              const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                              int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                              void *workspace, cudaStream_t stream);
int launch_dist_iou_rowcol_max(const PeerTable &t, unsigned long long epoch, size_t stage_off,
                               const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                               int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                               float *row_max, int64_t *row_argmax, float *col_max, int64_t *col_argmax,
                               void *workspace, cudaStream_t stream);
int launch_dist_push_keeps(const PeerTable &t, unsigned long long epoch, size_t out_off, const int64_t *keep,
                           const int64_t *tail, int nlocal, const int64_t *gid_host, int64_t cap,
                           int64_t ngroups_total, const float *rows_src, int row_floats, int row_stride,
                           const int64_t *group_base_host, cudaStream_t stream);
// rotgeo_assign.cu: pack + reduction sink only; the packed (value, ~index) keys stay in the workspace
int launch_iou_max_keys(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2, int stride2,
                        int mode, float min_wh, void *workspace, unsigned long long **row_key,
                        unsigned long long **col_key, cudaStream_t stream);

int launch_multiclass_nms(const float *multi_bboxes, int shared, const float *multi_scores,
                          const float *score_factors, int64_t rows, int classes, float score_thr,
                          float iou_threshold, int64_t cap, float *dets_out, int64_t *labels_out,
                          int64_t *keep_flat_out, int64_t *num_out, void *workspace,
                          size_t workspace_bytes, cudaStream_t stream);
size_t multiclass_nms_workspace_bytes(int64_t rows, int classes);

}  // namespace rotgeo
