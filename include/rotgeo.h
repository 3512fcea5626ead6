/*
 * rotgeo — B200-native (sm_100a) rotated-box IoU and rotated NMS.
 * C ABI of librotgeo.so: the drop-in boundary for the reference's rotated
 * geometry hot path.  Plain pointers and sizes only.
 *
 * Every entry point replaces a native op the reference reaches through
 * `mmcv.ops` (mmcv-full `_ext`, un-vendored; the reference-side bindings are
 * the python names imported at the file:line cited on each function).
 *
 * Conventions
 *   - all data pointers are DEVICE pointers unless a name ends in `_host`;
 *   - the caller owns every buffer (inputs, outputs, workspace); the library
 *     never allocates or frees device memory and keeps no pointer after return;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value: 0 = ok, negative = error (rotgeo_last_error() has text);
 *   - boxes are rows (cx, cy, w, h, theta_radians) with `stride` floats
 *     between rows (5, or 6 for det rows carrying a score column — the
 *     reference hands such rows over at mmrotate/core/evaluation/eval_map.py:60-62).
 */
#ifndef ROTGEO_H_
#define ROTGEO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ROTGEO_OK 0
#define ROTGEO_ERR_ARG (-1)        /* bad argument (null pointer, negative size, bad mode) */
#define ROTGEO_ERR_WORKSPACE (-2)  /* workspace too small / misaligned */
#define ROTGEO_ERR_CUDA (-3)       /* a CUDA runtime call or launch failed */
#define ROTGEO_ERR_RANGE (-4)      /* NMS front-end: ids of a group span >= 2^16 values / too many segments */

#define ROTGEO_MODE_IOU 0
#define ROTGEO_MODE_IOF 1          /* intersection / area(first box) */

/* ABI version (major*1000 + minor). */
int rotgeo_version(void);

/* Thread-local text of the last error on this thread ("" if none). */
const char *rotgeo_last_error(void);

/* ---- packed box layout -----------------------------------------------------
 * The kernels read boxes from a packed set of n_pad slots (n rounded up to a
 * multiple of 128), four planes back to back, 96 bytes per padded box:
 *   plane0  float4 (cx, cy, r, area)      r = circumscribed radius, area = w*h
 *   plane1  float4 (cos, sin, w/2, h/2)
 *   planeR  uint32  the box's occupancy as cyclic bin ranges over a modular grid
 *                   of 32 bins per axis (x, y; NMS packs: also log-area)
 *   planeT  1920 uint32 per block of 128 boxes: the block's inverted bin index
 *                   (which of its boxes occupy a run of bins), the column side of
 *                   the pair cull
 * (orientedobjectdetection_b200/csrc/rotgeo_device.cuh).  rotgeo_packed_bytes(n)
 * is the size of one packed set; the buffer must be 16-byte aligned. */
size_t rotgeo_packed_bytes(int64_t n);

/* AoS rows -> packed SoA.  `gather` (nullable, int64[n]) selects rows:
 * packed[i] = boxes[gather[i]].  `min_wh` > 0 clamps w,h from below first
 * (the reference clamps to 1e-3 at
 * mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:84-87);
 * pass 0 for no clamp.  `offsets` (nullable, float[n], indexed like the
 * SOURCE rows) is added to cx and cy before packing — the class/level offset
 * trick of mmrotate/core/post_processing/bbox_nms_rotated.py:73-77.
 * `cull_cell` is the bin width of the cull masks: two packed sets that are
 * compared with each other MUST be packed with the same positive value
 * (a good one is the extent of the box centres / 30); <= 0 derives it from this
 * set alone (fine for a set only ever compared with itself).  The value
 * affects speed only, never results. */
int rotgeo_pack_boxes(const float *boxes, int64_t n, int stride, const int64_t *gather,
                      const float *offsets, float min_wh, float cull_cell, void *packed,
                      void *stream);

/* ---- rotated IoU -----------------------------------------------------------
 * Replaces ext_module.box_iou_rotated (mmcv.ops.box_iou_rotated), bound by the
 * reference at mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:2,89,
 * mmrotate/core/evaluation/eval_map.py:6,60 and
 * mmrotate/datasets/pipelines/transforms.py:8,369.
 *
 * out[i * ld_out + j] = IoU (mode 0) or IoF (mode 1) of boxes1[i], boxes2[j].
 * Workspace: rotgeo_iou_workspace_bytes(n1, n2), 16-byte aligned. */
size_t rotgeo_iou_workspace_bytes(int64_t n1, int64_t n2);
int rotgeo_iou_pairwise(const float *boxes1, int64_t n1, int stride1, const float *boxes2,
                        int64_t n2, int stride2, int mode, float min_wh, float *out,
                        int64_t ld_out, void *workspace, size_t workspace_bytes, void *stream);

/* Same on already-packed sets (no workspace; lets a caller keep anchors or
 * ground truth packed across calls).  Both sets must have been packed with the
 * same positive `cull_cell`. */
int rotgeo_iou_pairwise_packed(const void *packed1, int64_t n1, const void *packed2, int64_t n2,
                               int mode, float *out, int64_t ld_out, void *stream);

/* ---- rotated IoU without any allocation by anybody -----------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same result as rotgeo_iou_pairwise (the same two kernels); the packed box
 * sets live in a file-scope __device__ array of ROTGEO_STATIC_SCRATCH_BYTES inside the
 * library instead of a caller-owned workspace, so the call takes inputs and the output only
 * and calls no cudaMalloc / cudaFree / allocating routine.  Every GPU that loads the library
 * holds the array until exit.
 *   limit   padded_count(n1) + padded_count(n2) <= 16384 boxes (each count rounded up to a
 *           multiple of 128; 8192 x 8192 fits); beyond it ROTGEO_ERR_WORKSPACE with text.
 *   state   all calls share the array and it keeps the last call's contents; nothing in it
 *           has to start from zero (every slot read is rewritten first).  Calls that OVERLAP
 *           on the GPU (different streams, unordered graph branches) race; calls on one
 *           stream are ordered and safe. */
#define ROTGEO_STATIC_SCRATCH_BYTES (2 * 8192 * 96)
int rotgeo_iou_pairwise_static(const float *boxes1, int64_t n1, int stride1, const float *boxes2,
                               int64_t n2, int stride2, int mode, float min_wh, float *out,
                               int64_t ld_out, void *stream);

/* aligned=True variant: out[i] = IoU(boxes1[i], boxes2[i]). */
int rotgeo_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1,
                       int stride2, int mode, float min_wh, float *out, void *stream);

/* ---- rotated NMS -----------------------------------------------------------
 * Replaces ext_module.nms_rotated (mmcv.ops.nms_rotated), bound by the
 * reference at mmrotate/core/post_processing/bbox_nms_rotated.py:3,80,122,
 * mmrotate/core/patch/merge_results.py:4,116, mmrotate/datasets/dota.py:16,379,
 * and reached through mmcv.ops.batched_nms at
 * mmrotate/models/dense_heads/oriented_rpn_head.py:275 and rotated_rpn_head.py:515.
 *
 * Greedy suppression with the reference's CPU semantics: visiting boxes in
 * `order`, a kept box i suppresses every later j of the same segment with
 * IoU(i, j) >= iou_threshold.
 *
 *   boxes / stride    AoS rows as above (n rows)
 *   offsets           nullable float[n]: added to cx, cy (class-offset trick)
 *   order             int64[n]: indices into boxes, score-descending over ALL boxes;
 *                     nullable when seg_order is given: kept indices then come out
 *                     in seg_order order (segment-major) and the caller re-orders
 *   seg_order         nullable int64[n]: the same indices sorted segment-major,
 *                     score-descending inside each segment (null = one segment,
 *                     `order` is used)
 *   seg_offsets_host  HOST int64[num_segments + 1]: segment boundaries in
 *                     seg_order (null when seg_order is null)
 *   keep_out          int64[n]: kept indices into boxes, in `order` order (see above)
 *   num_keep_out      int64[1] (device): how many were kept
 *
 * Workspace: rotgeo_nms_workspace_bytes(n, seg_offsets_host, num_segments). */
size_t rotgeo_nms_workspace_bytes(int64_t n, const int64_t *seg_offsets_host, int64_t num_segments);
int rotgeo_nms(const float *boxes, int64_t n, int stride, const float *offsets,
               const int64_t *order, const int64_t *seg_order, const int64_t *seg_offsets_host,
               int64_t num_segments, float iou_threshold, int64_t *keep_out,
               int64_t *num_keep_out, void *workspace, size_t workspace_bytes, void *stream);

/* ---- fused assigner reductions ------------------------------------------------
 * What every caller of the reference's IoU calculator does with the matrix, without the
 * matrix.  boxes1 = the rows (ground truth in the assigners),
 * boxes2 = the columns (anchors / proposals); mode, min_wh as in rotgeo_iou_pairwise.
 *
 * rotgeo_iou_rowcol_max: row_max[i], row_argmax[i] = max / first argmax over the columns of
 *   overlap(boxes1[i], boxes2[:]) and col_max[j], col_argmax[j] likewise over the rows — the
 *   pair of `overlaps.max(dim=1)` / `overlaps.max(dim=0)` every assigner starts with
 *   (mmdet MaxIoUAssigner.assign_wrt_overlaps; mmrotate/core/bbox/assigners/
 *   atss_obb_assigner.py:118-145; obb_label_assigner.py:128-143 per gt).  Any output may be
 *   null.  A line without a positive value reports 0 at index 0.
 *
 * rotgeo_max_iou_assign: the whole of MaxIoUAssigner.assign_wrt_overlaps after
 *   `overlaps = self.iou_calculator(gt_bboxes, bboxes)` (reference call:
 *   rotate_iou2d_calculator.py:44 <- mmdet max_iou_assigner.py): per box j
 *     -1, then 0 where neg_iou_lo <= max_j < neg_iou_hi (a float neg_iou_thr t is the
 *     interval [0, t)), then argmax_j + 1 where max_j >= pos_iou_thr; with
 *     match_low_quality, every gt i with gt_max_i >= min_pos_iou claims (ascending i, later
 *     wins) box gt_argmax_i, or with gt_max_assign_all every box whose overlap with i equals
 *     gt_max_i — including, as in mmdet, all boxes when gt_max_i == 0 >= min_pos_iou.
 *   assigned_gt_inds (num_boxes,) int64, max_overlaps (num_boxes,) float (nullable).
 *   gt_bboxes_ignore / ignore_iof_thr / gpu_assign_thr are not part of this entry.
 *
 * workspace: rotgeo_iou_max_workspace_bytes(n1, n2) bytes, 16-byte aligned. */
size_t rotgeo_iou_max_workspace_bytes(int64_t n1, int64_t n2);
int rotgeo_iou_rowcol_max(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2,
                          int stride2, int mode, float min_wh, float *row_max, int64_t *row_argmax,
                          float *col_max, int64_t *col_argmax, void *workspace, size_t workspace_bytes,
                          void *stream);
int rotgeo_max_iou_assign(const float *gt_boxes, int64_t num_gts, int stride_g, const float *boxes,
                          int64_t num_boxes, int stride_b, int mode, float min_wh, float pos_iou_thr,
                          float neg_iou_lo, float neg_iou_hi, float min_pos_iou, int match_low_quality,
                          int gt_max_assign_all, int64_t *assigned_gt_inds, float *max_overlaps,
                          void *workspace, size_t workspace_bytes, void *stream);

/* ---- mAP evaluation: true / false positives -------------------------------------
 * The per-image body of the reference's tpfp_default (mmrotate/core/evaluation/
 * eval_map.py:12-94) for ALL (class, image) pairs of eval_rbbox_map (eval_map.py:126-246)
 * in one launch.  Detections (rows of det_stride >= 6 floats:
 * box + score) and ground truths (rows of gt_stride >= 5 floats) are grouped by segment;
 * segment s owns dets [det_seg_offsets[s], det_seg_offsets[s+1]) and gts likewise (device
 * arrays of num_segments + 1 int64).  Inside a segment the ignored ground truths come after
 * the others, flagged in gt_ignore (nullable).  Per detection, as in the reference: IoU
 * (raw box_iou_rotated, no clamp) against the segment's gts, max and first argmax; then in
 * descending score order (ties: lower index first) tp = 1 if max >= iou_thr and the argmax
 * gt is neither ignored nor claimed yet, fp = 1 if it was claimed already or max < iou_thr,
 * both 0 if the argmax gt is an ignored one; a segment without gts is all fp.
 * tp_out, fp_out: (num_dets,) bytes; iou_max_out (num_dets,) float, nullable. */
size_t rotgeo_eval_tpfp_workspace_bytes(int64_t num_dets, int64_t num_gts);
int rotgeo_eval_tpfp(const float *dets, int64_t num_dets, int det_stride, const int64_t *det_seg_offsets,
                     const float *gts, int64_t num_gts, int gt_stride, const int64_t *gt_seg_offsets,
                     const uint8_t *gt_ignore, int64_t num_segments, float iou_thr, uint8_t *tp_out,
                     uint8_t *fp_out, float *iou_max_out, void *workspace, size_t workspace_bytes,
                     void *stream);

/* ---- differentiable aligned IoU ------------------------------------------------
 * Replaces mmcv.ops.diff_iou_rotated_2d as the reference's RotatedIoULoss calls it
 * (mmrotate/models/losses/rotated_iou_loss.py:47):
 * iou_out[i] = IoU(boxes1[i], boxes2[i]) and, when the pointers are not null,
 * grad1_out[i, :] = d iou_out[i] / d (cx, cy, w, h, theta) of boxes1[i] and grad2_out
 * likewise for boxes2[i] ((n, 5) contiguous).  The gradient is analytic (boundary
 * integrals over the clipped edges), not autograd through a vertex list; it is 0 where the
 * boxes do not intersect and where an area is below 1e-14. */
int rotgeo_diff_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1, int stride2,
                            float *iou_out, float *grad1_out, float *grad2_out, void *stream);

/* ---- rotated NMS front-end: one sort kernel, device-side plan -------------------
 * What the reference does in python/ATen around the native op, for one call or a whole
 * batch of calls at once:
 *   - mmcv/ops/nms.py nms_rotated: scores.sort(0, descending=True) + index_select;
 *   - mmcv/ops/nms.py batched_nms: boxes.max(), the id * (max + 1) offsets, the add;
 *   - the per-image python loop of the RPN heads around batched_nms
 *     (mmrotate/models/dense_heads/rotated_rpn_head.py:414-426,515, oriented_rpn_head.py:275);
 *   - the per-class loop of the patch merge (mmrotate/core/patch/merge_results.py:90-119).
 *
 * The n candidate rows are cut into GROUPS (contiguous row ranges, group_sizes_host[g] rows
 * each; null = one group): a group's kept boxes are emitted together in descending score
 * order (ties: lower row first), groups one after the other.  Inside a group, boxes of
 * equal seg_ids form a SEGMENT; suppression never crosses segments (seg_ids null = one
 * segment per group).  The ids of one group must span fewer than 65536 values and form at
 * most seg_cap segments.
 *
 *   rotgeo_nms_sort       ONE kernel (one thread-block cluster per group): score order,
 *                         segment-major order, with id_offsets != 0 the centre offsets
 *                         float(id) * (max(cx,cy) + max(w,h) + 1) of mmcv.ops.batched_nms
 *                         (maxima of the row's own group), and the segment table + slot /
 *                         tile / mask-word counts in a device header.  No synchronisation.
 *   rotgeo_nms_bounds     host-side UPPER BOUNDS of those counts from the group sizes and
 *                         seg_cap: sizes_out[6] = {slots, tiles, mask words, blocks of the
 *                         largest segment, segments, rows to emit}.  With them
 *                         rotgeo_nms_run needs no device->host read at all.
 *   rotgeo_nms_read_plan  the exact counts instead: copies the header back and
 *                         SYNCHRONISES `stream` (for jobs whose worst-case workspace is not
 *                         worth allocating).  ROTGEO_ERR_RANGE when the ids of a group
 *                         span 65536 values or more, or a group has more than seg_cap
 *                         segments (callers renumber / split).
 *   rotgeo_nms_run        pack + mask + resolve over `sizes` (bounds or exact).  Centre
 *                         offsets: `offsets` (n,) float added to cx, cy (nullable), or,
 *                         with sort_offsets != 0, the ones rotgeo_nms_sort computed.
 *                         keep_out (n,): kept rows, group after group, score-descending;
 *                         num_keep_out int64[2] (device): [0] how many, [1] the front-end's
 *                         error bits (non-zero = the result is void: same cases as
 *                         rotgeo_nms_read_plan's ROTGEO_ERR_RANGE).  group_counts_out
 *                         (nullable, (num_groups,) int64 device): when given, keep_out
 *                         holds rows INSIDE their group and group_counts_out the kept
 *                         boxes per group.  iou_threshold must be positive.
 *
 * sortws: rotgeo_nms_sort_bytes(...) bytes, 16-byte aligned, untouched between the calls;
 * workspace: rotgeo_nms_run_bytes(n, sizes).  group_sizes_host, num_groups and seg_cap must
 * be the same in all calls of one job. */
/*   rotgeo_nms_grouped    rotgeo_nms_sort + rotgeo_nms_bounds + rotgeo_nms_run in ONE call over worst-case
 *                         bounds (no device->host read), for hosts where every boundary crossing counts
 *                         (a python caller pays ~10 us per ctypes call).  tail_out int64[2 + num_groups]
 *                         (device): kept in all, front-end error bits, and with want_group_counts != 0
 *                         the kept boxes per group (keep_out then holds rows inside their group).
 *                         dets_out (nullable, float (n, 6)): the rows cat(boxes[keep], scores[keep]) that
 *                         mmcv's nms_rotated / batched_nms return, written by the same kernels.
 *                         need_out[2] (HOST) always receives the sizes of sortws and workspace this job
 *                         needs; ROTGEO_ERR_WORKSPACE (nothing launched) when the given ones are smaller,
 *                         so a caller keeps two growing buffers and retries once. */
int rotgeo_nms_grouped(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *seg_ids,
                       int id_offsets, const float *offsets, const int64_t *group_sizes_host, int64_t num_groups,
                       int64_t seg_cap, float iou_threshold, int64_t *keep_out, int64_t *tail_out,
                       int want_group_counts, float *dets_out, void *sortws, size_t sortws_bytes, void *workspace,
                       size_t workspace_bytes, size_t *need_out, void *stream);
size_t rotgeo_nms_sort_bytes(int64_t n, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap);
int rotgeo_nms_sort(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *seg_ids,
                    int id_offsets, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap,
                    void *sortws, size_t sortws_bytes, void *stream);
int rotgeo_nms_bounds(int64_t n, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap,
                      int64_t *sizes_out);
int rotgeo_nms_read_plan(const void *sortws, int64_t n, const int64_t *group_sizes_host, int64_t num_groups,
                         int64_t seg_cap, int64_t *sizes_out, void *stream);
size_t rotgeo_nms_run_bytes(int64_t n, const int64_t *sizes);
int rotgeo_nms_run(const float *boxes, int64_t n, int stride, const float *offsets, int sort_offsets,
                   const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap, const void *sortws,
                   const int64_t *sizes, float iou_threshold, int64_t *keep_out, int64_t *num_keep_out,
                   int64_t *group_counts_out, void *workspace, size_t workspace_bytes, void *stream);

/* ---- multi-GPU: results exchanged over peer memory (NVLink / NVSwitch) ---------------
 * The reference has no multi-GPU form of this path: it loops over the images of a batch and
 * over the classes of a patch merge in python inside one rank
 * (mmrotate/models/dense_heads/rotated_rpn_head.py:414-426,515,
 * mmrotate/core/patch/merge_results.py:90-119).  These entry points are what one rank of a
 * one-process-per-GPU job calls: it computes ITS shard and stores the result straight into
 * every peer's memory; no library collective sits in the data path.
 *
 * Window: one symmetric device buffer per rank, every rank's buffer mapped into every
 * process (the host layer allocates and exchanges them, e.g. torch symmetric memory);
 * peer_bases_host[r] = address of rank r's window IN THIS PROCESS (HOST array of `world`
 * pointers, world <= 16).  The first rotgeo_peer_pad_bytes() of a window are its signal pad
 * and must be zero before the first call; offsets passed below start behind it.
 * epoch: a counter the caller advances monotonically, identically on all ranks; every call
 * states how many values it consumes.  All ranks must issue the same calls in the same order.
 *
 *   rotgeo_peer_barrier           all ranks' earlier writes (remote ones included) are
 *                                 visible to all ranks behind it on `stream`.  1 epoch.
 *   rotgeo_dist_iou_allgather     the (n1_total, n2) matrix of rotgeo_iou_pairwise, rows
 *                                 [row_lo, row_lo + n1_local) computed here, complete in EVERY
 *                                 window at matrix_offset afterwards.  Zeros never travel:
 *                                 each rank zero-fills the foreign rows locally and pushes
 *                                 only the non-zero 16-byte groups of its block.  2 epochs.
 *                                 workspace: rotgeo_iou_workspace_bytes(n1_local, n2).
 *   rotgeo_dist_iou_rowcol_max    rotgeo_iou_rowcol_max of the row-sharded matrix: all four
 *                                 outputs complete on every rank (row_* (n1_total,), col_*
 *                                 (n2,), local buffers, any may be null).  Stage in the
 *                                 window: rotgeo_dist_rowcol_stage_bytes(...) at stage_offset.
 *                                 2 epochs.  workspace: rotgeo_iou_max_workspace_bytes(n1_local, n2).
 *   rotgeo_dist_push_keeps        the kept rows of this rank's NMS groups (keep / tail as
 *                                 rotgeo_nms_run leaves them: tail = num_keep_out[0..1] followed
 *                                 by group_counts_out) into every window at out_offset:
 *                                   int64 counts[num_groups_total] | int64 errors[world] |
 *                                   int64 keep[num_groups_total][cap] |
 *                                   float rows[num_groups_total][cap][row_floats]   (optional)
 *                                 group_ids_host[g] = global id of local group g (at most 64
 *                                 local groups); lists are cut to cap; rows (nullable) are
 *                                 gathered from rows_src at group_base_host[g] + kept index.
 *                                 Region size: rotgeo_dist_keeps_bytes(...).  1 epoch. */
size_t rotgeo_peer_pad_bytes(void);
int rotgeo_peer_barrier(void *const *peer_bases_host, int world, int rank, uint64_t epoch, void *stream);
int rotgeo_dist_iou_allgather(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t matrix_offset,
                              const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                              int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                              void *workspace, size_t workspace_bytes, void *stream);
size_t rotgeo_dist_rowcol_stage_bytes(int64_t n1_total, int64_t n2, int world);
int rotgeo_dist_iou_rowcol_max(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t stage_offset,
                               const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                               int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                               float *row_max, int64_t *row_argmax, float *col_max, int64_t *col_argmax,
                               void *workspace, size_t workspace_bytes, void *stream);
size_t rotgeo_dist_keeps_bytes(int64_t num_groups_total, int64_t cap, int world, int row_floats);
int rotgeo_dist_push_keeps(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t out_offset,
                           const int64_t *keep, const int64_t *tail, int num_local, const int64_t *group_ids_host,
                           int64_t cap, int64_t num_groups_total, const float *rows_src, int row_floats,
                           int row_stride, const int64_t *group_base_host, void *stream);

/* ---- class-batched rotated NMS, fused ----------------------------------------
 * Replaces the whole body of the reference's multiclass_nms_rotated
 * (mmrotate/core/post_processing/bbox_nms_rotated.py:33-92, called from the eleven
 * get_bboxes sites listed in SURVEY.md §3): candidate filter (raw score >
 * score_thr, tested before the factors), class separation by the reference's own
 * centre offset  label * (max(cx,cy) + max(w,h) + 1)  in fp32, greedy suppression
 * per class (IoU >= iou_threshold), global score order, cut to max_num, and the
 * gather of the output rows — with no device->host read inside the call.
 *
 *   multi_bboxes   (rows, 5) when bbox_cols == 5 (one box per row, shared by all
 *                  classes) or (rows, 5 * num_classes) when bbox_cols == 5 * num_classes;
 *                  contiguous
 *   multi_scores   (rows, num_classes + 1) contiguous; the last column (background)
 *                  is ignored
 *   score_factors  nullable (rows,): multiplied into the scores after the filter
 *   cap            rows of the output buffers: max_num if > 0, else rows * num_classes
 *   dets_out       float (cap, 6): kept boxes + score, score-descending
 *   labels_out     int64 (cap,)
 *   keep_flat_out  int64 (cap,): row * num_classes + class of every kept candidate
 *   num_out        int64[1] (device): how many rows were written (<= cap)
 *
 * num_classes <= 64 and rows * num_classes < 2^26.
 * Workspace: rotgeo_multiclass_nms_workspace_bytes(rows, num_classes) — sized for the
 * worst case (every class keeps every row), because no count reaches the host. */
size_t rotgeo_multiclass_nms_workspace_bytes(int64_t rows, int num_classes);
int rotgeo_multiclass_nms(const float *multi_bboxes, int bbox_cols, const float *multi_scores,
                          const float *score_factors, int64_t rows, int num_classes,
                          float score_thr, float iou_threshold, int64_t cap, float *dets_out,
                          int64_t *labels_out, int64_t *keep_flat_out, int64_t *num_out,
                          void *workspace, size_t workspace_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ROTGEO_H_ */
