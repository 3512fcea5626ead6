// extern "C" boundary of librotgeo.so (see include/rotgeo.h).  Argument
// checking, workspace carving and launches only; no allocation, no sync.
#include <stdarg.h>
#include <stdio.h>

#include <vector>

#include "../../include/rotgeo.h"
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace {

thread_local char g_err[256] = "";

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(const char *what) {
    cudaError_t e = cudaGetLastError();
    return fail(ROTGEO_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

namespace rotgeo {
// the same thread-local error text for entry points that live in other translation units
int abi_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace rotgeo

extern "C" {

int rotgeo_version(void) { return 2000; }   // 2000: rotgeo_nms_sort / _bounds / _read_plan / _run replace the prepare family

const char *rotgeo_last_error(void) { return g_err; }

size_t rotgeo_packed_bytes(int64_t n) {
    if (n < 0) return 0;
    return (size_t)rotgeo::padded_count(n) * rotgeo::kPackedBytesPerBox;
}

int rotgeo_pack_boxes(const float *boxes, int64_t n, int stride, const int64_t *gather,
                      const float *offsets, float min_wh, float cull_cell, void *packed,
                      void *stream) {
    if (n < 0 || stride < 5) return fail(ROTGEO_ERR_ARG, "pack: n=%lld stride=%d", (long long)n, stride);
    if (n == 0) return ROTGEO_OK;
    if (!boxes || !packed) return fail(ROTGEO_ERR_ARG, "pack: null pointer");
    if (!aligned16(packed)) return fail(ROTGEO_ERR_ARG, "pack: packed buffer must be 16-byte aligned");
    const rotgeo::PackSet s1 = rotgeo::make_pack_set(boxes, n, stride, gather, offsets, packed,
                                                     rotgeo::padded_count(n));
    const rotgeo::PackSet none = rotgeo::make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
    if (rotgeo::launch_pack(s1, none, min_wh, cull_cell, 0.f, (cudaStream_t)stream))
        return cuda_fail("pack launch");
    return ROTGEO_OK;
}

size_t rotgeo_iou_workspace_bytes(int64_t n1, int64_t n2) {
    return align_up(rotgeo_packed_bytes(n1), 256) + align_up(rotgeo_packed_bytes(n2), 256);
}

int rotgeo_iou_pairwise_packed(const void *packed1, int64_t n1, const void *packed2, int64_t n2,
                               int mode, float *out, int64_t ld_out, void *stream) {
    if (n1 < 0 || n2 < 0 || (mode != 0 && mode != 1) || ld_out < n2)
        return fail(ROTGEO_ERR_ARG, "iou: n1=%lld n2=%lld mode=%d ld_out=%lld", (long long)n1,
                    (long long)n2, mode, (long long)ld_out);
    if (n1 == 0 || n2 == 0) return ROTGEO_OK;
    if (!packed1 || !packed2 || !out) return fail(ROTGEO_ERR_ARG, "iou: null pointer");
    if (!aligned16(packed1) || !aligned16(packed2))
        return fail(ROTGEO_ERR_ARG, "iou: packed buffers must be 16-byte aligned");
    const int rc = rotgeo::launch_iou_pairwise(packed1, n1, packed2, n2, mode, out, ld_out, false,
                                               (cudaStream_t)stream);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "iou: too many tiles");
    if (rc) return cuda_fail("iou launch");
    return ROTGEO_OK;
}

int rotgeo_iou_pairwise(const float *boxes1, int64_t n1, int stride1, const float *boxes2,
                        int64_t n2, int stride2, int mode, float min_wh, float *out,
                        int64_t ld_out, void *workspace, size_t workspace_bytes, void *stream) {
    if (n1 < 0 || n2 < 0 || stride1 < 5 || stride2 < 5)
        return fail(ROTGEO_ERR_ARG, "iou: n1=%lld n2=%lld stride1=%d stride2=%d", (long long)n1,
                    (long long)n2, stride1, stride2);
    if (n1 == 0 || n2 == 0) return ROTGEO_OK;
    if (!boxes1 || !boxes2) return fail(ROTGEO_ERR_ARG, "iou: null pointer");
    if (!workspace || !aligned16(workspace) || workspace_bytes < rotgeo_iou_workspace_bytes(n1, n2))
        return fail(ROTGEO_ERR_WORKSPACE, "iou: workspace %zu < %zu bytes or misaligned",
                    workspace_bytes, rotgeo_iou_workspace_bytes(n1, n2));
    char *ws = reinterpret_cast<char *>(workspace);
    void *pk1 = ws;
    void *pk2 = ws + align_up(rotgeo_packed_bytes(n1), 256);
    if (mode != 0 && mode != 1) return fail(ROTGEO_ERR_ARG, "iou: mode=%d", mode);
    if (!out || ld_out < n2) return fail(ROTGEO_ERR_ARG, "iou: out=%p ld_out=%lld", (void *)out, (long long)ld_out);
    // both sets in ONE launch, so they share the automatically chosen bin width
    const rotgeo::PackSet s1 = rotgeo::make_pack_set(boxes1, n1, stride1, nullptr, nullptr, pk1,
                                                     rotgeo::padded_count(n1));
    const rotgeo::PackSet s2 = rotgeo::make_pack_set(boxes2, n2, stride2, nullptr, nullptr, pk2,
                                                     rotgeo::padded_count(n2));
    if (rotgeo::launch_pack(s1, s2, min_wh, 0.f, 0.f, (cudaStream_t)stream)) return cuda_fail("pack launch");
    // the tile kernel starts streaming zeros while the pack kernel is still running
    const int rc = rotgeo::launch_iou_pairwise(pk1, n1, pk2, n2, mode, out, ld_out, true,
                                               (cudaStream_t)stream);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "iou: too many tiles");
    if (rc) return cuda_fail("iou launch");
    return ROTGEO_OK;
}

int rotgeo_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1,
                       int stride2, int mode, float min_wh, float *out, void *stream) {
    if (n < 0 || stride1 < 5 || stride2 < 5 || (mode != 0 && mode != 1))
        return fail(ROTGEO_ERR_ARG, "iou_aligned: n=%lld mode=%d", (long long)n, mode);
    if (n == 0) return ROTGEO_OK;
    if (!boxes1 || !boxes2 || !out) return fail(ROTGEO_ERR_ARG, "iou_aligned: null pointer");
    if (rotgeo::launch_iou_aligned(boxes1, boxes2, n, stride1, stride2, mode, min_wh, out,
                                   (cudaStream_t)stream))
        return cuda_fail("iou_aligned launch");
    return ROTGEO_OK;
}

size_t rotgeo_iou_max_workspace_bytes(int64_t n1, int64_t n2) {
    return rotgeo::iou_max_workspace_bytes(n1, n2);
}

static int fill_async(void *dst, int value, size_t bytes, cudaStream_t st) {
    return bytes == 0 || cudaMemsetAsync(dst, value, bytes, st) == cudaSuccess ? 0 : -1;
}

int rotgeo_iou_rowcol_max(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2,
                          int stride2, int mode, float min_wh, float *row_max, int64_t *row_argmax,
                          float *col_max, int64_t *col_argmax, void *workspace, size_t workspace_bytes,
                          void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n1 < 0 || n2 < 0 || stride1 < 5 || stride2 < 5 || (mode != 0 && mode != 1))
        return fail(ROTGEO_ERR_ARG, "iou_rowcol_max: n1=%lld n2=%lld mode=%d", (long long)n1, (long long)n2, mode);
    if (n1 == 0 || n2 == 0) {
        // an empty line has maximum 0 at index 0 here (torch.max raises on an empty dimension)
        if ((n1 && row_max && fill_async(row_max, 0, (size_t)n1 * 4, st)) ||
            (n1 && row_argmax && fill_async(row_argmax, 0, (size_t)n1 * 8, st)) ||
            (n2 && col_max && fill_async(col_max, 0, (size_t)n2 * 4, st)) ||
            (n2 && col_argmax && fill_async(col_argmax, 0, (size_t)n2 * 8, st)))
            return cuda_fail("iou_rowcol_max memset");
        return ROTGEO_OK;
    }
    if (!boxes1 || !boxes2) return fail(ROTGEO_ERR_ARG, "iou_rowcol_max: null pointer");
    const size_t need = rotgeo::iou_max_workspace_bytes(n1, n2);
    if (!workspace || !aligned16(workspace) || workspace_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "iou_rowcol_max: workspace %zu < %zu bytes or misaligned",
                    workspace_bytes, need);
    const int rc = rotgeo::launch_iou_rowcol_max(boxes1, n1, stride1, boxes2, n2, stride2, mode, min_wh,
                                                 row_max, row_argmax, col_max, col_argmax, workspace, st);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "iou_rowcol_max: too many tiles");
    if (rc) return cuda_fail("iou_rowcol_max launch");
    return ROTGEO_OK;
}

int rotgeo_max_iou_assign(const float *gt_boxes, int64_t num_gts, int stride_g, const float *boxes,
                          int64_t num_boxes, int stride_b, int mode, float min_wh, float pos_iou_thr,
                          float neg_iou_lo, float neg_iou_hi, float min_pos_iou, int match_low_quality,
                          int gt_max_assign_all, int64_t *assigned_gt_inds, float *max_overlaps,
                          void *workspace, size_t workspace_bytes, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (num_gts < 0 || num_boxes < 0 || stride_g < 5 || stride_b < 5 || (mode != 0 && mode != 1) ||
        num_gts >= 0x7fffffffLL)
        return fail(ROTGEO_ERR_ARG, "max_iou_assign: gts=%lld boxes=%lld mode=%d", (long long)num_gts,
                    (long long)num_boxes, mode);
    if (num_boxes == 0) return ROTGEO_OK;
    if (!assigned_gt_inds) return fail(ROTGEO_ERR_ARG, "max_iou_assign: null assigned_gt_inds");
    if (num_gts == 0) {   // no ground truth: everything is background (max_iou_assigner.py, empty branch)
        if (fill_async(assigned_gt_inds, 0, (size_t)num_boxes * 8, st) ||
            (max_overlaps && fill_async(max_overlaps, 0, (size_t)num_boxes * 4, st)))
            return cuda_fail("max_iou_assign memset");
        return ROTGEO_OK;
    }
    if (!gt_boxes || !boxes) return fail(ROTGEO_ERR_ARG, "max_iou_assign: null pointer");
    const size_t need = rotgeo::iou_max_workspace_bytes(num_gts, num_boxes);
    if (!workspace || !aligned16(workspace) || workspace_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "max_iou_assign: workspace %zu < %zu bytes or misaligned",
                    workspace_bytes, need);
    const int rc = rotgeo::launch_max_iou_assign(gt_boxes, num_gts, stride_g, boxes, num_boxes, stride_b, mode,
                                                 min_wh, pos_iou_thr, neg_iou_lo, neg_iou_hi, min_pos_iou,
                                                 match_low_quality, gt_max_assign_all, assigned_gt_inds,
                                                 max_overlaps, workspace, st);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "max_iou_assign: too many tiles");
    if (rc) return cuda_fail("max_iou_assign launch");
    return ROTGEO_OK;
}

size_t rotgeo_eval_tpfp_workspace_bytes(int64_t num_dets, int64_t num_gts) {
    return rotgeo::eval_tpfp_workspace_bytes(num_dets, num_gts);
}

int rotgeo_eval_tpfp(const float *dets, int64_t num_dets, int det_stride, const int64_t *det_seg_offsets,
                     const float *gts, int64_t num_gts, int gt_stride, const int64_t *gt_seg_offsets,
                     const uint8_t *gt_ignore, int64_t num_segments, float iou_thr, uint8_t *tp_out,
                     uint8_t *fp_out, float *iou_max_out, void *workspace, size_t workspace_bytes,
                     void *stream) {
    if (num_dets < 0 || num_gts < 0 || num_segments < 0 || det_stride < 6 || gt_stride < 5 ||
        num_dets > 0x7fffffffLL || num_gts > 0x7fffffffLL)
        return fail(ROTGEO_ERR_ARG, "eval_tpfp: dets=%lld gts=%lld segments=%lld det_stride=%d gt_stride=%d",
                    (long long)num_dets, (long long)num_gts, (long long)num_segments, det_stride, gt_stride);
    if (num_dets == 0 || num_segments == 0) return ROTGEO_OK;
    if (!dets || !det_seg_offsets || !gt_seg_offsets || !tp_out || !fp_out || (num_gts > 0 && !gts))
        return fail(ROTGEO_ERR_ARG, "eval_tpfp: null pointer");
    const size_t need = rotgeo::eval_tpfp_workspace_bytes(num_dets, num_gts);
    if (!workspace || !aligned16(workspace) || workspace_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "eval_tpfp: workspace %zu < %zu bytes or misaligned", workspace_bytes, need);
    if (rotgeo::launch_eval_tpfp(dets, num_dets, det_stride, det_seg_offsets, gts, num_gts, gt_stride,
                                 gt_seg_offsets, gt_ignore, num_segments, iou_thr, tp_out, fp_out, iou_max_out,
                                 workspace, (cudaStream_t)stream))
        return cuda_fail("eval_tpfp launch");
    return ROTGEO_OK;
}

int rotgeo_diff_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1, int stride2,
                            float *iou_out, float *grad1_out, float *grad2_out, void *stream) {
    if (n < 0 || stride1 < 5 || stride2 < 5)
        return fail(ROTGEO_ERR_ARG, "diff_iou_aligned: n=%lld stride1=%d stride2=%d", (long long)n, stride1, stride2);
    if (n == 0) return ROTGEO_OK;
    if (!boxes1 || !boxes2 || !iou_out) return fail(ROTGEO_ERR_ARG, "diff_iou_aligned: null pointer");
    if (rotgeo::launch_diff_iou(boxes1, boxes2, n, stride1, stride2, iou_out, grad1_out, grad2_out,
                                (cudaStream_t)stream))
        return cuda_fail("diff_iou_aligned launch");
    return ROTGEO_OK;
}

size_t rotgeo_nms_workspace_bytes(int64_t n, const int64_t *seg_offsets_host, int64_t num_segments) {
    if (n <= 0) return 0;
    rotgeo::NmsPlan plan;
    rotgeo::nms_plan(n, seg_offsets_host, num_segments, &plan, nullptr);
    return plan.total_bytes;
}

int rotgeo_nms(const float *boxes, int64_t n, int stride, const float *offsets,
               const int64_t *order, const int64_t *seg_order, const int64_t *seg_offsets_host,
               int64_t num_segments, float iou_threshold, int64_t *keep_out,
               int64_t *num_keep_out, void *workspace, size_t workspace_bytes, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || stride < 5) return fail(ROTGEO_ERR_ARG, "nms: n=%lld stride=%d", (long long)n, stride);
    if (!num_keep_out) return fail(ROTGEO_ERR_ARG, "nms: null num_keep_out");
    if (n == 0) {
        if (cudaMemsetAsync(num_keep_out, 0, sizeof(int64_t), st) != cudaSuccess)
            return cuda_fail("nms memset");
        return ROTGEO_OK;
    }
    if (!boxes || !keep_out || (!order && !seg_order)) return fail(ROTGEO_ERR_ARG, "nms: null pointer");
    if (!(iou_threshold > 0.f)) return fail(ROTGEO_ERR_ARG, "nms: iou_threshold must be positive");
    const bool segmented = seg_order != nullptr;
    if (segmented) {
        if (!seg_offsets_host || num_segments < 1)
            return fail(ROTGEO_ERR_ARG, "nms: seg_order given without seg_offsets_host");
        if (seg_offsets_host[0] != 0 || seg_offsets_host[num_segments] != n)
            return fail(ROTGEO_ERR_ARG, "nms: seg_offsets_host must run from 0 to n");
        for (int64_t s = 0; s < num_segments; ++s)
            if (seg_offsets_host[s + 1] < seg_offsets_host[s])
                return fail(ROTGEO_ERR_ARG, "nms: seg_offsets_host not monotone");
    } else {
        seg_offsets_host = nullptr;
        num_segments = 1;
    }
    rotgeo::NmsPlan plan;
    std::vector<rotgeo::NmsSegment> segs((size_t)num_segments);
    rotgeo::nms_plan(n, seg_offsets_host, num_segments, &plan, segs.data());
    if (!workspace || !aligned16(workspace) || workspace_bytes < plan.total_bytes)
        return fail(ROTGEO_ERR_WORKSPACE, "nms: workspace %zu < %zu bytes or misaligned",
                    workspace_bytes, plan.total_bytes);
    for (const auto &g : segs)
        if (g.words / 2 >= 65536)     // the mask kernel carries the column block in 16 bits
            return fail(ROTGEO_ERR_ARG, "nms: a segment of more than %d boxes is not supported", 65535 * 128);
    char *ws = reinterpret_cast<char *>(workspace);
    void *packed = ws + plan.off_packed;
    rotgeo::NmsSegment *segs_dev = reinterpret_cast<rotgeo::NmsSegment *>(ws + plan.off_segments);
    uint64_t *mask = reinterpret_cast<uint64_t *>(ws + plan.off_mask);
    uint8_t *flags = reinterpret_cast<uint8_t *>(ws + plan.off_flags);
    rotgeo::NmsScratch scratch;
    scratch.kr = reinterpret_cast<unsigned long long *>(ws + plan.off_kr);
    scratch.summary = reinterpret_cast<unsigned long long *>(ws + plan.off_summary);
    scratch.sum_words = plan.sum_words;
    scratch.sync = reinterpret_cast<unsigned int *>(ws + plan.off_sync);
    scratch.block_counts = reinterpret_cast<int *>(ws + plan.off_counts);
    // pageable source: the runtime stages the bytes before returning, so `segs`
    // may die at scope exit
    if (cudaMemcpyAsync(segs_dev, segs.data(), segs.size() * sizeof(rotgeo::NmsSegment),
                        cudaMemcpyHostToDevice, st) != cudaSuccess)
        return cuda_fail("nms segment upload");
    const int64_t *sorder = segmented ? seg_order : order;
    {
        // slots follow the segment table: every segment starts on a 128-slot boundary
        rotgeo::PackSet s1 = rotgeo::make_pack_set(boxes, n, stride, sorder, offsets, packed, plan.n_pad);
        s1.segs = segs_dev;
        s1.nseg = (int)num_segments;
        s1.z_kr = scratch.kr;
        s1.z_summary = scratch.summary;
        s1.z_sum_words = scratch.sum_words;
        s1.z_sync = scratch.sync;
        const rotgeo::PackSet none = rotgeo::make_pack_set(nullptr, 0, 5, nullptr, nullptr, nullptr, 0);
        if (rotgeo::launch_pack(s1, none, 0.f, 0.f, rotgeo::nms_area_axis_ratio(iou_threshold), st)) return cuda_fail("nms pack launch");
    }
    // no global order given: kept indices come out segment-major (the caller re-orders)
    if (!order) order = sorder;
    const int rc = rotgeo::launch_nms(plan, packed, segs_dev, order, sorder, nullptr, nullptr,
                                      iou_threshold, mask, flags, scratch, keep_out, num_keep_out, st);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "nms: too many tiles");
    if (rc) return cuda_fail("nms launch");
    return ROTGEO_OK;
}

static int64_t max_group_of(const int64_t *group_sizes_host, int64_t num_groups, int64_t n) {
    int64_t m = 0;
    for (int64_t g = 0; g < num_groups; ++g) {
        const int64_t v = group_sizes_host ? group_sizes_host[g] : n;
        m = v > m ? v : m;
    }
    return m;
}
static bool groups_ok(const int64_t *group_sizes_host, int64_t num_groups, int64_t n) {
    if (num_groups < 1 || num_groups > (1 << 20) || (!group_sizes_host && num_groups != 1)) return false;
    if (!group_sizes_host) return true;
    int64_t s = 0;
    for (int64_t g = 0; g < num_groups; ++g) {
        if (group_sizes_host[g] < 0) return false;
        s += group_sizes_host[g];
    }
    return s == n;
}
static rotgeo::NmsSizes sizes_from(const int64_t *v) {
    rotgeo::NmsSizes z;
    z.n_pad = v[0]; z.tiles = v[1]; z.mask_words = v[2]; z.max_blocks = v[3]; z.nseg = v[4]; z.n_emit = v[5];
    return z;
}
static void sizes_to(const rotgeo::NmsSizes &z, int64_t *v) {
    v[0] = z.n_pad; v[1] = z.tiles; v[2] = z.mask_words; v[3] = z.max_blocks; v[4] = z.nseg; v[5] = z.n_emit;
}

size_t rotgeo_nms_sort_bytes(int64_t n, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap) {
    if (n <= 0 || seg_cap < 1 || !groups_ok(group_sizes_host, num_groups, n)) return 0;
    return rotgeo::nms_sort_bytes(n, num_groups, seg_cap, max_group_of(group_sizes_host, num_groups, n));
}

int rotgeo_nms_sort(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *seg_ids,
                    int id_offsets, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap,
                    void *sortws, size_t sortws_bytes, void *stream) {
    if (n <= 0 || n > 0x7fffffffLL || stride < 5 || seg_cap < 1 || seg_cap > 65536)
        return fail(ROTGEO_ERR_ARG, "nms_sort: n=%lld stride=%d seg_cap=%lld", (long long)n, stride, (long long)seg_cap);
    if (!groups_ok(group_sizes_host, num_groups, n))
        return fail(ROTGEO_ERR_ARG, "nms_sort: group sizes must be non-negative and add up to n");
    if (!boxes || !scores || !sortws) return fail(ROTGEO_ERR_ARG, "nms_sort: null pointer");
    const size_t need = rotgeo_nms_sort_bytes(n, group_sizes_host, num_groups, seg_cap);
    if (!aligned16(sortws) || sortws_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "nms_sort: buffer %zu < %zu bytes or misaligned", sortws_bytes, need);
    const int rc = rotgeo::nms_sort(boxes, n, stride, scores, seg_ids, id_offsets, group_sizes_host, num_groups,
                                    seg_cap, sortws, (cudaStream_t)stream);
    if (rc == -6) return fail(ROTGEO_ERR_RANGE, "nms_sort: a group of more than sms * 8192 boxes");
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "nms_sort: bad group sizes");
    if (rc) return cuda_fail("nms_sort");
    return ROTGEO_OK;
}

int rotgeo_nms_bounds(int64_t n, const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap,
                      int64_t *sizes_out) {
    if (n <= 0 || seg_cap < 1 || !sizes_out || !groups_ok(group_sizes_host, num_groups, n))
        return fail(ROTGEO_ERR_ARG, "nms_bounds: bad argument");
    rotgeo::NmsSizes z;
    rotgeo::nms_bounds(group_sizes_host, num_groups, n, seg_cap, &z);
    sizes_to(z, sizes_out);
    return ROTGEO_OK;
}

int rotgeo_nms_read_plan(const void *sortws, int64_t n, const int64_t *group_sizes_host, int64_t num_groups,
                         int64_t seg_cap, int64_t *sizes_out, void *stream) {
    if (n <= 0 || seg_cap < 1 || !sortws || !sizes_out || !groups_ok(group_sizes_host, num_groups, n))
        return fail(ROTGEO_ERR_ARG, "nms_read_plan: bad argument");
    rotgeo::NmsSizes z;
    unsigned int err = 0;
    if (rotgeo::nms_read_plan(sortws, n, num_groups, seg_cap, max_group_of(group_sizes_host, num_groups, n), &z, &err,
                              (cudaStream_t)stream))
        return cuda_fail("nms_read_plan");
    sizes_to(z, sizes_out);
    if (err & rotgeo::kSortErrRange) return fail(ROTGEO_ERR_RANGE, "nms: segment ids of a group span 65536 values or more");
    if (err & rotgeo::kSortErrSegments) return fail(ROTGEO_ERR_RANGE, "nms: more segments than seg_cap in a group");
    return ROTGEO_OK;
}

size_t rotgeo_nms_run_bytes(int64_t n, const int64_t *sizes) {
    if (n <= 0 || !sizes) return 0;
    rotgeo::NmsRunLayout l;
    rotgeo::nms_run_layout(n, sizes_from(sizes), true, &l);     // with the regions of a spatially packed launch
    return l.total_bytes;
}

static int nms_run_checked(const float *boxes, int64_t n, int stride, const float *offsets, int sort_offsets,
                           const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap, const void *sortws,
                           const int64_t *sizes, float iou_threshold, int64_t *keep_out, int64_t *num_keep_out,
                           int64_t *group_counts_out, void *workspace, size_t workspace_bytes, void *stream,
                           const float *det_scores, float *dets_out) {
    if (n <= 0 || stride < 5 || seg_cap < 1 || !groups_ok(group_sizes_host, num_groups, n))
        return fail(ROTGEO_ERR_ARG, "nms_run: n=%lld stride=%d", (long long)n, stride);
    if (!(iou_threshold > 0.f))
        return fail(ROTGEO_ERR_ARG, "nms_run: iou_threshold must be positive (<= 0 keeps only the best box: "
                                    "the caller answers that case, the pair culls never form disjoint pairs)");
    if (!boxes || !sortws || !sizes || !keep_out || !num_keep_out) return fail(ROTGEO_ERR_ARG, "nms_run: null pointer");
    const rotgeo::NmsSizes z = sizes_from(sizes);
    if (z.n_pad < 128 || z.n_pad % 128 || z.tiles < 0 || z.mask_words < 0 || z.max_blocks < 1 || z.nseg < 0)
        return fail(ROTGEO_ERR_ARG, "nms_run: bad sizes");
    if (z.max_blocks >= 65536)     // the mask kernel carries the column block in 16 bits
        return fail(ROTGEO_ERR_ARG, "nms_run: a segment of more than %d boxes is not supported", 65535 * 128);
    const size_t need = rotgeo_nms_run_bytes(n, sizes);
    if (!workspace || !aligned16(workspace) || workspace_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "nms_run: workspace %zu < %zu bytes or misaligned", workspace_bytes, need);
    const int rc = rotgeo::nms_run(boxes, n, stride, offsets, sort_offsets, group_sizes_host, num_groups, seg_cap,
                                   sortws, z, iou_threshold, keep_out, num_keep_out, group_counts_out, workspace,
                                   (cudaStream_t)stream, det_scores, dets_out);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "nms_run: too many tiles");
    if (rc) return cuda_fail("nms_run");
    return ROTGEO_OK;
}

int rotgeo_nms_run(const float *boxes, int64_t n, int stride, const float *offsets, int sort_offsets,
                   const int64_t *group_sizes_host, int64_t num_groups, int64_t seg_cap, const void *sortws,
                   const int64_t *sizes, float iou_threshold, int64_t *keep_out, int64_t *num_keep_out,
                   int64_t *group_counts_out, void *workspace, size_t workspace_bytes, void *stream) {
    return nms_run_checked(boxes, n, stride, offsets, sort_offsets, group_sizes_host, num_groups, seg_cap, sortws, sizes,
                           iou_threshold, keep_out, num_keep_out, group_counts_out, workspace, workspace_bytes, stream,
                           nullptr, nullptr);
}

int rotgeo_nms_grouped(const float *boxes, int64_t n, int stride, const float *scores, const int64_t *seg_ids,
                       int id_offsets, const float *offsets, const int64_t *group_sizes_host, int64_t num_groups,
                       int64_t seg_cap, float iou_threshold, int64_t *keep_out, int64_t *tail_out,
                       int want_group_counts, float *dets_out, void *sortws, size_t sortws_bytes, void *workspace,
                       size_t workspace_bytes, size_t *need_out, void *stream) {
    if (n <= 0 || n > 0x7fffffffLL || stride < 5 || seg_cap < 1 || seg_cap > 65536 || !need_out)
        return fail(ROTGEO_ERR_ARG, "nms_grouped: n=%lld stride=%d seg_cap=%lld", (long long)n, stride, (long long)seg_cap);
    if (!groups_ok(group_sizes_host, num_groups, n))
        return fail(ROTGEO_ERR_ARG, "nms_grouped: group sizes must be non-negative and add up to n");
    int64_t sizes[6];
    rotgeo::NmsSizes z;
    rotgeo::nms_bounds(group_sizes_host, num_groups, n, seg_cap, &z);
    sizes_to(z, sizes);
    need_out[0] = rotgeo_nms_sort_bytes(n, group_sizes_host, num_groups, seg_cap);
    need_out[1] = rotgeo_nms_run_bytes(n, sizes);
    if (sortws_bytes < need_out[0] || workspace_bytes < need_out[1] || !sortws || !workspace)
        return fail(ROTGEO_ERR_WORKSPACE, "nms_grouped: buffers %zu / %zu bytes, need %zu / %zu", sortws_bytes,
                    workspace_bytes, need_out[0], need_out[1]);
    int rc = rotgeo_nms_sort(boxes, n, stride, scores, seg_ids, id_offsets, group_sizes_host, num_groups, seg_cap,
                             sortws, sortws_bytes, stream);
    if (rc) return rc;
    return nms_run_checked(boxes, n, stride, offsets, (id_offsets && seg_ids) ? 1 : 0, group_sizes_host, num_groups,
                           seg_cap, sortws, sizes, iou_threshold, keep_out, tail_out,
                           want_group_counts ? tail_out + 2 : nullptr, workspace, workspace_bytes, stream, scores,
                           dets_out);
}

size_t rotgeo_multiclass_nms_workspace_bytes(int64_t rows, int num_classes) {
    if (rows <= 0 || num_classes <= 0 || num_classes > rotgeo::kMcMaxClasses) return 0;
    return rotgeo::multiclass_nms_workspace_bytes(rows, num_classes);
}

int rotgeo_multiclass_nms(const float *multi_bboxes, int bbox_cols, const float *multi_scores,
                          const float *score_factors, int64_t rows, int num_classes,
                          float score_thr, float iou_threshold, int64_t cap, float *dets_out,
                          int64_t *labels_out, int64_t *keep_flat_out, int64_t *num_out,
                          void *workspace, size_t workspace_bytes, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (rows < 0 || num_classes < 1 || num_classes > rotgeo::kMcMaxClasses ||
        (bbox_cols != 5 && bbox_cols != 5 * num_classes) || cap < 0)
        return fail(ROTGEO_ERR_ARG, "multiclass_nms: rows=%lld classes=%d bbox_cols=%d cap=%lld",
                    (long long)rows, num_classes, bbox_cols, (long long)cap);
    if (rows * num_classes >= (1ll << 26))
        return fail(ROTGEO_ERR_ARG, "multiclass_nms: rows * classes must stay below 2^26");
    if (!num_out) return fail(ROTGEO_ERR_ARG, "multiclass_nms: null num_out");
    if (rows == 0 || cap == 0) {
        if (cudaMemsetAsync(num_out, 0, sizeof(int64_t), st) != cudaSuccess)
            return cuda_fail("multiclass_nms memset");
        return ROTGEO_OK;
    }
    if (!multi_bboxes || !multi_scores || !dets_out || !labels_out || !keep_flat_out)
        return fail(ROTGEO_ERR_ARG, "multiclass_nms: null pointer");
    if (!(iou_threshold > 0.f)) return fail(ROTGEO_ERR_ARG, "multiclass_nms: iou_threshold must be positive");
    const size_t need = rotgeo::multiclass_nms_workspace_bytes(rows, num_classes);
    if (!workspace || !aligned16(workspace) || workspace_bytes < need)
        return fail(ROTGEO_ERR_WORKSPACE, "multiclass_nms: workspace %zu < %zu bytes or misaligned",
                    workspace_bytes, need);
    const int shared = bbox_cols == 5 ? 1 : 0;   // one box per row, used by every class
    const int rc = rotgeo::launch_multiclass_nms(multi_bboxes, shared, multi_scores, score_factors,
                                                 rows, num_classes, score_thr, iou_threshold, cap,
                                                 dets_out, labels_out, keep_flat_out, num_out,
                                                 workspace, workspace_bytes, st);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "multiclass_nms: too many tiles");
    if (rc == -2) return fail(ROTGEO_ERR_WORKSPACE, "multiclass_nms: workspace too small");
    if (rc) return cuda_fail("multiclass_nms launch");
    return ROTGEO_OK;
}

// ---- multi-GPU exchange over peer memory -----------------------------------------------------------
static int make_peer_table(void *const *peer_bases_host, int world, int rank, rotgeo::PeerTable *t, const char *who) {
    if (!peer_bases_host || world < 1 || world > rotgeo::kPeerMax || rank < 0 || rank >= world)
        return fail(ROTGEO_ERR_ARG, "%s: world=%d rank=%d (at most %d ranks)", who, world, rank, rotgeo::kPeerMax);
    for (int r = 0; r < rotgeo::kPeerMax; ++r) t->base[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (!peer_bases_host[r] || !aligned16(peer_bases_host[r]))
            return fail(ROTGEO_ERR_ARG, "%s: window of rank %d is null or misaligned", who, r);
        t->base[r] = peer_bases_host[r];
    }
    t->world = world;
    t->rank = rank;
    return ROTGEO_OK;
}

size_t rotgeo_peer_pad_bytes(void) { return rotgeo::kPeerPadBytes; }

int rotgeo_peer_barrier(void *const *peer_bases_host, int world, int rank, uint64_t epoch, void *stream) {
    rotgeo::PeerTable t;
    if (int rc = make_peer_table(peer_bases_host, world, rank, &t, "peer_barrier")) return rc;
    if (rotgeo::launch_peer_barrier(t, epoch, (cudaStream_t)stream)) return cuda_fail("peer_barrier launch");
    return ROTGEO_OK;
}

int rotgeo_dist_iou_allgather(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t matrix_offset,
                              const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                              int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                              void *workspace, size_t workspace_bytes, void *stream) {
    rotgeo::PeerTable t;
    if (int rc = make_peer_table(peer_bases_host, world, rank, &t, "dist_iou_allgather")) return rc;
    if (n1_local < 0 || row_lo < 0 || row_lo + n1_local > n1_total || n2 <= 0 || stride1 < 5 || stride2 < 5 ||
        (mode != 0 && mode != 1) || matrix_offset < rotgeo::kPeerPadBytes || matrix_offset % 16)
        return fail(ROTGEO_ERR_ARG, "dist_iou_allgather: rows [%lld, +%lld) of %lld, n2=%lld, mode=%d, offset=%zu",
                    (long long)row_lo, (long long)n1_local, (long long)n1_total, (long long)n2, mode, matrix_offset);
    if (n1_local > 0 && (!boxes1_local || !boxes2)) return fail(ROTGEO_ERR_ARG, "dist_iou_allgather: null pointer");
    const size_t need = n1_local > 0 ? rotgeo_iou_workspace_bytes(n1_local, n2) : 0;
    if (need && (!workspace || !aligned16(workspace) || workspace_bytes < need))
        return fail(ROTGEO_ERR_WORKSPACE, "dist_iou_allgather: workspace %zu < %zu bytes or misaligned", workspace_bytes, need);
    const int rc = rotgeo::launch_dist_iou_allgather(t, epoch, matrix_offset, boxes1_local, row_lo, n1_local, n1_total,
                                                     stride1, boxes2, n2, stride2, mode, min_wh, workspace,
                                                     (cudaStream_t)stream);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "dist_iou_allgather: too many tiles");
    if (rc) return cuda_fail("dist_iou_allgather launch");
    return ROTGEO_OK;
}

size_t rotgeo_dist_rowcol_stage_bytes(int64_t n1_total, int64_t n2, int world) {
    if (n1_total < 0 || n2 < 0 || world < 1) return 0;
    return align_up((size_t)n1_total * 8, 256) + align_up((size_t)world * n2 * 8, 256);
}

int rotgeo_dist_iou_rowcol_max(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t stage_offset,
                               const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                               int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                               float *row_max, int64_t *row_argmax, float *col_max, int64_t *col_argmax,
                               void *workspace, size_t workspace_bytes, void *stream) {
    rotgeo::PeerTable t;
    if (int rc = make_peer_table(peer_bases_host, world, rank, &t, "dist_iou_rowcol_max")) return rc;
    if (n1_local < 0 || row_lo < 0 || row_lo + n1_local > n1_total || n2 <= 0 || stride1 < 5 || stride2 < 5 ||
        (mode != 0 && mode != 1) || stage_offset < rotgeo::kPeerPadBytes || stage_offset % 16 ||
        n1_total >= 0x7fffffffLL)
        return fail(ROTGEO_ERR_ARG, "dist_iou_rowcol_max: rows [%lld, +%lld) of %lld, n2=%lld, mode=%d",
                    (long long)row_lo, (long long)n1_local, (long long)n1_total, (long long)n2, mode);
    if (n1_local > 0 && (!boxes1_local || !boxes2)) return fail(ROTGEO_ERR_ARG, "dist_iou_rowcol_max: null pointer");
    const size_t need = n1_local > 0 ? rotgeo::iou_max_workspace_bytes(n1_local, n2) : 0;
    if (need && (!workspace || !aligned16(workspace) || workspace_bytes < need))
        return fail(ROTGEO_ERR_WORKSPACE, "dist_iou_rowcol_max: workspace %zu < %zu bytes or misaligned", workspace_bytes, need);
    const int rc = rotgeo::launch_dist_iou_rowcol_max(t, epoch, stage_offset, boxes1_local, row_lo, n1_local, n1_total,
                                                      stride1, boxes2, n2, stride2, mode, min_wh, row_max, row_argmax,
                                                      col_max, col_argmax, workspace, (cudaStream_t)stream);
    if (rc == -1) return fail(ROTGEO_ERR_ARG, "dist_iou_rowcol_max: too many tiles");
    if (rc) return cuda_fail("dist_iou_rowcol_max launch");
    return ROTGEO_OK;
}

size_t rotgeo_dist_keeps_bytes(int64_t num_groups_total, int64_t cap, int world, int row_floats) {
    if (num_groups_total < 0 || cap < 0 || world < 1 || row_floats < 0) return 0;
    return align_up((size_t)num_groups_total * 8 + (size_t)world * 8 + (size_t)num_groups_total * cap * 8 +
                        (size_t)num_groups_total * cap * row_floats * 4, 256);
}

int rotgeo_dist_push_keeps(void *const *peer_bases_host, int world, int rank, uint64_t epoch, size_t out_offset,
                           const int64_t *keep, const int64_t *tail, int num_local, const int64_t *group_ids_host,
                           int64_t cap, int64_t num_groups_total, const float *rows_src, int row_floats,
                           int row_stride, const int64_t *group_base_host, void *stream) {
    rotgeo::PeerTable t;
    if (int rc = make_peer_table(peer_bases_host, world, rank, &t, "dist_push_keeps")) return rc;
    if (num_local < 0 || num_local > rotgeo::kSortGroupsInline || cap < 1 || num_groups_total < 1 || row_floats < 0 ||
        row_floats > 8 || out_offset < rotgeo::kPeerPadBytes || out_offset % 16)
        return fail(ROTGEO_ERR_ARG, "dist_push_keeps: local groups=%d (at most %d), cap=%lld, row_floats=%d", num_local,
                    rotgeo::kSortGroupsInline, (long long)cap, row_floats);
    if (num_local > 0 && (!keep || !tail || !group_ids_host)) return fail(ROTGEO_ERR_ARG, "dist_push_keeps: null pointer");
    if (rows_src && (!group_base_host || row_stride < row_floats))
        return fail(ROTGEO_ERR_ARG, "dist_push_keeps: payload rows need group_base_host and row_stride >= row_floats");
    for (int g = 0; g < num_local; ++g)
        if (group_ids_host[g] < 0 || group_ids_host[g] >= num_groups_total)
            return fail(ROTGEO_ERR_ARG, "dist_push_keeps: group id out of range");
    const int rc = rotgeo::launch_dist_push_keeps(t, epoch, out_offset, keep, tail, num_local, group_ids_host, cap,
                                                  num_groups_total, rows_src, row_floats, row_stride, group_base_host,
                                                  (cudaStream_t)stream);
    if (rc) return cuda_fail("dist_push_keeps launch");
    return ROTGEO_OK;
}

}  // extern "C"
