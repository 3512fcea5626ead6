// Fused assigner reductions: what every caller of the IoU calculator does next, without the
// N x M matrix.  sm_100a.
//
// The reference's assigners take the (num_gts, num_bboxes) matrix from
// RBboxOverlaps2D (mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:44) and
// immediately reduce it: max / argmax over gts for every box, max / argmax over boxes for
// every gt, thresholds, low-quality matches (mmdet's MaxIoUAssigner.assign_wrt_overlaps,
// the assigner of the fork's RetinaNet / RPN / R-CNN configs; ATSS and the fork's
// OBBLabelAssigner, mmrotate/core/bbox/assigners/atss_obb_assigner.py:70,118-145 and
// obb_label_assigner.py:128-143, do the same kind of reduction).  At the C3 shape the matrix
// is 268 MB written and read again several times; here the tile kernel of rotgeo_iou.cu
// runs with a reduction sink instead (atomicMax on value|index keys for the few pairs that
// overlap at all) and nothing of size N x M ever exists.
//
//   pack (both sets, clamp fused) -> iou_tile_kernel<kSinkMax>      row / column maxima + argmax
//                                 -> decode_keys_kernel             keys -> float max, int64 argmax
//   for the assigner additionally
//                                 -> gt_lowq_kernel                 per gt: low-quality match target(s)
//                                 -> iou_tile_kernel<kSinkRowMaxHits>  (gt_max_assign_all) every box that
//                                                                   attains a gt's maximum
//                                 -> assign_finalize_kernel         thresholds + overrides per box
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

static inline size_t align_up_sz(size_t x, size_t a) { return (x + a - 1) / a * a; }

// key = float bits << 32 | (0xffffffff - index); 0 = no positive value: max 0 at index 0 (the
// first maximal element, which is what torch.max returns for an all-zero line)
__global__ void decode_keys_kernel(const unsigned long long *__restrict__ keys, int64_t n,
                                   float *__restrict__ vmax, int64_t *__restrict__ arg) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long k = keys[i];
    if (vmax) vmax[i] = __uint_as_float((unsigned)(k >> 32));
    if (arg) arg[i] = k ? (int64_t)(0xffffffffu - (unsigned)k) : 0;
}

// per gt i (ascending i overrides, so the final owner of a box is the LARGEST such i):
//   not assign_all: box gt_argmax[i] <- i + 1   if gt_max[i] >= min_pos_iou
//   assign_all, gt_max[i] == 0 >= min_pos_iou: EVERY box has overlap 0 == gt_max[i] with it
__global__ void gt_lowq_kernel(const unsigned long long *__restrict__ row_key, int64_t num_gts,
                               float min_pos_iou, int assign_all, int *__restrict__ col_lowq,
                               int *__restrict__ all_cols_lowq) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= num_gts) return;
    const unsigned long long k = row_key[i];
    const float m = __uint_as_float((unsigned)(k >> 32));
    if (!(m >= min_pos_iou)) return;
    if (assign_all) {
        if (k == 0ull) atomicMax(all_cols_lowq, (int)i + 1);
    } else {
        const int64_t j = k ? (int64_t)(0xffffffffu - (unsigned)k) : 0;
        atomicMax(col_lowq + j, (int)i + 1);
    }
}

// mmdet MaxIoUAssigner.assign_wrt_overlaps, steps 2-4, per box
__global__ void assign_finalize_kernel(const unsigned long long *__restrict__ col_key, int64_t num_boxes,
                                       float pos_iou_thr, float neg_lo, float neg_hi,
                                       int match_low_quality, const int *__restrict__ col_lowq,
                                       const int *__restrict__ all_cols_lowq,
                                       int64_t *__restrict__ assigned, float *__restrict__ max_overlaps) {
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j >= num_boxes) return;
    const unsigned long long k = col_key[j];
    const float mo = __uint_as_float((unsigned)(k >> 32));
    const int64_t am = k ? (int64_t)(0xffffffffu - (unsigned)k) : 0;
    int64_t a = -1;
    if (mo >= neg_lo && mo < neg_hi) a = 0;
    if (mo >= pos_iou_thr) a = am + 1;
    if (match_low_quality) {
        int lq = col_lowq[j];
        const int all = *all_cols_lowq;
        lq = lq > all ? lq : all;
        if (lq > 0) a = lq;
    }
    assigned[j] = a;
    if (max_overlaps) max_overlaps[j] = mo;
}

struct AssignLayout {
    size_t off_pk1, off_pk2, off_row_key, off_col_key, off_row_max, off_lowq, off_all, total_bytes;
};
static void assign_layout(int64_t n1, int64_t n2, AssignLayout *l) {
    size_t off = 0;
    l->off_pk1 = off;     off = align_up_sz(off + (size_t)padded_count(n1) * kPackedBytesPerBox, 256);
    l->off_pk2 = off;     off = align_up_sz(off + (size_t)padded_count(n2) * kPackedBytesPerBox, 256);
    // the three zero-initialised arrays are contiguous: one memset
    l->off_row_key = off; off += (size_t)n1 * 8;
    l->off_col_key = off; off += (size_t)n2 * 8;
    l->off_lowq = off;    off += (size_t)n2 * 4;
    l->off_all = off;     off = align_up_sz(off + 4, 256);
    l->off_row_max = off; off = align_up_sz(off + (size_t)n1 * 4, 256);
    l->total_bytes = off;
}
size_t iou_max_workspace_bytes(int64_t n1, int64_t n2) {
    AssignLayout l;
    assign_layout(n1 > 0 ? n1 : 0, n2 > 0 ? n2 : 0, &l);
    return l.total_bytes;
}

static int pack_and_reduce(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2,
                           int stride2, int mode, float min_wh, char *ws, const AssignLayout &l,
                           cudaStream_t stream) {
    if (cudaMemsetAsync(ws + l.off_row_key, 0, l.off_all + 4 - l.off_row_key, stream) != cudaSuccess) return -3;
    const PackSet s1 = make_pack_set(boxes1, n1, stride1, nullptr, nullptr, ws + l.off_pk1, padded_count(n1));
    const PackSet s2 = make_pack_set(boxes2, n2, stride2, nullptr, nullptr, ws + l.off_pk2, padded_count(n2));
    if (launch_pack(s1, s2, min_wh, 0.f, 0.f, stream)) return -3;
    return launch_iou_max(ws + l.off_pk1, n1, ws + l.off_pk2, n2, mode,
                          reinterpret_cast<unsigned long long *>(ws + l.off_row_key),
                          reinterpret_cast<unsigned long long *>(ws + l.off_col_key), true, stream);
}

int launch_iou_max_keys(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2, int stride2,
                        int mode, float min_wh, void *workspace, unsigned long long **row_key,
                        unsigned long long **col_key, cudaStream_t stream) {
    AssignLayout l;
    assign_layout(n1, n2, &l);
    char *ws = reinterpret_cast<char *>(workspace);
    *row_key = reinterpret_cast<unsigned long long *>(ws + l.off_row_key);
    *col_key = reinterpret_cast<unsigned long long *>(ws + l.off_col_key);
    return pack_and_reduce(boxes1, n1, stride1, boxes2, n2, stride2, mode, min_wh, ws, l, stream);
}

int launch_iou_rowcol_max(const float *boxes1, int64_t n1, int stride1, const float *boxes2, int64_t n2,
                          int stride2, int mode, float min_wh, float *row_max, int64_t *row_argmax,
                          float *col_max, int64_t *col_argmax, void *workspace, cudaStream_t stream) {
    AssignLayout l;
    assign_layout(n1, n2, &l);
    char *ws = reinterpret_cast<char *>(workspace);
    const int rc = pack_and_reduce(boxes1, n1, stride1, boxes2, n2, stride2, mode, min_wh, ws, l, stream);
    if (rc) return rc;
    if (row_max || row_argmax)
        decode_keys_kernel<<<(unsigned)((n1 + 255) / 256), 256, 0, stream>>>(
            reinterpret_cast<const unsigned long long *>(ws + l.off_row_key), n1, row_max, row_argmax);
    if (col_max || col_argmax)
        decode_keys_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, stream>>>(
            reinterpret_cast<const unsigned long long *>(ws + l.off_col_key), n2, col_max, col_argmax);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int launch_max_iou_assign(const float *gts, int64_t num_gts, int stride_g, const float *boxes,
                          int64_t num_boxes, int stride_b, int mode, float min_wh, float pos_iou_thr,
                          float neg_lo, float neg_hi, float min_pos_iou, int match_low_quality,
                          int gt_max_assign_all, int64_t *assigned, float *max_overlaps, void *workspace,
                          cudaStream_t stream) {
    AssignLayout l;
    assign_layout(num_gts, num_boxes, &l);
    char *ws = reinterpret_cast<char *>(workspace);
    int rc = pack_and_reduce(gts, num_gts, stride_g, boxes, num_boxes, stride_b, mode, min_wh, ws, l, stream);
    if (rc) return rc;
    unsigned long long *row_key = reinterpret_cast<unsigned long long *>(ws + l.off_row_key);
    unsigned long long *col_key = reinterpret_cast<unsigned long long *>(ws + l.off_col_key);
    int *lowq = reinterpret_cast<int *>(ws + l.off_lowq);
    int *all_lowq = reinterpret_cast<int *>(ws + l.off_all);
    if (match_low_quality) {
        gt_lowq_kernel<<<(unsigned)((num_gts + 255) / 256), 256, 0, stream>>>(
            row_key, num_gts, min_pos_iou, gt_max_assign_all, lowq, all_lowq);
        if (gt_max_assign_all) {
            float *row_max = reinterpret_cast<float *>(ws + l.off_row_max);
            decode_keys_kernel<<<(unsigned)((num_gts + 255) / 256), 256, 0, stream>>>(row_key, num_gts,
                                                                                     row_max, nullptr);
            rc = launch_iou_row_max_hits(ws + l.off_pk1, num_gts, ws + l.off_pk2, num_boxes, mode, row_max,
                                         min_pos_iou, lowq, stream);
            if (rc) return rc;
        }
    }
    assign_finalize_kernel<<<(unsigned)((num_boxes + 255) / 256), 256, 0, stream>>>(
        col_key, num_boxes, pos_iou_thr, neg_lo, neg_hi, match_low_quality, lowq, all_lowq, assigned,
        max_overlaps);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
