// Detection-vs-ground-truth matching of the mAP evaluation, batched over every
// (class, image) pair in one launch.  sm_100a.
//
// Replaces the per-image body of the reference's tpfp_default
// (mmrotate/core/evaluation/eval_map.py:12-94), which eval_rbbox_map (eval_map.py:126-246)
// farms out to a 4-process spawn pool once per class: box_iou_rotated(dets, gts) on the CPU,
// max / argmax per detection, then a greedy pass over the detections in descending score
// order that marks each one true positive (its best ground truth has IoU >= thr, is not
// ignored and was not claimed before), false positive, or neither (best ground truth is an
// ignored one).
//
// One CTA per segment (= class x image), segments given by offsets into detections and
// ground truths that are already grouped by segment:
//   1. ground truths are converted to the clipper's form (centre, cos/sin, half extents)
//      in shared memory, 128 at a time; every thread owns detections and scans them —
//      circle pre-test, exact clipper — keeping max and FIRST argmax (numpy's argmax)
//   2. rank of every detection by (score descending, index ascending) — counting, the
//      segments are small
//   3. warp 0 walks the detections 32 ranks at a time: lanes that hit a live ground truth
//      are grouped by ground truth (__match_any_sync); the lowest rank of a group claims it.
// Raw boxes in, no packing: the segments are tiny (tens of boxes), the launch is the cost.
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kEvalThreads = 128;

__device__ __forceinline__ void eval_load_box(const float *b, float4 &p0, float4 &p1) {
    const float w = b[2], h = b[3];
    double sd, cd;
    sincos((double)b[4], &sd, &cd);
    const float a = 0.5f * w, hb = 0.5f * h;
    p0 = make_float4(b[0], b[1], circum_radius(a, hb), w * h);
    p1 = make_float4((float)cd, (float)sd, a, hb);
}

__global__ void __launch_bounds__(kEvalThreads)
eval_tpfp_kernel(const float *__restrict__ dets, int det_stride, const int64_t *__restrict__ det_off,
                 const float *__restrict__ gts, int gt_stride, const int64_t *__restrict__ gt_off,
                 const uint8_t *__restrict__ gt_ignore, int64_t num_segments, float iou_thr,
                 uint8_t *__restrict__ tp, uint8_t *__restrict__ fp, float *__restrict__ iou_max,
                 int *__restrict__ iou_arg, int *__restrict__ order, uint8_t *__restrict__ covered) {
    __shared__ float4 s_g0[kEvalThreads], s_g1[kEvalThreads];
    const int tid = threadIdx.x, lane = tid & 31;
    for (int64_t s = blockIdx.x; s < num_segments; s += gridDim.x) {
        const int64_t d0 = det_off[s], nd = det_off[s + 1] - d0;
        const int64_t g0 = gt_off[s], ng = gt_off[s + 1] - g0;
        if (nd == 0) continue;
        // ---- 1. max / first argmax over the segment's ground truths -------------------------
        for (int64_t db = 0; db < nd; db += kEvalThreads) {
            const int64_t d = db + tid;
            float4 A0 = make_float4(0.f, 0.f, 0.f, 0.f), A1 = A0;
            if (d < nd) eval_load_box(dets + (d0 + d) * det_stride, A0, A1);
            float best = 0.f;            // numpy: max of an empty row does not happen (ng == 0 is handled below)
            int arg = 0;
            bool first = true;
            for (int64_t gb = 0; gb < ng; gb += kEvalThreads) {
                __syncthreads();
                if (gb + tid < ng) eval_load_box(gts + (g0 + gb + tid) * gt_stride, s_g0[tid], s_g1[tid]);
                __syncthreads();
                const int cnt = (int)min((int64_t)kEvalThreads, ng - gb);
                if (d < nd) {
                    for (int j = 0; j < cnt; ++j) {
                        const float4 B0 = s_g0[j];
                        float v = 0.f;
                        if (!circles_disjoint(A0.x, A0.y, A0.z, B0.x, B0.y, B0.z))
                            v = rect_overlap<0>(A0, A1, B0, s_g1[j]);
                        if (first || v > best) {
                            best = v;
                            arg = (int)gb + j;
                            first = false;
                        }
                    }
                }
            }
            if (d < nd) {
                iou_max[d0 + d] = best;
                iou_arg[d0 + d] = arg;
            }
        }
        if (ng == 0) {
            // no ground truth of this class in this image: every detection is a false positive
            for (int64_t d = tid; d < nd; d += kEvalThreads) {
                tp[d0 + d] = 0;
                fp[d0 + d] = 1;
            }
            continue;
        }
        // ---- 2. rank by (score descending, index ascending) ---------------------------------
        for (int64_t d = tid; d < nd; d += kEvalThreads) {
            const float sc = dets[(d0 + d) * det_stride + 5];
            int rank = 0;
            for (int64_t e = 0; e < nd; ++e) {
                const float se = dets[(d0 + e) * det_stride + 5];
                rank += (se > sc) || (se == sc && e < d);
            }
            order[d0 + rank] = (int)d;
        }
        __syncthreads();
        // ---- 3. greedy claim in rank order, 32 ranks at a time --------------------------------
        if (tid < 32) {
            for (int64_t rb = 0; rb < nd; rb += 32) {
                const int64_t r = rb + lane;
                int d = -1, g = -1;
                bool hit = false, live = false;
                if (r < nd) {
                    d = order[d0 + r];
                    hit = iou_max[d0 + d] >= iou_thr;
                    g = iou_arg[d0 + d];
                    live = hit && !(gt_ignore && gt_ignore[g0 + g]);
                }
                // lanes that hit the same live ground truth: the lowest rank claims it
                const unsigned grp = __match_any_sync(0xffffffffu, live ? g : -1 - lane);
                const bool leader = live && (__ffs(grp) - 1 == lane);
                bool is_tp = false;
                if (leader) {
                    is_tp = covered[g0 + g] == 0;
                    covered[g0 + g] = 1;
                }
                if (r < nd) {
                    // hit a live gt: tp if claimed now, else fp; hit an ignored gt: neither; no hit: fp
                    tp[d0 + d] = is_tp ? 1 : 0;
                    fp[d0 + d] = hit ? ((live && !is_tp) ? 1 : 0) : 1;
                }
                __syncwarp();
            }
        }
        __syncthreads();
    }
}

struct EvalLayout {
    size_t off_max, off_arg, off_order, off_covered, total_bytes;
};
static void eval_layout(int64_t num_dets, int64_t num_gts, EvalLayout *l) {
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    size_t off = 0;
    l->off_max = off;     off = up(off + (size_t)num_dets * 4);
    l->off_arg = off;     off = up(off + (size_t)num_dets * 4);
    l->off_order = off;   off = up(off + (size_t)num_dets * 4);
    l->off_covered = off; off = up(off + (size_t)num_gts);
    l->total_bytes = off;
}
size_t eval_tpfp_workspace_bytes(int64_t num_dets, int64_t num_gts) {
    EvalLayout l;
    eval_layout(num_dets > 0 ? num_dets : 0, num_gts > 0 ? num_gts : 0, &l);
    return l.total_bytes;
}

int launch_eval_tpfp(const float *dets, int64_t num_dets, int det_stride, const int64_t *det_off,
                     const float *gts, int64_t num_gts, int gt_stride, const int64_t *gt_off,
                     const uint8_t *gt_ignore, int64_t num_segments, float iou_thr, uint8_t *tp,
                     uint8_t *fp, float *iou_max_out, void *workspace, cudaStream_t stream) {
    EvalLayout l;
    eval_layout(num_dets, num_gts, &l);
    char *ws = reinterpret_cast<char *>(workspace);
    if (num_gts > 0 && cudaMemsetAsync(ws + l.off_covered, 0, (size_t)num_gts, stream) != cudaSuccess) return -3;
    float *vmax = reinterpret_cast<float *>(ws + l.off_max);
    int *arg = reinterpret_cast<int *>(ws + l.off_arg);
    const int64_t blocks = num_segments < 148 * 16 ? num_segments : 148 * 16;
    eval_tpfp_kernel<<<(unsigned)blocks, kEvalThreads, 0, stream>>>(
        dets, det_stride, det_off, gts, gt_stride, gt_off, gt_ignore, num_segments, iou_thr, tp, fp, vmax, arg,
        reinterpret_cast<int *>(ws + l.off_order), reinterpret_cast<uint8_t *>(ws + l.off_covered));
    if (cudaGetLastError() != cudaSuccess) return -3;
    if (iou_max_out &&
        cudaMemcpyAsync(iou_max_out, vmax, (size_t)num_dets * 4, cudaMemcpyDeviceToDevice, stream) != cudaSuccess)
        return -3;
    return 0;
}

}  // namespace rotgeo
