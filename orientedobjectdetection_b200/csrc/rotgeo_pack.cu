// AoS (cx, cy, w, h, theta) rows -> packed SoA float4 planes (see rotgeo_device.cuh).
// One thread per box, one launch for up to two box sets.  The optional gather /
// centre offset / size clamp are the pieces of the reference's python front-end
// that would otherwise be separate torch kernels (clone + clamp_ at
// rotate_iou2d_calculator.py:84-87, the offset add at bbox_nms_rotated.py:73-77,
// index_select by score order in mmcv's nms_rotated wrapper).
//
// The bin width of the cull masks is derived in a prologue every CTA repeats on
// the same strided sample of <= 256 boxes (so all CTAs agree without a second
// launch): 32 bins span the sampled extent of the RAW centres — before the class
// offsets, because a class segment only ever meets itself — and, for NMS packs, 30 of
// the 32 area bins span the sampled range of log2(area).  The choices affect speed
// only; bin_range() and area_bin() are conservative for any width.
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kPackThreads = 256;

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}

// position (index into gather[]) of slot i of a set, or -1 for padding.  Plain sets:
// slot i = position i.  Segmented sets (NMS): every segment starts at a slot that is
// a multiple of 128 (so each 128-slot table block lies inside one segment) and is
// padded up to the next one; slots past the last segment are padding too.
// seg_hint >= 0: the segment is already known (cooperative lookup of the CTA, see pack_kernel)
__device__ __forceinline__ int64_t slot_position(const PackSet &s, int64_t i, int64_t n, int seg_hint = -1) {
    if (s.segs) {
        int lo = 0, hi = s.nseg - 1;   // last segment with pstart <= i
        if (seg_hint >= 0) lo = hi = seg_hint;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s.segs[mid].pstart <= i) lo = mid; else hi = mid - 1;
        }
        const int64_t pos = i - s.segs[lo].pstart;
        if (pos >= s.segs[lo].len) return -1;
        return s.segs[lo].start + pos;
    }
    return i < n ? i : -1;
}

// source row of position `pos` and the centre offset that goes with it
__device__ __forceinline__ int64_t position_source(const PackSet &s, int64_t pos, float &offset) {
    const int64_t v = s.gather32 ? (int64_t)s.gather32[pos] : (s.gather ? s.gather[pos] : pos);
    offset = 0.f;
    if (s.mc_classes > 0) {
        const int64_t row = v / s.mc_classes;
        const int cls = (int)(v - row * s.mc_classes);
        offset = __fmul_rn((float)cls, __fadd_rn(__fadd_rn(s.mc_max[0], s.mc_max[1]), 1.f));
        return s.mc_shared ? row : v;
    }
    if (s.offsets) offset = s.offsets[v];
    // mmcv.ops.batched_nms: id * (max(cx,cy) + max(w,h) + 1), all in fp32
    // (no fused multiply-add anywhere here: torch rounds the product, then the sum)
    if (s.offset_ids)
        offset = __fmul_rn((float)s.offset_ids[v], __fadd_rn(__fadd_rn(s.mc_max[0], s.mc_max[1]), 1.f));
    return v;
}

struct BinScale {
    float inv_cell;          // 1 / (spatial bin width)
    float a_lo, a_inv;       // area bin = floor((log2(area) - a_lo) * a_inv) + 1, clamped to 0 .. 31
};

// identical in every CTA of the launch: every CTA reduces the same strided sample of <= 256
// raw boxes (centres and log2 areas)
__device__ BinScale sample_bin_scale(const PackSet &s1, const PackSet &s2) {
    __shared__ float red[8][kPackThreads / 32];
    const int tid = threadIdx.x;
    const bool two = s2.n > 0;
    const int64_t n1 = s1.n_dev ? *s1.n_dev : s1.n;
    // first half of the block samples set 1, second half set 2 (all of it set 1 if alone)
    const PackSet s = (two && tid >= kPackThreads / 2) ? s2 : s1;
    const int slots = two ? kPackThreads / 2 : kPackThreads;
    const int t = two ? (tid & (kPackThreads / 2 - 1)) : tid;
    float lo_x = INFINITY, hi_x = -INFINITY, lo_y = INFINITY, hi_y = -INFINITY;
    float lo_a = INFINITY, hi_a = -INFINITY;
    float sum_d = 0.f, cnt_d = 0.f;                  // sampled box diagonals (spatially packed sets)
    const int64_t sn = (two && tid >= kPackThreads / 2) ? s2.n : n1;   // only set 1 is ever device-counted
    if (sn > 0) {
        const int64_t i = (sn >= slots) ? (int64_t)t * (sn / slots) : (t % sn);
        float unused;
        const float *b = s.boxes + position_source(s, i, unused) * s.stride;
        const float cx = b[0], cy = b[1];
        if (fabsf(cx) < 1e30f && fabsf(cy) < 1e30f) {
            lo_x = hi_x = cx;
            lo_y = hi_y = cy;
        }
        const float la = __log2f(b[2] * b[3]);
        if (fabsf(la) < 1e30f) lo_a = hi_a = la;     // also rejects NaN / -inf (zero area)
        const float dm = sqrtf(fmaf(b[2], b[2], b[3] * b[3]));
        if (dm < 1e30f) { sum_d = dm; cnt_d = 1.f; }
    }
    lo_x = warp_min(lo_x); hi_x = warp_max(hi_x);
    lo_y = warp_min(lo_y); hi_y = warp_max(hi_y);
    lo_a = warp_min(lo_a); hi_a = warp_max(hi_a);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum_d += __shfl_xor_sync(0xffffffffu, sum_d, d);
        cnt_d += __shfl_xor_sync(0xffffffffu, cnt_d, d);
    }
    if ((tid & 31) == 0) {
        red[6][tid >> 5] = sum_d; red[7][tid >> 5] = cnt_d;
        red[0][tid >> 5] = lo_x; red[1][tid >> 5] = hi_x;
        red[2][tid >> 5] = lo_y; red[3][tid >> 5] = hi_y;
        red[4][tid >> 5] = lo_a; red[5][tid >> 5] = hi_a;
    }
    __syncthreads();
    sum_d = 0.f; cnt_d = 0.f;
#pragma unroll
    for (int w = 0; w < kPackThreads / 32; ++w) {
        sum_d += red[6][w]; cnt_d += red[7][w];
        lo_x = fminf(lo_x, red[0][w]); hi_x = fmaxf(hi_x, red[1][w]);
        lo_y = fminf(lo_y, red[2][w]); hi_y = fmaxf(hi_y, red[3][w]);
        lo_a = fminf(lo_a, red[4][w]); hi_a = fmaxf(hi_a, red[5][w]);
    }
    BinScale sc;
    const float extent = fmaxf(hi_x - lo_x, hi_y - lo_y);
    // 30 bins across the sampled extent (the sample misses a little at the edges)
    float cell = extent * (1.f / 30.f);
    // Spatially packed sets: a tile pairs two NEIGHBOURING blocks, whose boxes all sit in the same two or
    // three bins of an extent-wide grid (C5: a quarter of a tile's pairs passed the bins).  The bins are
    // modular, so any width is correct; one mean box diagonal keeps a box in two or three of the 32 bins
    // wherever it lies: (3 / 32)^2 of the pairs of a local tile survive.
    if (s1.aabb_out && cnt_d > 0.f) cell = fminf(cell, fmaxf(sum_d / cnt_d, extent * (1.f / 30.f / 64.f)));
    if (!(cell > 1e-30f) || !(cell < 1e30f)) cell = 1.f;
    sc.inv_cell = 1.f / cell;
    const float span = hi_a - lo_a;
    sc.a_lo = (span >= 0.f) ? lo_a : 0.f;
    sc.a_inv = (span > 1e-6f && span < 1e30f) ? 30.f / span : 1.f;
    return sc;
}

// area bin of log2(area) value `la` (monotone, clamped; NaN -> 0)
__device__ __forceinline__ int area_bin(float la, const BinScale &sc) {
    const float t = floorf((la - sc.a_lo) * sc.a_inv) + 1.f;
    return (int)fminf(fmaxf(t, 0.f), 31.f);
}

// One CTA packs 256 consecutive slots = two table blocks (slot counts are
// multiples of 128, so a block never straddles the two sets); warp v packs the 32
// boxes of word v % 4 of block v / 4.
__global__ void __launch_bounds__(kPackThreads)
pack_kernel(PackSet s1, PackSet s2, float min_wh, float cell, float area_ratio) {
    __shared__ uint32_t s_t0[2][3][32][4];   // level 0: [block][axis x, y, area][bin][word]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the consumer kernel may start now: it streams zeros until it needs the packed set
    pdl_launch_dependents();
    const int64_t g = blockIdx.x * (int64_t)kPackThreads + tid;
    // device-planned sets: blocks past the slots actually in use have nothing to do (block-uniform)
    if (s1.n_pad_dev && blockIdx.x * (int64_t)kPackThreads >= *s1.n_pad_dev) return;
    const bool second = g >= s1.n_pad;
    PackSet s = second ? s2 : s1;
    if (s.nseg_dev) s.nseg = (int)*s.nseg_dev;
    const int64_t i = second ? g - s1.n_pad : g;
    const bool in_set = i < s.n_pad;

    // Segmented sets: a 128-slot block lies inside one segment, so the 128 threads of a block find
    // it together — every thread tests a few table entries, one count per block — instead of each
    // thread walking a binary search of dependent loads (22 % of this kernel's stall samples on the
    // 320-segment RPN batch).  Tables of more than 512 segments keep the per-thread search.
    __shared__ int s_segcnt[2];
    const int hb = tid >> 7;
    const bool coop = in_set && s.segs != nullptr && s.nseg <= 512;
    if (tid < 2) s_segcnt[tid] = 0;
    __syncthreads();
    if (coop) {
        const int64_t b0 = i & ~(int64_t)127;
        int c = 0;
        for (int j = tid & 127; j < s.nseg; j += 128) c += s.segs[j].pstart <= b0 ? 1 : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
        if (lane == 0 && c) atomicAdd(&s_segcnt[hb], c);
    }
    __syncthreads();
    const int seg_hint = coop ? max(s_segcnt[hb] - 1, 0) : -1;

    // the box first: its loads are in flight while the bin width is being derived
    const int64_t pos = in_set ? slot_position(s, i, s.n_dev ? *s.n_dev : s.n, seg_hint) : -1;
    int64_t src = -1;
    float cx = 0.f, cy = 0.f, w = 0.f, h = 0.f, th = 0.f;
    if (pos >= 0) {
        float o;
        src = position_source(s, pos, o);
        const float *b = s.boxes + src * s.stride;
        cx = __fadd_rn(b[0], o); cy = __fadd_rn(b[1], o); w = b[2]; h = b[3]; th = b[4];
    }
    BinScale sc;
    if (cell > 0.f) {
        sc.inv_cell = 1.f / cell;
        sc.a_lo = 0.f;
        sc.a_inv = 1.f;
    } else {
        sc = sample_bin_scale(s1, s2);
    }
    const float inv_cell = sc.inv_cell;

    // NMS: scratch of the resolve stage starts from zero (the mask kernel only touches it
    // after its griddepcontrol.wait, i.e. after this grid has completed)
    if (in_set && !second && s1.z_kr) {
        for (int k = 0; k < s1.z_sum_words; ++k) s1.z_summary[i * s1.z_sum_words + k] = 0ull;
        if ((i & 63) == 0) {
            s1.z_kr[2 * (i >> 6)] = 0ull;
            s1.z_kr[2 * (i >> 6) + 1] = 0ull;
        }
        if (i < kNmsSyncWords) s1.z_sync[i] = 0u;
    }

    uint32_t xm = 0u, ym = 0u, am = 0u;
    float bx0 = INFINITY, by0 = INFINITY, bx1 = -INFINITY, by1 = -INFINITY;   // this slot's part of the block's box
    if (in_set && s.rank_out) s.rank_out[i] = pos >= 0 ? s.rank_src[pos] : 0xffffffffu;
    if (in_set) {
        if (src < 0) {
            // padding: occupies no bin, so no pair with it survives the cull
            s.p0[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            s.p1[i] = make_float4(1.f, 0.f, 0.f, 0.f);
            s.pr[i] = 0u;
        } else {
            if (min_wh > 0.f) {
                w = fmaxf(w, min_wh);
                h = fmaxf(h, min_wh);
            }
            // the reference takes cos/sin of theta in double and narrows (Spec A step 3)
            double sd, cd;
            sincos((double)th, &sd, &cd);
            const float a = 0.5f * w, hb = 0.5f * h;
            // radius for the bins: the circle-test radius plus the rounding of cx -/+ r
            const float r = circum_radius(a, hb);
            const float rx = r + fabsf(cx) * 2.4e-7f, ry = r + fabsf(cy) * 2.4e-7f;
            int sx, nx, sy, ny;
            bin_range(cx - rx, cx + rx, inv_cell, sx, nx);
            bin_range(cy - ry, cy + ry, inv_cell, sy, ny);
            xm = range_mask(sx, nx);
            ym = range_mask(sy, ny);
            // area axis (NMS): the box sits in one bin; a partner can reach the IoU threshold only
            // with an area inside [area * ratio, area / ratio] (IoU <= min(area) / max(area))
            int sa = 0, na = 0;
            if (area_ratio > 0.f && area_ratio < 1.f) {
                const float la = __log2f(w * h), dl = -__log2f(area_ratio);
                am = 1u << area_bin(la, sc);
                sa = area_bin(la - dl, sc);
                na = area_bin(la + dl, sc) - sa + 1;
                if (na >= 32 || !(fabsf(la) < 1e30f)) { sa = 0; na = 0; }     // no constraint
            }
            s.p0[i] = make_float4(cx, cy, r, w * h);
            s.p1[i] = make_float4((float)cd, (float)sd, a, hb);
            s.pr[i] = pack_ranges(sx, nx, sy, ny, sa, na);
            bx0 = cx - rx; bx1 = cx + rx; by0 = cy - ry; by1 = cy + ry;     // the radius the bins use (rounding included)
        }
    }

    // ---- bounding box of every 128-slot block (spatially packed NMS sets only; launch-uniform) ----
    if (s1.aabb_out) {
        __shared__ float s_box[kPackThreads / 32][4];
        bx0 = warp_min(bx0); by0 = warp_min(by0); bx1 = warp_max(bx1); by1 = warp_max(by1);
        if (lane == 0) { s_box[warp][0] = bx0; s_box[warp][1] = by0; s_box[warp][2] = bx1; s_box[warp][3] = by1; }
        __syncthreads();
        if (tid < 2) {
            const int64_t blk = blockIdx.x * 2 + tid;                       // block of set 1 (NMS packs have no set 2)
            if (blk * kBlock < s1.n_pad) {
                float4 bb = make_float4(INFINITY, INFINITY, -INFINITY, -INFINITY);
                for (int w4 = 0; w4 < 4; ++w4) {
                    bb.x = fminf(bb.x, s_box[4 * tid + w4][0]); bb.y = fminf(bb.y, s_box[4 * tid + w4][1]);
                    bb.z = fmaxf(bb.z, s_box[4 * tid + w4][2]); bb.w = fmaxf(bb.w, s_box[4 * tid + w4][3]);
                }
                s1.aabb_out[blk] = bb;
            }
        }
    }

    // ---- inverted index, level 0: transpose the warp's 32 masks with ballots -----------
    {
        uint32_t tx = 0u, ty = 0u;
#pragma unroll
        for (int b = 0; b < 32; ++b) {
            const uint32_t bx = __ballot_sync(0xffffffffu, (xm >> b) & 1u);
            const uint32_t by = __ballot_sync(0xffffffffu, (ym >> b) & 1u);
            if (lane == b) { tx = bx; ty = by; }
        }
        s_t0[warp >> 2][0][lane][warp & 3] = tx;   // bin = lane
        s_t0[warp >> 2][1][lane][warp & 3] = ty;
        if (area_ratio > 0.f && area_ratio < 1.f) {   // NMS packs only (launch-uniform)
            uint32_t ta = 0u;
#pragma unroll
            for (int b = 0; b < 32; ++b) {
                const uint32_t ba = __ballot_sync(0xffffffffu, (am >> b) & 1u);
                if (lane == b) ta = ba;
            }
            s_t0[warp >> 2][2][lane][warp & 3] = ta;
        }
    }
    __syncthreads();
    // ---- levels 0..4 by prefix OR over 16 consecutive bins, straight to global ----------
    // task (block h, axis a, bin b, word w): T[j][b][w] = OR of level 0 at bins b .. b+2^j-1
    const int axes = (area_ratio > 0.f && area_ratio < 1.f) ? 3 : 2;   // IoU packs: the area axis is never read
    for (int task = tid; task < 2 * axes * 32 * 4; task += kPackThreads) {
        const int tw = task & 3, b = (task >> 2) & 31, hb = (task >> 7) & 1, a = task >> 8;
        const int64_t g0 = blockIdx.x * (int64_t)kPackThreads + hb * kBlock;   // first slot of the block
        const bool sec = g0 >= s1.n_pad;
        const PackSet sb = sec ? s2 : s1;
        const int64_t i0 = sec ? g0 - s1.n_pad : g0;
        if (i0 >= sb.n_pad) continue;
        uint32_t *dst = sb.pt + (i0 / kBlock) * kTableWords + a * kTableAxisWords + b * 4 + tw;
        uint32_t acc = 0u;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            acc |= s_t0[hb][a][(b + k) & 31][tw];
            if (k == 0) dst[0] = acc;
            if (k == 1) dst[128] = acc;
            if (k == 3) dst[256] = acc;
            if (k == 7) dst[384] = acc;
            if (k == 15) dst[512] = acc;
        }
    }
}

int launch_pack(const PackSet &s1, const PackSet &s2, float min_wh, float cell, float area_ratio,
                cudaStream_t stream) {
    const int64_t total = s1.n_pad + s2.n_pad;
    if (total == 0) return 0;
    const int64_t blocks = (total + kPackThreads - 1) / kPackThreads;
    pack_kernel<<<(unsigned)blocks, kPackThreads, 0, stream>>>(s1, s2, min_wh, cell, area_ratio);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

PackSet make_pack_set(const float *boxes, int64_t n, int stride, const int64_t *gather,
                      const float *offsets, void *packed, int64_t n_pad) {
    PackSet s;
    s.segs = nullptr;
    s.nseg = 0;
    s.nseg_dev = nullptr;
    s.n_pad_dev = nullptr;
    s.n_dev = nullptr;
    s.gather32 = nullptr;
    s.offset_ids = nullptr;
    s.mc_classes = 0;
    s.mc_shared = 0;
    s.mc_max = nullptr;
    s.rank_src = nullptr;
    s.rank_out = nullptr;
    s.aabb_out = nullptr;
    s.z_kr = nullptr;
    s.z_summary = nullptr;
    s.z_sum_words = 0;
    s.z_sync = nullptr;
    s.boxes = boxes;
    s.n = n;
    s.stride = stride;
    s.gather = gather;
    s.offsets = offsets;
    s.p0 = reinterpret_cast<float4 *>(packed);
    s.p1 = s.p0 ? s.p0 + n_pad : nullptr;
    s.pr = s.p0 ? reinterpret_cast<uint32_t *>(s.p1 + n_pad) : nullptr;
    s.pt = s.p0 ? s.pr + n_pad : nullptr;
    s.n_pad = n_pad;
    return s;
}

}  // namespace rotgeo
