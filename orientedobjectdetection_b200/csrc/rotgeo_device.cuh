// rotgeo device-side geometry: packed box layout, the centre-distance reject
// test and the exact rotated-rectangle intersection.  sm_100a only.
//
// Replaces (does not translate) mmcv's single_box_iou_rotated
// (box_iou_rotated_utils.hpp; reached by the reference at
// mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:89).  mmcv
// collects up to 24 candidate points, Graham-scans them and sums a triangle
// fan — variable-length local arrays and an in-thread sort.  Here the
// intersection area comes from Green's theorem over the CLIPPED EDGES of both
// rectangles (Sutherland-Hodgman's half-plane clipping applied per edge in
// Liang-Barsky form): fixed work, registers only, no sort, no local memory.
#pragma once
#include <stdio.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rotgeo {

constexpr int kPackPad = 128;  // packed sets are padded to a multiple of this many boxes

__host__ __device__ inline int64_t padded_count(int64_t n) {
    return (n + kPackPad - 1) / kPackPad * kPackPad;
}

// A packed set of n_pad slots (n_pad a multiple of 128) is four planes, back to back:
//   plane0 float4 (cx, cy, r, area)   r = circumscribed radius (inflated), area = w*h
//   plane1 float4 (cos, sin, a, b)    a = w/2, b = h/2
//   planeR uint32 the box's occupancy as bin ranges [s, s+n): x and y cyclic over a modular
//                                     grid of 32 bins per axis (n = 0: padding), plus, for
//                                     NMS packs, the range of AREA bins a partner must fall in
//                                     to reach the IoU threshold (pack_ranges below)
//   planeT per block of 128 boxes, kTableWords uint32: the block's INVERTED bin
//          index as a sparse table — T[axis][level j][bin b][word w] has bit c set
//          iff box 32w+c of the block occupies one of the bins b .. b+2^j-1 (mod 32);
//          axes x, y and (NMS packs) log-area, where a box occupies exactly one bin.
// The pair cull reads planeR of the row boxes and planeT of the column block: the
// 128-bit set of columns a row can touch is two table lookups per axis (a range OR)
// and one AND — about one instruction per pair instead of one test per pair.
// planes 0/1 are read only for the pairs that survive.
constexpr int kBlock = 128;                                  // boxes per table block
constexpr int kTableLevels = 5;                              // run lengths 1, 2, 4, 8, 16
constexpr int kTableAxisWords = kTableLevels * 32 * (kBlock / 32);   // 640
constexpr int kTableXYWords = 2 * kTableAxisWords;           // 1280: what the IoU kernel stages (5 KB)
constexpr int kTableWords = 3 * kTableAxisWords;             // 1920 (7680 bytes)
constexpr int kPackedBytesPerBox = 2 * 16 + 4 + kTableWords * 4 / kBlock;   // 96

struct PackedView {
    const float4 *p0;
    const float4 *p1;
    const uint32_t *pr;
    const uint32_t *pt;
};

// n_pad = number of slots per plane (padded_count(n) for the IoU entry points; the
// NMS packs pad every segment to whole blocks)
__host__ __device__ inline PackedView packed_view(const void *base, int64_t n_pad) {
    PackedView v;
    v.p0 = reinterpret_cast<const float4 *>(base);
    v.p1 = v.p0 + n_pad;
    v.pr = reinterpret_cast<const uint32_t *>(v.p1 + n_pad);
    v.pt = v.pr + n_pad;
    return v;
}

// planeR: sx:5 | nx:6 | sy:5 | ny:6 | sa:5 | na:5  (na = 0: no area constraint)
__host__ __device__ inline uint32_t pack_ranges(int sx, int nx, int sy, int ny, int sa, int na) {
    return (uint32_t)sx | ((uint32_t)nx << 5) | ((uint32_t)sy << 11) | ((uint32_t)ny << 16) |
           ((uint32_t)sa << 22) | ((uint32_t)na << 27);
}
struct BinRanges {
    int sx, nx, sy, ny, sa, na;
};
__host__ __device__ inline BinRanges unpack_ranges(uint32_t rr) {
    BinRanges r;
    r.sx = rr & 31; r.nx = (rr >> 5) & 63; r.sy = (rr >> 11) & 31; r.ny = (rr >> 16) & 63;
    r.sa = (rr >> 22) & 31; r.na = rr >> 27;
    return r;
}

// ---- cull level 1: modular bin masks ------------------------------------------
// The interval [lo, hi] covers the bins floor(lo/cell) .. floor(hi/cell); bin k is
// bit (k mod 32).  Two intervals that share a point share a bin, hence a bit —
// whatever the bin width and wherever the boxes are (class-offset coordinates
// included), so the test `(xa & xb) && (ya & yb)` never rejects a touching pair.
// Intervals spanning >= 32 bins, or so far out that fp32 no longer resolves bins,
// occupy all 32.
// cyclic bin range of [lo, hi]: start bin (0..31) and length (1..32)
__device__ __forceinline__ void bin_range(float lo, float hi, float inv_cell, int &start, int &len) {
    const float fl = floorf(lo * inv_cell), fh = floorf(hi * inv_cell);
    const float span = fh - fl;
    if (!(span < 31.f) || !(fabsf(fl) < 8388608.f)) {  // also NaN / inf
        start = 0;
        len = 32;
    } else {
        start = (int)fl & 31;
        len = (int)span + 1;
    }
}
__device__ __forceinline__ uint32_t range_mask(int start, int len) {
    const uint32_t m = len >= 32 ? 0xffffffffu : ((1u << len) - 1u);
    return __funnelshift_l(m, m, start);
}

// circumscribed radius from the half extents, inflated so that every later use
// (bin masks, circle test) stays conservative under fp32 rounding
__device__ __forceinline__ float circum_radius(float a, float b) {
    return sqrtf(fmaf(a, a, b * b)) * 1.000001f + 1e-30f;
}

// cull level 2: true when the circumscribed circles are disjoint, so the
// rectangles cannot touch and the reference value is exactly 0.
__device__ __forceinline__ bool circles_disjoint(float ax, float ay, float ar, float bx, float by,
                                                 float br) {
    float dx = ax - bx, dy = ay - by;
    float rs = ar + br;
    return fmaf(dx, dx, dy * dy) > rs * rs;
}

// Upper bound of the intersection area of A and B: in A's frame A is the slab |x| <= aA, |y| <= bA,
// and the intersection lies inside A clipped by the axis-aligned bounding box of B in that frame
// (half extents aB|c| + bB|s| and aB|s| + bB|c| around B's centre).  <= 0: the boxes are separated
// along one of A's axes.  ~25 FP32 instructions against the clipper's ~170.
__device__ __forceinline__ float intersection_upper_bound(float4 A0, float4 A1, float4 B0, float4 B1) {
    const float dx = B0.x - A0.x, dy = B0.y - A0.y;
    const float X = fmaf(dx, A1.x, dy * A1.y), Y = fmaf(dy, A1.x, -dx * A1.y);
    const float c = fabsf(fmaf(A1.x, B1.x, A1.y * B1.y)), s = fabsf(fmaf(B1.y, A1.x, -B1.x * A1.y));
    const float px = fmaf(B1.z, c, B1.w * s), py = fmaf(B1.z, s, B1.w * c);
    const float ox = fminf(A1.z, X + px) - fmaxf(-A1.z, X - px);
    const float oy = fminf(A1.w, Y + py) - fmaxf(-A1.w, Y - py);
    return fmaxf(ox, 0.f) * fmaxf(oy, 0.f);
}

__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
__device__ __forceinline__ float fmin3(float a, float b, float c) { return fminf(fminf(a, b), c); }

// Intersection area of rectangle A and rectangle B.
//
// Frame: B's own (origin at B's centre, axes along B's sides), so B is the
// slab |x|<=aB, |y|<=bB.  A's centre is (X, Y), A's long-axis direction is
// e = (c, s) with c, s forced >= tiny > 0 by two exact symmetries (A rotated by
// pi is the same box; mirroring the frame about B's x axis keeps B).  Exactly
// parallel sides are thereby treated as an infinitesimal rotation — a valid
// nearby configuration, so the area is continuous there (no special cases).
//
// Each of the 16 line/line crossings is evaluated ONCE: its parameter on A's
// side clips A's edge, and the same point re-expressed on B's side clips B's
// edge.  That keeps the clipped boundary a closed polygon even when two sides
// are nearly parallel and the crossing itself is ill-conditioned.
//
// 2*area = sum over the 8 edges of (clipped length) * cross(edge dir, point-origin);
// the origin is put at the centre of the SMALLER box to limit cancellation.
__device__ __forceinline__ float rect_intersection_area(float4 A0, float4 A1, float4 B0,
                                                        float4 B1, float areaA, float areaB) {
    const float aA = A1.z, bA = A1.w, aB = B1.z, bB = B1.w;
    const float dx = A0.x - B0.x, dy = A0.y - B0.y;
    float c = fmaf(A1.x, B1.x, A1.y * B1.y);
    float s = fmaf(A1.y, B1.x, -A1.x * B1.y);
    const float X = fmaf(dx, B1.x, dy * B1.y);
    float Y = fmaf(dy, B1.x, -dx * B1.y);
    if (c < 0.f) { c = -c; s = -s; }
    if (s < 0.f) { s = -s; Y = -Y; }
    c = fmaxf(c, 1e-12f);
    s = fmaxf(s, 1e-12f);
    const float rc = __frcp_rn(c), rs = __frcp_rn(s);
    const float n2 = fmaf(c, c, s * s);
    const bool a_small = areaA < areaB;
    // origin O = lam*(X,Y): lam = 1 -> A's centre, 0 -> B's centre
    const float kE = a_small ? 0.f : fmaf(c, Y, -s * X);
    const float kF = a_small ? 0.f : fmaf(c, X, s * Y);
    const float lX = a_small ? X : 0.f;
    const float lY = a_small ? Y : 0.f;
    const float bAs = bA * s, bAc = bA * c, aAc = aA * c, aAs = aA * s;
    const float wE = bA * n2, wF = aA * n2;

    float acc = 0.f;
    float yE[2][2], xE[2][2], yF[2][2], xF[2][2];  // [sigma: 0 = +, 1 = -][rho: 0 = +, 1 = -]
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float g = k ? -1.f : 1.f;
        // A's two long sides: q + t*(c, s), t in [-aA, aA]
        {
            const float qx = fmaf(-g, bAs, X), qy = fmaf(g, bAc, Y);
            const float txm = (-aB - qx) * rc, txp = (aB - qx) * rc;
            const float tym = (-bB - qy) * rs, typ = (bB - qy) * rs;
            const float lo = fmax3(-aA, txm, tym), hi = fmin3(aA, txp, typ);
            acc = fmaf(fmaxf(hi - lo, 0.f), fmaf(g, kE, wE), acc);
            yE[k][1] = fmaf(txm, s, qy);
            yE[k][0] = fmaf(txp, s, qy);
            xE[k][1] = fmaf(tym, c, qx);
            xE[k][0] = fmaf(typ, c, qx);
        }
        // A's two short sides: p + t*(-s, c), t in [-bA, bA]
        {
            const float px = fmaf(g, aAc, X), py = fmaf(g, aAs, Y);
            const float txp = (px - aB) * rs, txm = (px + aB) * rs;
            const float tym = (-bB - py) * rc, typ = (bB - py) * rc;
            const float lo = fmax3(-bA, txp, tym), hi = fmin3(bA, txm, typ);
            acc = fmaf(fmaxf(hi - lo, 0.f), fmaf(g, kF, wF), acc);
            yF[k][0] = fmaf(txp, c, py);
            yF[k][1] = fmaf(txm, c, py);
            xF[k][1] = fmaf(-tym, s, px);
            xF[k][0] = fmaf(-typ, s, px);
        }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const float g = r ? -1.f : 1.f;
        // B's side x = g*aB, parameter y in [-bB, bB]
        {
            const float lo = fmax3(-bB, yE[1][r], yF[1][r]);
            const float hi = fmin3(bB, yE[0][r], yF[0][r]);
            acc = fmaf(fmaxf(hi - lo, 0.f), fmaf(-g, lX, aB), acc);
        }
        // B's side y = g*bB, parameter x in [-aB, aB]
        {
            const float lo = fmax3(-aB, xE[0][r], xF[1][r]);
            const float hi = fmin3(aB, xE[1][r], xF[0][r]);
            acc = fmaf(fmaxf(hi - lo, 0.f), fmaf(-g, lY, bB), acc);
        }
    }
    float inter = fmaxf(0.5f * acc, 0.f);
    return fminf(inter, fminf(areaA, areaB));
}

// IoU (mode 0) or IoF (mode 1, divide by A's area).  Areas below 1e-14 give 0
// like the reference's native code (SURVEY.md §8(c) Spec A step 2).
template <int MODE>
__device__ __forceinline__ float rect_overlap(float4 A0, float4 A1, float4 B0, float4 B1) {
    const float areaA = A0.w, areaB = B0.w;
    const float inter = rect_intersection_area(A0, A1, B0, B1, areaA, areaB);
    const float base = MODE == 0 ? (areaA + areaB - inter) : areaA;
    const float v = __fdividef(inter, base);
    return (areaA < 1e-14f || areaB < 1e-14f) ? 0.f : v;
}

// level-2 cull + exact value of one pair that survived the bin masks
template <int MODE>
__device__ __forceinline__ float pair_value(float4 A0, float4 A1, float4 B0, float4 B1) {
    if (circles_disjoint(A0.x, A0.y, A0.z, B0.x, B0.y, B0.z)) return 0.f;
    return rect_overlap<MODE>(A0, A1, B0, B1);
}

// ---- TMA (bulk async copy) + mbarrier helpers, raw PTX -----------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
#ifdef ROTGEO_DEBUG_WAIT
// probe builds: a wait that does not end reports where it hangs and stops the kernel
__device__ __forceinline__ bool mbar_try(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
#define mbar_wait(bar, parity)                                                                      \
    do {                                                                                            \
        unsigned long long spins_ = 0;                                                              \
        while (!mbar_try((bar), (parity))) {                                                        \
            if (++spins_ > 2000000ull) {                                                            \
                printf("stuck line %d block %d thread %d bar %u parity %u\n", __LINE__, blockIdx.x, \
                       threadIdx.x, smem_u32(bar) & 0xffffu, (unsigned)(parity));                   \
                __trap();                                                                           \
            }                                                                                       \
        }                                                                                           \
    } while (0)
#else
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
#endif
// 1-D bulk copy global -> shared through the TMA unit (SASS: UBLKCP); bytes and
// both addresses must be multiples of 16.
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- programmatic dependent launch ---------------------------------------------------
// A kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still running (once every CTA of the
// predecessor has called pdl_launch_dependents or exited); pdl_wait blocks until the
// predecessor has completed and its writes are visible.  Both are no-ops otherwise.
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- cluster launch control (sm_100 work stealing) ------------------------------
// A running CTA asks the hardware scheduler to cancel the launch of a CTA of the
// same grid that has not started yet and takes over its blockIdx: persistent CTAs
// with the hardware's own dynamic load balancing and no global work counter.
// The 16-byte response lands in shared memory and completes `bar` (16 tx bytes).
__device__ __forceinline__ void clc_try_cancel(void *resp_smem, uint64_t *bar) {
    asm volatile(
        "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];"
        ::"r"(smem_u32(resp_smem)), "r"(smem_u32(bar))
        : "memory");
}
// true + the cancelled CTA's blockIdx.x, or false when nothing was left to cancel
__device__ __forceinline__ bool clc_decode(const void *resp_smem, unsigned &ctaid_x) {
    unsigned ok, x = 0;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b128 r;\n"
        "ld.shared.b128 r, [%2];\n"
        "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "@p clusterlaunchcontrol.query_cancel.get_first_ctaid::x.b32.b128 %1, r;\n"
        "}\n"
        : "=r"(ok), "+r"(x)
        : "r"(smem_u32(resp_smem))
        : "memory");
    ctaid_x = x;
    return ok != 0;
}
// orders this thread's generic-proxy reads of shared memory before later
// async-proxy writes to it (a response slot about to be reused)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

}  // namespace rotgeo
