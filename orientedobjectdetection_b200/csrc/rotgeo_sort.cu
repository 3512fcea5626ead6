// NMS front-end: the orders and the segment table of a rotated NMS, made by ONE kernel.  sm_100a.
//
// Replaces what the reference does in python / ATen around the native op — the score filter and
// nonzero() of multiclass_nms_rotated (mmrotate/core/post_processing/bbox_nms_rotated.py:49-58), the
// two global max reductions and the centre offsets (:73-77; mmcv.ops.batched_nms), scores.sort() +
// index_select inside mmcv.ops.nms_rotated, the per-image python loop of the RPN heads
// (mmrotate/models/dense_heads/rotated_rpn_head.py:414-426,515, oriented_rpn_head.py:275) — and what
// round 1 of this package did with 5-7 library radix-sort passes, a run-length encode and a host
// read of the segment sizes.
//
// Unit of work: a GROUP = a contiguous range of candidate rows whose kept boxes are emitted
// together in descending score order (a whole nms_rotated / batched_nms / multiclass call, one image
// of an RPN batch, one class of a patch merge).  Inside a group, boxes of equal id form a SEGMENT
// (class, FPN level): suppression never crosses segments.  One thread-block CLUSTER (1, 2, 4 or 8
// CTAs, 12 288 elements each) owns a group and keeps it in (distributed) shared memory from the
// first load to the last store:
//
//   load     32-bit key = descending image of the score (multiclass: candidate filter on the raw
//            score, stable compaction of the candidates across the cluster), extremes of the ids,
//            max(cx, cy) / max(w, h) of the offset formula, all reduced across the cluster
//   sort 1   4 LSD radix passes (8 bits) over the keys, payload = index: stable, so equal scores
//            keep their index order.  Per pass: warp-level ranking with match.any (no atomics),
//            per-CTA digit histogram, histograms exchanged through DISTRIBUTED SHARED MEMORY,
//            scatter straight into the owning CTA's shared memory (st.shared::cluster), two
//            cluster barriers.  -> `order` (emission order)
//   sort 2   1-2 more stable passes on (id - min id) with payload = score position: segment-major
//            order, score order inside a segment.  -> `seg_order`, `seg_rank`, segment starts
//   plan     the cluster that finishes last (atomic ticket) turns the segment sizes of ALL groups
//            into the segment table, tile / mask / slot totals the host used to compute after a
//            device->host read (nms_plan() in rotgeo_nms.cu), and leaves them in a device header.
//
// Groups beyond one cluster's capacity (98 304 elements) take the same code with the element
// buffers and the histogram exchange in global memory and a cooperative grid barrier in place of
// the cluster barrier (template parameter GLOBAL).
#include <cooperative_groups.h>

#include <vector>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace cg = cooperative_groups;

namespace rotgeo {

namespace {

constexpr int kThreads = kSortThreads;          // 1024
constexpr int kWarps = kThreads / 32;
constexpr int kItems = kSortItems;              // 12
constexpr int kCap = kSortCap;                  // elements per CTA
constexpr int kTile = 128;

// float -> unsigned whose integer order is the float order
__device__ __forceinline__ unsigned int ordered_bits(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_float(unsigned int o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ unsigned long long id_image(int64_t id) {   // order-preserving
    return (unsigned long long)id ^ 0x8000000000000000ull;
}

// group-wide reductions, all expressed as maxima (minima through the complemented image)
struct Red {
    unsigned long long id_max, id_min_inv;
    unsigned int max_xy, max_wh;
    unsigned int count;      // elements (multiclass: candidates) of this CTA
    unsigned int pad;
};

struct __align__(16) SortSmem {
    unsigned short cnt[kWarps][256];    // per-warp digit counters, then exclusive prefixes over the warps
    unsigned int hist[256];             // this CTA's digit histogram (read by the whole cluster)
    unsigned int tot[256];              // digit totals of the group
    unsigned int dbase[256];            // first destination of (digit, this CTA)
    unsigned int wsum[kWarps];
    Red red;
    unsigned int bcount;                // segment boundaries in this CTA
    int skip;                           // every key of the group has the same digit: nothing moves
    int last;                           // this cluster took the last ticket: it makes the plan
    long long carry[4];
};

// ---- team = the CTAs that own one group: a cluster (shared memory) or the cooperative grid (global) ----
template <bool GLOBAL>
struct Team {
    int rank, size;
    SortSmem *sm;
    // element buffers: CLUSTER -> this CTA's shared arrays of kCap entries; GLOBAL -> global arrays
    // of the whole group (index = position)
    uint32_t *key[2], *val[2];
    unsigned int *ghist;     // GLOBAL: [size][256]
    Red *gred;               // GLOBAL: [size]
    unsigned int *gcount;    // GLOBAL: [size] boundary counts

    __device__ __forceinline__ void sync() const {
        if (GLOBAL) {
            __threadfence();
            cg::this_grid().sync();
        } else {
            cg::this_cluster().sync();
        }
    }
    // element at local position lp of this CTA (share = positions per CTA)
    __device__ __forceinline__ uint32_t ld(const uint32_t *buf, uint32_t share, uint32_t lp) const {
        return GLOBAL ? __ldcg(buf + (size_t)rank * share + lp) : buf[lp];
    }
    __device__ __forceinline__ void st_local(uint32_t *buf, uint32_t share, uint32_t lp, uint32_t v) const {
        if (GLOBAL) __stcg(buf + (size_t)rank * share + lp, v); else buf[lp] = v;
    }
    // element at group position p (lands in the shared memory of CTA p / share)
    __device__ __forceinline__ void st_pos(uint32_t *buf, uint32_t share, uint32_t p, uint32_t v) const {
        if (GLOBAL) {
            __stcg(buf + p, v);
        } else {
            const uint32_t c = p / share;
            uint32_t *dst = cg::this_cluster().map_shared_rank(buf, c);
            dst[p - c * share] = v;
        }
    }
    __device__ __forceinline__ uint32_t ld_pos(const uint32_t *buf, uint32_t share, uint32_t p) const {
        if (GLOBAL) return __ldcg(buf + p);
        const uint32_t c = p / share;
        const uint32_t *src = cg::this_cluster().map_shared_rank(const_cast<uint32_t *>(buf), c);
        return src[p - c * share];
    }
    __device__ __forceinline__ void publish_hist(int d, unsigned int v) const {
        sm->hist[d] = v;
        if (GLOBAL) __stcg(ghist + (size_t)rank * 256 + d, v);
    }
    __device__ __forceinline__ unsigned int peer_hist(int c, int d) const {
        if (GLOBAL) return __ldcg(ghist + (size_t)c * 256 + d);
        return cg::this_cluster().map_shared_rank(sm->hist, c)[d];
    }
    __device__ __forceinline__ void publish_red() const {
        if (GLOBAL) {
            Red r = sm->red;
            gred[rank] = r;
        }
    }
    __device__ __forceinline__ Red peer_red(int c) const {
        if (GLOBAL) {
            Red r;
            const volatile Red *p = gred + c;
            r.id_max = p->id_max; r.id_min_inv = p->id_min_inv; r.max_xy = p->max_xy; r.max_wh = p->max_wh;
            r.count = p->count; r.pad = 0;
            return r;
        }
        return *cg::this_cluster().map_shared_rank(&sm->red, c);
    }
    __device__ __forceinline__ void publish_bcount(unsigned int v) const {
        sm->bcount = v;
        if (GLOBAL) __stcg(gcount + rank, v);
    }
    __device__ __forceinline__ unsigned int peer_bcount(int c) const {
        if (GLOBAL) return __ldcg(gcount + c);
        return *cg::this_cluster().map_shared_rank(&sm->bcount, c);
    }
};

__device__ __forceinline__ uint32_t share_of(uint32_t n, int size) {
    uint32_t s = (n + (uint32_t)size - 1u) / (uint32_t)size;
    s = (s + 31u) & ~31u;
    return s < 32u ? 32u : s;
}
__device__ __forceinline__ uint32_t count_of(uint32_t n, uint32_t share, int rank) {
    const uint64_t lo = (uint64_t)rank * share;
    return lo >= n ? 0u : (n - lo < share ? (uint32_t)(n - lo) : share);
}

// exclusive scan of one value per thread over the block; returns the exclusive prefix, total in `total`
__device__ __forceinline__ long long block_exclusive_scan(long long v, long long *s_warp /* kWarps */, long long &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const long long t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    __syncthreads();
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    long long before = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        const long long c = s_warp[w];
        before += w < warp ? c : 0;
        tot += c;
    }
    total = tot;
    return before + incl - v;
}

// One stable LSD pass over the 8 key bits at `shift`: the n elements of the group move from buffer
// `cur` to buffer `cur ^ 1`.  Position order inside a CTA = (warp, item, lane), iu items per thread.
// Returns true when the buffers were swapped (false: all keys share the digit, nothing moved).
template <bool GLOBAL>
__device__ bool radix_pass(const Team<GLOBAL> &T, int cur, int shift, uint32_t n, uint32_t share) {
    SortSmem &sm = *T.sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t count = count_of(n, share, T.rank);
    const int iu = (int)((share + kThreads - 1) / kThreads);
    const uint32_t lt = (1u << lane) - 1u;
    // 1. zero the counters
    {
        uint4 *z = reinterpret_cast<uint4 *>(&sm.cnt[0][0]);
        z[tid] = make_uint4(0u, 0u, 0u, 0u);      // 32 * 256 * 2 B = 1024 * 16 B
        if (tid == 0) sm.skip = 0;
    }
    __syncthreads();
    // 2. rank every element among the elements of its warp with the same digit
    uint32_t dr[kItems];       // digit << 16 | rank inside (warp, digit); 0xffffffff = no element
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        dr[i] = 0xffffffffu;
        if (i < iu) {                                  // warp-uniform
            const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
            const bool active = lp < count;
            const uint32_t d = active ? ((T.ld(T.key[cur], share, lp) >> shift) & 0xffu) : (0x10000u + lane);
            const uint32_t peers = __match_any_sync(0xffffffffu, d);
            unsigned int base = 0;
            if (active) base = sm.cnt[warp][d];
            __syncwarp();
            if (active && (peers & lt) == 0u) sm.cnt[warp][d] = (unsigned short)(base + __popc(peers));
            __syncwarp();
            if (active) dr[i] = (d << 16) | (base + __popc(peers & lt));
        }
    }
    __syncthreads();
    // 3. exclusive prefix over the warps, CTA histogram
    if (tid < 256) {
        unsigned int run = 0;
#pragma unroll 8
        for (int w = 0; w < kWarps; ++w) {
            const unsigned int c = sm.cnt[w][tid];
            sm.cnt[w][tid] = (unsigned short)run;
            run += c;
        }
        T.publish_hist(tid, run);
    }
    T.sync();
    // 4. totals of the group and what lies before this CTA
    unsigned int tot = 0, before = 0;
    if (tid < 256) {
        for (int c = 0; c < T.size; ++c) {
            const unsigned int h = T.peer_hist(c, tid);
            tot += h;
            before += c < T.rank ? h : 0u;
        }
        if (tot == n) sm.skip = 1;
    }
    {
        // exclusive scan of the 256 digit totals (warps 0..7)
        unsigned int incl = tot;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (tid < 256 && lane == 31) sm.wsum[warp] = incl;
        __syncthreads();
        if (tid < 256) {
            unsigned int pre = 0;
            for (int w = 0; w < warp; ++w) pre += sm.wsum[w];
            sm.dbase[tid] = pre + incl - tot + before;
        }
    }
    __syncthreads();
    const bool moved = sm.skip == 0;
    // 5. scatter
    if (moved) {
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            if (i < iu && dr[i] != 0xffffffffu) {
                const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                const uint32_t d = dr[i] >> 16, r = dr[i] & 0xffffu;
                const uint32_t p = sm.dbase[d] + sm.cnt[warp][d] + r;
                T.st_pos(T.key[cur ^ 1], share, p, T.ld(T.key[cur], share, lp));
                T.st_pos(T.val[cur ^ 1], share, p, T.ld(T.val[cur], share, lp));
            }
        }
    }
    T.sync();     // also: nobody overwrites its histogram while a peer still reads it
    return moved;
}

__device__ __forceinline__ int64_t group_start(const SortArgs &a, int g) {
    return a.gstart ? a.gstart[g] : a.gstart_inline[g];
}

// the plan nms_plan() makes on the host, made by the last CTA to finish (1024 threads)
__device__ void make_plan(const SortArgs &a, SortSmem &sm) {
    const int tid = threadIdx.x;
    __shared__ long long s_warp[kWarps];
    __shared__ int s_maxc[kWarps];
    // segments before each group
    long long carry = 0;
    for (int base = 0; base < a.ngroups; base += kThreads) {
        const int g = base + tid;
        long long v = 0;
        if (g < a.ngroups) v = __ldcg(a.gseg_count + g);
        long long tot;
        const long long ex = block_exclusive_scan(v, s_warp, tot);
        if (g < a.ngroups) a.gseg_off[g] = (int)(carry + ex);
        carry += tot;
    }
    const long long T = carry;
    __threadfence_block();
    __syncthreads();
    unsigned int err = 0;
    if (T > a.segs_cap) err |= kSortErrSegments;
    long long slots = 0, words = 0, tiles = 0;
    int maxc = 0;
    const long long Tc = T > a.segs_cap ? a.segs_cap : T;
    for (long long base = 0; base < Tc; base += kThreads) {
        const long long j = base + tid;
        long long len = 0, C = 0, start = 0;
        if (j < Tc) {
            int lo = 0, hi = a.ngroups - 1;     // last group with gseg_off <= j (groups without segments share offsets)
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (a.gseg_off[mid] <= j) lo = mid; else hi = mid - 1;
            }
            const int k = (int)(j - a.gseg_off[lo]);
            const uint32_t *gs = a.gseg_start + (size_t)lo * (a.seg_cap + 1);
            const long long s0 = __ldcg(gs + k), s1 = __ldcg(gs + k + 1);
            len = s1 - s0;
            C = (len + kTile - 1) / kTile;
            start = (a.mode == 1 ? 0 : group_start(a, lo)) + s0;
        }
        long long t1, t2, t3;
        const long long e1 = block_exclusive_scan(C * kTile, s_warp, t1);
        const long long e2 = block_exclusive_scan(len * 2 * C, s_warp, t2);
        const long long e3 = block_exclusive_scan(C * (C + 1) / 2, s_warp, t3);
        if (j < Tc) {
            NmsSegment &sg = a.segs[j];
            sg.start = start;
            sg.pstart = slots + e1;
            sg.len = len;
            sg.mask_base = words + e2;
            sg.tile_base = tiles + e3;
            sg.words = (int32_t)(2 * C);
            sg.row_blocks = (int32_t)((len + 63) / 64);
        }
        slots += t1; words += t2; tiles += t3;
        maxc = max(maxc, (int)C);
    }
    // largest segment
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) maxc = max(maxc, __shfl_xor_sync(0xffffffffu, maxc, d));
    __syncthreads();
    if ((tid & 31) == 0) s_maxc[tid >> 5] = maxc;
    __syncthreads();
    if (tid == 0) {
        for (int w = 0; w < kWarps; ++w) maxc = max(maxc, s_maxc[w]);
        SortHeader *h = a.hdr;
        h->nseg = Tc;
        h->n_pad = slots > 0 ? slots : kTile;
        h->total_tiles = tiles;
        h->mask_words = words;
        h->max_blocks = maxc;
        if (err) atomicOr(&h->error, err);
    }
}

// everything one team does for one group
template <bool GLOBAL>
__device__ void sort_group(const SortArgs &a, int g, Team<GLOBAL> &T) {
    SortSmem &sm = *T.sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    const int64_t g0 = group_start(a, g);
    const uint32_t n_in = (uint32_t)(group_start(a, g + 1) - g0);    // rows (mode 1: flat candidates)
    const bool mc = a.mode == 1;

    if (tid == 0) {
        sm.red.id_max = 0ull; sm.red.id_min_inv = 0ull; sm.red.max_xy = 0u; sm.red.max_wh = 0u;
        sm.red.count = 0u; sm.red.pad = 0u;
        sm.bcount = 0u;
        sm.last = 0;
    }
    __syncthreads();

    // ---------------------------------------------------------------- load
    uint32_t share = share_of(n_in, T.size);
    uint32_t n = n_in;
    int cur = 0;
    {
        const uint32_t count = count_of(n_in, share, T.rank);
        const int iu = (int)((share + kThreads - 1) / kThreads);
        unsigned long long idmx = 0ull, idmn = 0ull;
        unsigned int mxy = 0u, mwh = 0u;
        uint32_t kv[kItems];          // keys of my elements (multiclass: 0xffffffff = not a candidate)
        unsigned int nvalid = 0;
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            kv[i] = 0xffffffffu;
            if (i < iu) {
                const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                if (lp < count) {
                    const uint32_t q = (uint32_t)T.rank * share + lp;       // index inside the group
                    float sc;
                    bool valid = true;
                    const float *b;
                    if (mc) {
                        const uint32_t row = q / (uint32_t)a.classes, cls = q - row * (uint32_t)a.classes;
                        const float raw = a.scores[(size_t)row * (a.classes + 1) + cls];
                        valid = raw > a.score_thr;               // tested BEFORE the factors, like the reference
                        sc = raw;
                        if (valid && a.factors) sc *= a.factors[row];
                        b = a.boxes + (size_t)(a.shared ? row : q) * 5;
                    } else {
                        sc = a.scores[g0 + q];
                        b = a.boxes + (size_t)(g0 + q) * a.stride;
                    }
                    if (valid) {
                        // descending score: complement of the order-preserving image (+0.0f: -0 -> +0);
                        // equal scores keep their index order because every pass is stable
                        uint32_t key = 0xffffffffu - ordered_bits(sc + 0.0f);
                        if (mc && key == 0xffffffffu) key = 0xfffffffeu;     // -NaN payloads only
                        kv[i] = key;
                        ++nvalid;
                        if (mc || a.ids) {
                            mxy = max(mxy, ordered_bits(fmaxf(b[0], b[1])));
                            mwh = max(mwh, ordered_bits(fmaxf(b[2], b[3])));
                        }
                        if (!mc && a.ids) {
                            const unsigned long long u = id_image(a.ids[g0 + q]);
                            idmx = max(idmx, u);
                            idmn = max(idmn, ~u);
                        }
                    }
                    if (!mc) {
                        T.st_local(T.key[0], share, lp, kv[i]);
                        T.st_local(T.val[0], share, lp, q);
                    }
                }
            }
        }
        // CTA reductions
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            idmx = max(idmx, __shfl_xor_sync(0xffffffffu, idmx, d));
            idmn = max(idmn, __shfl_xor_sync(0xffffffffu, idmn, d));
            mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, d));
            mwh = max(mwh, __shfl_xor_sync(0xffffffffu, mwh, d));
        }
        if (lane == 0) {
            atomicMax(&sm.red.id_max, idmx);
            atomicMax(&sm.red.id_min_inv, idmn);
            atomicMax(&sm.red.max_xy, mxy);
            atomicMax(&sm.red.max_wh, mwh);
        }
        if (mc) {
            // stable compaction of the candidates: rank inside the CTA in position order
            // (per item a ballot; per-warp running count; prefix over the warps)
            unsigned int run = 0;
            uint32_t rk[kItems];
#pragma unroll
            for (int i = 0; i < kItems; ++i) {
                rk[i] = 0u;
                if (i < iu) {
                    const bool v = kv[i] != 0xffffffffu;
                    const uint32_t bal = __ballot_sync(0xffffffffu, v);
                    rk[i] = run + __popc(bal & lt);
                    run += __popc(bal);
                }
            }
            if (lane == 0) sm.wsum[warp] = run;
            __syncthreads();
            unsigned int wbefore = 0, ctot = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const unsigned int c = sm.wsum[w];
                wbefore += w < warp ? c : 0u;
                ctot += c;
            }
            if (tid == 0) sm.red.count = ctot;
            __syncthreads();
            T.publish_red();
            T.sync();
            unsigned int before = 0, total = 0;
            for (int c = 0; c < T.size; ++c) {
                const unsigned int h = T.peer_red(c).count;
                total += h;
                before += c < T.rank ? h : 0u;
            }
            n = total;
            const uint32_t share2 = share_of(n, T.size);
#pragma unroll
            for (int i = 0; i < kItems; ++i) {
                if (i < iu && kv[i] != 0xffffffffu) {
                    const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                    const uint32_t q = (uint32_t)T.rank * share + lp;
                    const uint32_t p = before + wbefore + rk[i];
                    T.st_pos(T.key[0], share2, p, kv[i]);
                    T.st_pos(T.val[0], share2, p, q);
                }
            }
            share = share2;
            T.sync();
        } else {
            __syncthreads();
            T.publish_red();
            T.sync();
        }
        (void)nvalid;
    }
    // group-wide reductions
    Red R;
    {
        R = T.peer_red(0);
        for (int c = 1; c < T.size; ++c) {
            const Red o = T.peer_red(c);
            R.id_max = max(R.id_max, o.id_max);
            R.id_min_inv = max(R.id_min_inv, o.id_min_inv);
            R.max_xy = max(R.max_xy, o.max_xy);
            R.max_wh = max(R.max_wh, o.max_wh);
        }
    }
    const bool has_ids = mc || a.ids != nullptr;
    const unsigned long long id_lo = ~R.id_min_inv;
    unsigned long long id_range = mc ? (unsigned long long)a.classes : (a.ids && n > 0 ? R.id_max - id_lo + 1ull : 1ull);
    bool range_bad = false;
    if (id_range > 65536ull) {
        range_bad = true;          // the caller renumbers the ids and splits the job (ROTGEO_ERR_RANGE)
        id_range = 1ull;
    }
    const float off_unit = __fadd_rn(__fadd_rn(ordered_float(R.max_xy), ordered_float(R.max_wh)), 1.0f);

    // ---------------------------------------------------------------- sort 1: score order
    for (int shift = 0; shift < 32; shift += 8)
        if (radix_pass(T, cur, shift, n, share)) cur ^= 1;

    const uint32_t count = count_of(n, share, T.rank);
    const int iu = (int)((share + kThreads - 1) / kThreads);
    const int64_t obase = mc ? 0 : g0;       // `order` holds global rows (multiclass: flat candidate indices)
    // -> order; sort-2 keys in place: key = id - min id, val = position in the score order
    const bool two = has_ids && !range_bad && id_range > 1ull;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        if (i < iu) {
            const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
            if (lp < count) {
                const uint32_t p = (uint32_t)T.rank * share + lp;
                const uint32_t idx = T.ld(T.val[cur], share, lp);
                const uint32_t src = (uint32_t)(obase + idx);
                a.order[obase + p] = src;
                uint32_t idk = 0u;
                if (mc) {
                    idk = idx % (uint32_t)a.classes;
                } else if (a.ids) {
                    const int64_t id = a.ids[g0 + idx];
                    idk = range_bad ? 0u : (uint32_t)(id_image(id) - id_lo);
                    // mmcv.ops.batched_nms: float(id) * (max(cx,cy) + max(w,h) + 1), separate roundings
                    if (a.id_offsets) a.offsets_out[g0 + idx] = __fmul_rn((float)id, off_unit);
                }
                if (two) {
                    T.st_local(T.key[cur], share, lp, idk);
                    T.st_local(T.val[cur], share, lp, p);
                } else {
                    a.seg_order[obase + p] = src;
                    a.seg_rank[obase + p] = p;
                }
            }
        }
    }
    __threadfence();
    T.sync();

    // ---------------------------------------------------------------- sort 2: segment-major order
    unsigned int nseg_here = n > 0 ? 1u : 0u;
    if (two) {
        if (radix_pass(T, cur, 0, n, share)) cur ^= 1;
        if (id_range > 256ull)
            if (radix_pass(T, cur, 8, n, share)) cur ^= 1;
        // seg_order / seg_rank, segment boundaries (count, then write)
        uint32_t bmask = 0u;      // bit i: my item i starts a segment
        unsigned int run = 0;
        uint32_t ordl[kItems];
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            ordl[i] = 0u;
            if (i < iu) {
                const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                bool isb = false;
                if (lp < count) {
                    const uint32_t q = (uint32_t)T.rank * share + lp;
                    const uint32_t p = T.ld(T.val[cur], share, lp);
                    const uint32_t k = T.ld(T.key[cur], share, lp);
                    a.seg_rank[obase + q] = p;
                    a.seg_order[obase + q] = __ldcg(a.order + obase + p);
                    isb = q == 0u || T.ld_pos(T.key[cur], share, q - 1u) != k;
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, isb);
                if (isb) bmask |= 1u << i;
                ordl[i] = run + __popc(bal & lt);
                run += __popc(bal);
            }
        }
        if (lane == 0) sm.wsum[warp] = run;
        __syncthreads();
        unsigned int wbefore = 0, ctot = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const unsigned int c = sm.wsum[w];
            wbefore += w < warp ? c : 0u;
            ctot += c;
        }
        __syncthreads();
        if (tid == 0) T.publish_bcount(ctot);
        T.sync();
        unsigned int before = 0, total = 0;
        for (int c = 0; c < T.size; ++c) {
            const unsigned int h = T.peer_bcount(c);
            total += h;
            before += c < T.rank ? h : 0u;
        }
        nseg_here = total;
        uint32_t *gs = a.gseg_start + (size_t)g * (a.seg_cap + 1);
        if (total <= (unsigned int)a.seg_cap) {
#pragma unroll
            for (int i = 0; i < kItems; ++i) {
                if (i < iu && (bmask >> i) & 1u) {
                    const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                    gs[before + wbefore + ordl[i]] = (uint32_t)T.rank * share + lp;
                }
            }
        }
    } else if (T.rank == 0 && tid == 0 && n > 0) {
        a.gseg_start[(size_t)g * (a.seg_cap + 1)] = 0u;
    }
    // ---------------------------------------------------------------- group summary, ticket, plan
    if (T.rank == 0 && tid == 0) {
        unsigned int err = range_bad ? kSortErrRange : 0u;
        unsigned int s = nseg_here;
        if (s > (unsigned int)a.seg_cap) {
            err |= kSortErrSegments;
            s = 0u;
        }
        a.gseg_start[(size_t)g * (a.seg_cap + 1) + s] = n;      // end of the last segment
        a.gseg_count[g] = (int)s;
        if (err) atomicOr(&a.hdr->error, err);
        atomicAdd(reinterpret_cast<unsigned long long *>(&a.hdr->n_valid), (unsigned long long)n);
        if (mc) {
            a.hdr->mc_max[0] = ordered_float(R.max_xy);
            a.hdr->mc_max[1] = ordered_float(R.max_wh);
        }
    }
    __threadfence();
    T.sync();       // every CTA of the team has written its share
    if (T.rank == 0) {
        if (tid == 0) {
            __threadfence();
            sm.last = atomicAdd(&a.hdr->done, 1u) == (unsigned int)a.ngroups - 1u ? 1 : 0;
        }
        __syncthreads();
        if (sm.last) {
            __threadfence();
            make_plan(a, sm);
        }
    }
}

}  // namespace

// ---- kernels ---------------------------------------------------------------------------------------
// one cluster per group; groups too large for the cluster are left to sort_global_kernel
__global__ void __launch_bounds__(kSortThreads, 1)
sort_cluster_kernel(const SortArgs a) {
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ SortSmem sm;
    cg::cluster_group cluster = cg::this_cluster();
    const int K = (int)cluster.num_blocks();
    const int g = (int)(blockIdx.x / K);
    const int64_t n_in = group_start(a, g + 1) - group_start(a, g);
    if (n_in > (int64_t)K * kCap) return;            // whole cluster returns: nobody waits at a barrier
    Team<false> T;
    T.rank = (int)cluster.block_rank();
    T.size = K;
    T.sm = &sm;
    uint32_t *buf = reinterpret_cast<uint32_t *>(dyn);
    T.key[0] = buf;
    T.val[0] = buf + kCap;
    T.key[1] = buf + 2 * kCap;
    T.val[1] = buf + 3 * kCap;
    T.ghist = nullptr; T.gred = nullptr; T.gcount = nullptr;
    cluster.sync();        // every CTA of the cluster has started: its shared memory may be written
    sort_group<false>(a, g, T);
    cluster.sync();        // nobody exits while a peer may still read its shared memory
}

// cooperative grid: the groups the clusters skipped, one after the other
__global__ void __launch_bounds__(kSortThreads, 1)
sort_global_kernel(const SortArgs a, int K_cluster) {
    __shared__ SortSmem sm;
    for (int g = 0; g < a.ngroups; ++g) {
        const int64_t n_in = group_start(a, g + 1) - group_start(a, g);
        if (n_in <= (int64_t)K_cluster * kCap) continue;
        Team<true> T;
        T.rank = (int)blockIdx.x;
        T.size = (int)gridDim.x;
        T.sm = &sm;
        T.key[0] = a.big_buf;
        T.val[0] = a.big_buf + a.big_cap;
        T.key[1] = a.big_buf + 2 * a.big_cap;
        T.val[1] = a.big_buf + 3 * a.big_cap;
        T.ghist = a.big_hist;
        T.gred = reinterpret_cast<Red *>(a.big_red);
        T.gcount = a.big_count;
        sort_group<true>(a, g, T);
        T.sync();
    }
}

// ---- host side --------------------------------------------------------------------------------------
static inline size_t align_up(size_t x, size_t al) { return (x + al - 1) / al * al; }

void sort_layout(int64_t n, int64_t ngroups, int64_t seg_cap, int64_t max_group, SortLayout *l) {
    size_t off = 0;
    l->off_hdr = off;        off = align_up(off + sizeof(SortHeader), 256);
    l->off_order = off;      off = align_up(off + (size_t)n * 4, 256);
    l->off_seg_order = off;  off = align_up(off + (size_t)n * 4, 256);
    l->off_seg_rank = off;   off = align_up(off + (size_t)n * 4, 256);
    l->off_offsets = off;    off = align_up(off + (size_t)n * 4, 256);
    l->off_gseg_count = off; off = align_up(off + (size_t)ngroups * 4, 256);
    l->off_gseg_off = off;   off = align_up(off + (size_t)ngroups * 4, 256);
    l->off_gseg_start = off; off = align_up(off + (size_t)ngroups * (seg_cap + 1) * 4, 256);
    l->segs_cap = ngroups * seg_cap < kSortSegsMax ? ngroups * seg_cap : kSortSegsMax;
    l->off_segs = off;       off = align_up(off + (size_t)l->segs_cap * sizeof(NmsSegment), 256);
    l->off_gstart = off;     off = align_up(off + (size_t)(ngroups + 1) * 8, 256);
    l->big_cap = max_group > (int64_t)kSortMaxCluster * kSortCap ? max_group : 0;
    l->off_big_buf = off;    off = align_up(off + (size_t)l->big_cap * 16, 256);
    l->off_big_hist = off;   off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * 256 * 4 : 0), 256);
    l->off_big_red = off;    off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * sizeof(Red) : 0), 256);
    l->off_big_count = off;  off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * 4 : 0), 256);
    l->total_bytes = off;
}

static int sort_cluster_size(int64_t max_group) {
    int K = 1;
    while (K < kSortMaxCluster && (int64_t)K * kSortCap < max_group) K *= 2;
    return K;
}

int launch_sort(SortArgs &a, const int64_t *group_sizes_host, int64_t n, void *sortws, const SortLayout &l,
                cudaStream_t stream) {
    // group table: inline for small batches, else one small upload (pageable source: staged by the
    // runtime before the call returns)
    int64_t max_group = 0;
    char *ws = reinterpret_cast<char *>(sortws);
    if (a.ngroups <= kSortGroupsInline) {
        a.gstart = nullptr;
        int64_t s = 0;
        for (int g = 0; g < a.ngroups; ++g) {
            a.gstart_inline[g] = s;
            const int64_t len = group_sizes_host ? group_sizes_host[g] : n;
            if (len < 0) return -1;
            s += len;
            max_group = len > max_group ? len : max_group;
        }
        a.gstart_inline[a.ngroups] = s;
        if (s != n) return -1;
    } else {
        std::vector<int64_t> st((size_t)a.ngroups + 1, 0);
        for (int g = 0; g < a.ngroups; ++g) {
            const int64_t len = group_sizes_host[g];
            if (len < 0) return -1;
            st[(size_t)g + 1] = st[(size_t)g] + len;
            max_group = len > max_group ? len : max_group;
        }
        if (st[(size_t)a.ngroups] != n) return -1;
        int64_t *dst = reinterpret_cast<int64_t *>(ws + l.off_gstart);
        if (cudaMemcpyAsync(dst, st.data(), st.size() * 8, cudaMemcpyHostToDevice, stream) != cudaSuccess) return -3;
        a.gstart = dst;
    }
    if (max_group > 0xfffffff0LL || n > 0xfffffff0LL) return -1;
    a.hdr = reinterpret_cast<SortHeader *>(ws + l.off_hdr);
    a.order = reinterpret_cast<uint32_t *>(ws + l.off_order);
    a.seg_order = reinterpret_cast<uint32_t *>(ws + l.off_seg_order);
    a.seg_rank = reinterpret_cast<uint32_t *>(ws + l.off_seg_rank);
    a.offsets_out = reinterpret_cast<float *>(ws + l.off_offsets);
    a.gseg_count = reinterpret_cast<int *>(ws + l.off_gseg_count);
    a.gseg_off = reinterpret_cast<int *>(ws + l.off_gseg_off);
    a.gseg_start = reinterpret_cast<uint32_t *>(ws + l.off_gseg_start);
    a.segs = reinterpret_cast<NmsSegment *>(ws + l.off_segs);
    a.segs_cap = (int)l.segs_cap;
    a.big_buf = reinterpret_cast<uint32_t *>(ws + l.off_big_buf);
    a.big_cap = (uint32_t)l.big_cap;
    a.big_hist = reinterpret_cast<unsigned int *>(ws + l.off_big_hist);
    a.big_red = ws + l.off_big_red;
    a.big_count = reinterpret_cast<unsigned int *>(ws + l.off_big_count);
    if (cudaMemsetAsync(a.hdr, 0, sizeof(SortHeader), stream) != cudaSuccess) return -3;

    const int K = sort_cluster_size(max_group);
    const size_t dyn = (size_t)4 * kSortCap * sizeof(uint32_t);
    // (set on every call: the attribute belongs to the current device's copy of the function)
    if (cudaFuncSetAttribute(sort_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn) != cudaSuccess)
        return -3;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(a.ngroups * K));
        cfg.blockDim = dim3(kSortThreads);
        cfg.dynamicSmemBytes = dyn;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)K;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, sort_cluster_kernel, a) != cudaSuccess) return -3;
    }
    if (max_group > (int64_t)K * kSortCap) {
        // the oversize groups: cooperative grid, buffers in global memory
        int dev = 0, sms = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            return -3;
        int64_t want = (max_group + kSortCap - 1) / kSortCap;
        const int64_t limit = sms < kNmsMaxBlocks ? sms : kNmsMaxBlocks;
        if (want > limit) return -6;              // beyond sms * 12 288 elements in one group
        if (want < 8) want = 8;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)want);
        cfg.blockDim = dim3(kSortThreads);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, sort_global_kernel, a, K) != cudaSuccess) return -3;
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
