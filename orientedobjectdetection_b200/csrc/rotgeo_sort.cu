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
// CTAs, up to 8192 elements each) owns a group and keeps it in (distributed) shared memory from
// the first load to the last store; an element is one 8-byte (key, source index) pair:
//
//   load     32-bit key = descending image of the score (multiclass: candidate filter on the raw
//            score, stable compaction of the candidates across the cluster), extremes of the ids,
//            max(cx, cy) / max(w, h) of the offset formula, all reduced across the cluster
//   sort 1   4 stable LSD radix passes (8 bits) over the keys -> `order` (emission order).  Per
//            pass: every element is ranked inside its warp with 8 ballots (no atomics; match.any
//            is serialised by the hardware over the ~30 distinct digits of a warp), per-warp
//            counters in a padded digit-major table are prefix-summed by (digit, quarter) threads,
//            the CTA histograms are exchanged through DISTRIBUTED SHARED MEMORY, the elements are
//            first reordered by digit inside the CTA and only then copied to their owners, so
//            that a warp's remote stores are a few contiguous runs instead of 32 scattered words
//            (measured: 7500 scattered 4-byte st.shared::cluster per CTA and pass took 5.4 us)
//   sort 2   1-2 more passes on key' = (id - min id) << 16 | score position: segment-major order,
//            score order inside a segment -> `seg_order`, `seg_rank`, segment starts
//   plan     the segment table and the slot / tile / mask-word totals the host used to compute
//            after a device->host read (nms_plan() in rotgeo_nms.cu), left in a device header:
//            from shared memory when the call is one group, else by the cluster that finishes
//            last (atomic ticket) from the per-group results.
//
// Groups beyond one cluster's capacity (65 536 elements) take the same code with the element
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
constexpr int kItems = kSortItems;              // 8
constexpr int kCap = kSortCap;                  // elements per CTA
constexpr int kTile = 128;
constexpr int kCntRow = kWarps + 2;             // u16 per digit row: 17 words, so digits fall into different banks
constexpr int kPlanLocal = 320;                 // segments of a single-group call planned from shared memory

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
    unsigned int max_x, min_x_inv, max_y, min_y_inv;   // spatial mode: extremes of the centres (ordered images)
    unsigned int pad;
};

struct __align__(16) SortSmem {
    unsigned short cnt[256][kCntRow];   // [digit][warp]: per-warp counters, then exclusive prefixes over the warps
    unsigned int hist[256];             // this CTA's digit histogram (read by the whole cluster)
    unsigned int lstart[256];           // first staging slot of a digit inside this CTA
    unsigned int dbase[256];            // first destination of (digit, this CTA) in the group
    unsigned int wsum[2][kWarps + 1];
    unsigned int gs_local[kPlanLocal + 1];   // rank 0 of a single-group call: the segment starts
    Red red;
    unsigned int bcount;                // segment boundaries in this CTA
    int skip;                           // every key of the group has the same digit: nothing moves
    int last;                           // this cluster took the last ticket: it makes the plan
};

// ---- team = the CTAs that own one group: a cluster (shared memory) or the cooperative grid (global) ----
template <bool GLOBAL>
struct Team {
    int rank, size;
    SortSmem *sm;
    // two element buffers: CLUSTER -> this CTA's shared arrays of `stride` entries; GLOBAL -> global
    // arrays of the whole group (index = position).  stage: always this CTA's shared memory.
    uint2 *base;
    uint32_t stride;
    uint2 *stage;
    unsigned int *ghist;     // GLOBAL: [size][256]
    Red *gred;               // GLOBAL: [size]
    unsigned int *gcount;    // GLOBAL: [size] boundary counts

    __device__ __forceinline__ uint2 *buf(int c) const { return base + (size_t)c * stride; }
    __device__ __forceinline__ void sync() const {
        if (GLOBAL) {
            __threadfence();
            cg::this_grid().sync();
        } else {
            cg::this_cluster().sync();
        }
    }
    // element at local position lp of this CTA (share = positions per CTA)
    __device__ __forceinline__ uint2 ld(int c, uint32_t share, uint32_t lp) const {
        return GLOBAL ? __ldcg(buf(c) + (size_t)rank * share + lp) : buf(c)[lp];
    }
    __device__ __forceinline__ void st_local(int c, uint32_t share, uint32_t lp, uint2 v) const {
        if (GLOBAL) __stcg(buf(c) + (size_t)rank * share + lp, v); else buf(c)[lp] = v;
    }
    // element at group position p (lands in the shared memory of CTA p / share)
    __device__ __forceinline__ void st_pos(int c, uint32_t share, uint32_t p, uint2 v) const {
        if (GLOBAL) {
            __stcg(buf(c) + p, v);
        } else {
            const uint32_t r = p / share;
            uint2 *dst = (int)r == rank ? buf(c) : cg::this_cluster().map_shared_rank(buf(c), r);
            dst[p - r * share] = v;
        }
    }
    __device__ __forceinline__ uint32_t ld_pos_key(int c, uint32_t share, uint32_t p) const {
        if (GLOBAL) return __ldcg(&buf(c)[p].x);
        const uint32_t r = p / share;
        const uint2 *src = cg::this_cluster().map_shared_rank(buf(c), r);
        return src[p - r * share].x;
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
            r.max_x = p->max_x; r.min_x_inv = p->min_x_inv; r.max_y = p->max_y; r.min_y_inv = p->min_y_inv;
            return r;
        }
        return *cg::this_cluster().map_shared_rank(&sm->red, c);
    }
    __device__ __forceinline__ unsigned int peer_count(int c) const {
        if (GLOBAL) return *reinterpret_cast<const volatile unsigned int *>(&gred[c].count);
        return cg::this_cluster().map_shared_rank(&sm->red, c)->count;
    }
    __device__ __forceinline__ void publish_bcount(unsigned int v) const {
        sm->bcount = v;
        if (GLOBAL) __stcg(gcount + rank, v);
    }
    __device__ __forceinline__ unsigned int peer_bcount(int c) const {
        if (GLOBAL) return __ldcg(gcount + c);
        return *cg::this_cluster().map_shared_rank(&sm->bcount, c);
    }
    // rank 0's copy of the segment starts (single-group calls, cluster path only)
    __device__ __forceinline__ void st_gs_rank0(uint32_t idx, unsigned int v) const {
        if (GLOBAL) return;
        unsigned int *dst = rank == 0 ? sm->gs_local : cg::this_cluster().map_shared_rank(sm->gs_local, 0);
        dst[idx] = v;
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

// exclusive scan of one value per thread over the block; returns the exclusive prefix, total in `total`.
// Two levels: warp scans, then warp 0 scans the 32 warp sums (every thread reading all 32 sums cost
// ~1 us of shared-memory bandwidth per scan with 1024 threads).  s_warp: kWarps + 1 entries.
template <typename V>
__device__ __forceinline__ V block_exclusive_scan(V v, V *s_warp, V &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    V incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const V t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    __syncthreads();                       // s_warp may still be read from the previous scan
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const V w = s_warp[lane];          // kWarps == 32
        V wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const V t = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += t;
        }
        s_warp[lane] = wi - w;
        if (lane == 31) s_warp[kWarps] = wi;
    }
    __syncthreads();
    total = s_warp[kWarps];
    return s_warp[warp] + incl - v;
}

// `run` is a warp-uniform count; returns the sum over the warps before mine, `total` = all warps.
// s: kWarps + 1 entries.
__device__ __forceinline__ unsigned int warp_totals_exclusive(unsigned int run, unsigned int *s, unsigned int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s[warp] = run;
    __syncthreads();
    if (warp == 0) {
        const unsigned int w = s[lane];
        unsigned int wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += t;
        }
        s[lane] = wi - w;
        if (lane == 31) s[kWarps] = wi;
    }
    __syncthreads();
    total = s[kWarps];
    return s[warp];
}

#ifdef ROTGEO_SORT_TIMING   // probe builds (tools/sort_phases.py): clocks of one pass and of the phases of group 0
#define PT_DECL unsigned long long pt_[12]; int ptn_ = 0;
#define PT_MARK do { if (blockIdx.x == 0 && threadIdx.x == 0 && shift == 8 && ptn_ < 12) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pt_[ptn_])); ++ptn_; } } while (0)
#define PT_PRINT do { if (blockIdx.x == 0 && threadIdx.x == 0 && shift == 8) { printf("  pass steps (us):"); for (int i_ = 1; i_ < ptn_; ++i_) printf(" %.2f", (double)(pt_[i_] - pt_[i_ - 1]) * 1e-3); printf("\n"); } } while (0)
#define ST_DECL unsigned long long st_[24]; int stn_ = 0;
#define ST_MARK do { if (g == 0 && T.rank == 0 && threadIdx.x == 0 && stn_ < 24) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(st_[stn_])); ++stn_; } } while (0)
#define ST_PRINT do { if (g == 0 && T.rank == 0 && threadIdx.x == 0) { printf("sort phases (us), n=%u size=%d:", n, T.size); for (int i_ = 1; i_ < stn_; ++i_) printf(" %.1f", (double)(st_[i_] - st_[i_ - 1]) * 1e-3); printf("\n"); } } while (0)
#else
#define PT_DECL
#define PT_MARK do {} while (0)
#define PT_PRINT do {} while (0)
#define ST_DECL
#define ST_MARK do {} while (0)
#define ST_PRINT do {} while (0)
#endif

// One stable LSD pass over the 8 key bits at `shift`: the n elements of the group move from buffer
// `cur` to buffer `cur ^ 1`.  Position order inside a CTA = (warp, item, lane), iu items per thread.
// Returns true when the buffers were swapped (false: all keys share the digit, nothing moved).
template <bool GLOBAL>
__device__ bool radix_pass(const Team<GLOBAL> &T, int cur, int shift, uint32_t n, uint32_t share) {
    SortSmem &sm = *T.sm;
    PT_DECL
    PT_MARK;   // 0
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t count = count_of(n, share, T.rank);
    const int iu = (int)((share + kThreads - 1) / kThreads);
    const uint32_t lt = (1u << lane) - 1u;
    // 1. zero the counters and the match masks (they alias the staging array, dirty from the last pass)
    unsigned int *const mm = reinterpret_cast<unsigned int *>(T.stage) + warp * 256;   // [warp][digit] lane masks
    {
        uint4 *z = reinterpret_cast<uint4 *>(&sm.cnt[0][0]);
        constexpr int kZ = 256 * kCntRow * 2 / 16;
        for (int i = tid; i < kZ; i += kThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
        uint4 *zm = reinterpret_cast<uint4 *>(T.stage);
        for (int i = tid; i < kWarps * 256 * 4 / 16; i += kThreads) zm[i] = make_uint4(0u, 0u, 0u, 0u);
        if (tid == 0) sm.skip = 0;
    }
    __syncthreads();
    PT_MARK;   // 1 zeroed
    // 2. rank every element among the elements of its warp with the same digit: the lanes that share
    //    a digit meet in one shared-memory word (atomicOr of the lane bit, read back) — 3 shared-memory
    //    operations per element where 8 ballots cost 0.75 us per item (the VOTE pipe) and match.any is
    //    serialised over the ~30 distinct digits of a warp; then the warp's running count of the digit
    uint32_t dg[kItems], pm[kItems];       // digit (0xffffffff = no element), rank inside (warp, digit)
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        dg[i] = 0xffffffffu;
        pm[i] = 0u;
        if (i < iu) {                                  // warp-uniform
            const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
            const bool active = lp < count;
            const uint32_t d = active ? ((T.ld(cur, share, lp).x >> shift) & 0xffu) : 0u;
            if (active) atomicOr(mm + d, 1u << lane);
            __syncwarp();
            uint32_t peers = 0u;
            unsigned int base = 0u;
            if (active) {
                peers = mm[d];
                base = sm.cnt[d][warp];
            }
            __syncwarp();
            if (active) {
                if ((peers & lt) == 0u) {              // lowest lane of the digit
                    mm[d] = 0u;
                    sm.cnt[d][warp] = (unsigned short)(base + __popc(peers));
                }
                dg[i] = d;
                pm[i] = base + __popc(peers & lt);
            }
            __syncwarp();
        }
    }
    __syncthreads();
    PT_MARK;   // 2 ranked
    // 3. exclusive prefix over the warps + the CTA histogram: thread = (digit, quarter of the warps),
    //    its 8 counters are 4 consecutive words of the digit's row, the quarters joined by shuffles
    {
        const int d = tid >> 2, q = tid & 3;
        uint32_t *row = reinterpret_cast<uint32_t *>(&sm.cnt[d][0]) + 4 * q;
        uint32_t w[4];
        unsigned int c[8], sum = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            w[k] = row[k];
            c[2 * k] = w[k] & 0xffffu;
            c[2 * k + 1] = w[k] >> 16;
            sum += c[2 * k] + c[2 * k + 1];
        }
        unsigned int incl = sum;
#pragma unroll
        for (int o = 1; o < 4; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o, 4);
            if (q >= o) incl += t;
        }
        unsigned int run = incl - sum;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned int lo = run;
            run += c[2 * k];
            const unsigned int hi = run;
            run += c[2 * k + 1];
            row[k] = lo | (hi << 16);
        }
        if (q == 3) T.publish_hist(d, incl);
    }
    PT_MARK;   // 3 warp prefix
    T.sync();
    PT_MARK;   // 4 cluster sync
    // 4. totals of the group, what lies before this CTA, and two exclusive scans over the digits:
    //    of the totals (destinations in the group) and of my own histogram (staging slots)
    {
        unsigned int tot = 0, before = 0, mine = 0;
        if (tid < 256) {
            mine = sm.hist[tid];
            if (!GLOBAL) {
                unsigned int h[kSortMaxCluster];
#pragma unroll
                for (int c = 0; c < kSortMaxCluster; ++c) h[c] = c < T.size ? T.peer_hist(c, tid) : 0u;
#pragma unroll
                for (int c = 0; c < kSortMaxCluster; ++c) {
                    tot += h[c];
                    before += c < T.rank ? h[c] : 0u;
                }
            } else {
                for (int c = 0; c < T.size; ++c) {
                    const unsigned int h = T.peer_hist(c, tid);
                    tot += h;
                    before += c < T.rank ? h : 0u;
                }
            }
            if (tot == n) sm.skip = 1;
        }
        unsigned int i1 = tot, i2 = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int t1 = __shfl_up_sync(0xffffffffu, i1, d);
            const unsigned int t2 = __shfl_up_sync(0xffffffffu, i2, d);
            if (lane >= d) { i1 += t1; i2 += t2; }
        }
        if (tid < 256 && lane == 31) { sm.wsum[0][warp] = i1; sm.wsum[1][warp] = i2; }
        __syncthreads();
        if (tid < 256) {
            unsigned int p1 = 0, p2 = 0;
            for (int w = 0; w < warp; ++w) { p1 += sm.wsum[0][w]; p2 += sm.wsum[1][w]; }
            sm.dbase[tid] = p1 + i1 - tot + before;
            sm.lstart[tid] = p2 + i2 - mine;
        }
    }
    __syncthreads();
    PT_MARK;   // 5 totals + digit scans
    const bool moved = sm.skip == 0;
    if (moved) {
        // 5. reorder by digit inside the CTA (stable: warp-major, then item, then lane) ...
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            if (i < iu && dg[i] != 0xffffffffu) {
                const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                const uint32_t d = dg[i];
                T.stage[sm.lstart[d] + sm.cnt[d][warp] + pm[i]] = T.ld(cur, share, lp);
            }
        }
        __syncthreads();
        // 6. ... and copy the runs to their owners: consecutive threads, consecutive destinations
        for (uint32_t j = tid; j < count; j += kThreads) {
            const uint2 e = T.stage[j];
            const uint32_t d = (e.x >> shift) & 0xffu;
            T.st_pos(cur ^ 1, share, sm.dbase[d] + (j - sm.lstart[d]), e);
        }
    }
    PT_MARK;   // 6 scatter issued
    T.sync();     // also: nobody overwrites its histogram while a peer still reads it
    PT_MARK;   // 7 cluster sync
    PT_PRINT;
    return moved;
}

__device__ __forceinline__ int64_t group_start(const SortArgs &a, int g) {
    return a.gstart ? a.gstart[g] : a.gstart_inline[g];
}

// The plan nms_plan() makes on the host: segment table + totals, by one CTA (1024 threads).
// fetch(j, start, len): first position (in seg_order) and length of segment j of all Tc segments.
template <typename Fetch>
__device__ void emit_plan(const SortArgs &a, long long Tc, unsigned int err, Fetch fetch) {
    const int tid = threadIdx.x;
    __shared__ long long s_warp[kWarps + 1];
    __shared__ int s_maxc[kWarps];
    long long slots = 0, words = 0, tiles = 0;
    int maxc = 0;
    for (long long base = 0; base < Tc; base += kThreads) {
        const long long j = base + tid;
        long long len = 0, C = 0, start = 0;
        if (j < Tc) {
            fetch(j, start, len);
            C = (len + kTile - 1) / kTile;
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

// several groups: from the per-group results in global memory, by the cluster that finished last
__device__ void make_plan(const SortArgs &a) {
    const int tid = threadIdx.x;
    __shared__ long long s_warp2[kWarps + 1];
    long long carry = 0;                  // segments before each group
    for (int base = 0; base < a.ngroups; base += kThreads) {
        const int g = base + tid;
        long long v = 0;
        if (g < a.ngroups) v = __ldcg(a.gseg_count + g);
        long long tot;
        const long long ex = block_exclusive_scan(v, s_warp2, tot);
        if (g < a.ngroups) a.gseg_off[g] = (int)(carry + ex);
        carry += tot;
    }
    const long long Tall = carry;
    __threadfence_block();
    __syncthreads();
    const long long Tc = Tall > a.segs_cap ? a.segs_cap : Tall;
    emit_plan(a, Tc, Tall > a.segs_cap ? kSortErrSegments : 0u, [&](long long j, long long &start, long long &len) {
        int lo = 0, hi = a.ngroups - 1;     // last group with gseg_off <= j (groups without segments share offsets)
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (a.gseg_off[mid] <= j) lo = mid; else hi = mid - 1;
        }
        const int k = (int)(j - a.gseg_off[lo]);
        const uint32_t *gs = a.gseg_start + (size_t)lo * (a.seg_cap + 1);
        const long long s0 = __ldcg(gs + k), s1 = __ldcg(gs + k + 1);
        len = s1 - s0;
        start = (a.mode == 1 ? 0 : group_start(a, lo)) + s0;
    });
}

// one group: the segment starts are in rank 0's shared memory (gs_local[0 .. nseg], the last = n)
__device__ void make_plan_local(const SortArgs &a, const SortSmem &sm, int nseg) {
    emit_plan(a, nseg, 0u, [&](long long j, long long &start, long long &len) {
        start = sm.gs_local[j];
        len = (long long)sm.gs_local[j + 1] - start;
    });
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
    const bool fast_plan = !GLOBAL && a.ngroups == 1;
    ST_DECL
    ST_MARK;   // 0 start

    if (tid == 0) {
        sm.red.id_max = 0ull; sm.red.id_min_inv = 0ull; sm.red.max_xy = 0u; sm.red.max_wh = 0u;
        sm.red.count = 0u; sm.red.pad = 0u;
        sm.red.max_x = sm.red.min_x_inv = sm.red.max_y = sm.red.min_y_inv = 0u;
        sm.bcount = 0u;
        sm.last = 0;
        sm.gs_local[0] = 0u;
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
        unsigned int cxmx = 0u, cxmn = 0u, cymx = 0u, cymn = 0u;
        uint32_t kv[kItems];          // keys of my elements (multiclass: 0xffffffff = not a candidate)
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
                        if (mc || a.ids) {
                            mxy = max(mxy, ordered_bits(fmaxf(b[0], b[1])));
                            mwh = max(mwh, ordered_bits(fmaxf(b[2], b[3])));
                        }
                        if (!mc && a.ids) {
                            const unsigned long long u = id_image(a.ids[g0 + q]);
                            idmx = max(idmx, u);
                            idmn = max(idmn, ~u);
                        }
                        if (a.spatial) {
                            const unsigned int ux = ordered_bits(b[0]), uy = ordered_bits(b[1]);
                            cxmx = max(cxmx, ux); cxmn = max(cxmn, ~ux);
                            cymx = max(cymx, uy); cymn = max(cymn, ~uy);
                        }
                    }
                    if (!mc) T.st_local(0, share, lp, make_uint2(kv[i], q));
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
        if (a.spatial) {
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                cxmx = max(cxmx, __shfl_xor_sync(0xffffffffu, cxmx, d));
                cxmn = max(cxmn, __shfl_xor_sync(0xffffffffu, cxmn, d));
                cymx = max(cymx, __shfl_xor_sync(0xffffffffu, cymx, d));
                cymn = max(cymn, __shfl_xor_sync(0xffffffffu, cymn, d));
            }
            if (lane == 0) {
                atomicMax(&sm.red.max_x, cxmx);
                atomicMax(&sm.red.min_x_inv, cxmn);
                atomicMax(&sm.red.max_y, cymx);
                atomicMax(&sm.red.min_y_inv, cymn);
            }
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
            unsigned int ctot;
            const unsigned int wbefore = warp_totals_exclusive(run, sm.wsum[0], ctot);
            if (tid == 0) sm.red.count = ctot;
            __syncthreads();
            T.publish_red();
            T.sync();
            unsigned int before = 0, total = 0;
            {
                unsigned int h[kSortMaxCluster];
                if (!GLOBAL) {
#pragma unroll
                    for (int c = 0; c < kSortMaxCluster; ++c) h[c] = c < T.size ? T.peer_count(c) : 0u;
#pragma unroll
                    for (int c = 0; c < kSortMaxCluster; ++c) {
                        total += h[c];
                        before += c < T.rank ? h[c] : 0u;
                    }
                } else {
                    for (int c = 0; c < T.size; ++c) {
                        const unsigned int v = T.peer_count(c);
                        total += v;
                        before += c < T.rank ? v : 0u;
                    }
                }
            }
            n = total;
            const uint32_t share2 = share_of(n, T.size);
#pragma unroll
            for (int i = 0; i < kItems; ++i) {
                if (i < iu && kv[i] != 0xffffffffu) {
                    const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                    const uint32_t q = (uint32_t)T.rank * share + lp;
                    T.st_pos(0, share2, before + wbefore + rk[i], make_uint2(kv[i], q));
                }
            }
            share = share2;
            T.sync();
        } else {
            __syncthreads();
            T.publish_red();
            T.sync();
        }
    }
    // group-wide reductions
    Red R;
    if (!GLOBAL) {
        Red o[kSortMaxCluster];       // all remote reads in flight together
#pragma unroll
        for (int c = 0; c < kSortMaxCluster; ++c) o[c] = T.peer_red(c < T.size ? c : 0);
        R = o[0];
#pragma unroll
        for (int c = 1; c < kSortMaxCluster; ++c) {
            R.id_max = max(R.id_max, o[c].id_max);
            R.id_min_inv = max(R.id_min_inv, o[c].id_min_inv);
            R.max_xy = max(R.max_xy, o[c].max_xy);
            R.max_wh = max(R.max_wh, o[c].max_wh);
            R.max_x = max(R.max_x, o[c].max_x); R.min_x_inv = max(R.min_x_inv, o[c].min_x_inv);
            R.max_y = max(R.max_y, o[c].max_y); R.min_y_inv = max(R.min_y_inv, o[c].min_y_inv);
        }
    } else {
        R = T.peer_red(0);
        for (int c = 1; c < T.size; ++c) {
            const Red o = T.peer_red(c);
            R.id_max = max(R.id_max, o.id_max);
            R.id_min_inv = max(R.id_min_inv, o.id_min_inv);
            R.max_xy = max(R.max_xy, o.max_xy);
            R.max_wh = max(R.max_wh, o.max_wh);
            R.max_x = max(R.max_x, o.max_x); R.min_x_inv = max(R.min_x_inv, o.min_x_inv);
            R.max_y = max(R.max_y, o.max_y); R.min_y_inv = max(R.min_y_inv, o.min_y_inv);
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
    ST_MARK;   // 1 load + reductions

    // ---------------------------------------------------------------- sort 1: score order
    for (int shift = 0; shift < 32; shift += 8) {
        if (radix_pass(T, cur, shift, n, share)) cur ^= 1;
        ST_MARK;   // 2..5 passes
    }

    const uint32_t count = count_of(n, share, T.rank);
    const int iu = (int)((share + kThreads - 1) / kThreads);
    const int64_t obase = mc ? 0 : g0;       // `order` holds global rows (multiclass: flat candidate indices)
    // -> order; the elements become those of sort 2, in place.  Cluster path: key' = (id - min id) << 16 |
    // score position (a group in shared memory has at most 65 536 elements), the source index stays
    // the payload, so nothing has to be gathered afterwards.  Global path: key' = id - min id, payload =
    // score position, and the source index is gathered from `order` after the passes.
    const bool two = has_ids && !range_bad && id_range > 1ull;
    // Spatial mode (large sparse segments): inside a segment the boxes are packed in Morton order of
    // their centres instead of score order, so that a 128-box block covers a small area and only
    // neighbouring blocks form mask tiles; key' = (id - min id) << 16 | Morton16, payload = score position
    // (the rank the mask kernel compares), source index gathered from `order` like on the global path.
    const bool spatial = a.spatial != 0 && !mc;
    const bool resort = two || spatial;
    const bool gather = GLOBAL || spatial;
    const float sp_x0 = ordered_float(~R.min_x_inv), sp_y0 = ordered_float(~R.min_y_inv);
    const float sp_ext = fmaxf(fmaxf(ordered_float(R.max_x) - sp_x0, ordered_float(R.max_y) - sp_y0), 1e-20f);
    const float sp_scale = 255.999f / sp_ext;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        if (i < iu) {
            const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
            if (lp < count) {
                const uint32_t p = (uint32_t)T.rank * share + lp;
                const uint32_t idx = T.ld(cur, share, lp).y;
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
                if (spatial) {
                    const float *b = a.boxes + (size_t)(g0 + idx) * a.stride;
                    // 8 bits per axis, interleaved (NaN / inf centres land in cell 0 or 255: order only)
                    uint32_t ix = (uint32_t)fminf(fmaxf((b[0] - sp_x0) * sp_scale, 0.f), 255.f);
                    uint32_t iy = (uint32_t)fminf(fmaxf((b[1] - sp_y0) * sp_scale, 0.f), 255.f);
                    ix = (ix | (ix << 4)) & 0x0f0fu; ix = (ix | (ix << 2)) & 0x3333u; ix = (ix | (ix << 1)) & 0x5555u;
                    iy = (iy | (iy << 4)) & 0x0f0fu; iy = (iy | (iy << 2)) & 0x3333u; iy = (iy | (iy << 1)) & 0x5555u;
                    T.st_local(cur, share, lp, make_uint2(((two ? idk : 0u) << 16) | ix | (iy << 1), p));
                } else if (two) {
                    T.st_local(cur, share, lp, GLOBAL ? make_uint2(idk, p) : make_uint2((idk << 16) | p, idx));
                } else {
                    a.seg_order[obase + p] = src;
                    a.seg_rank[obase + p] = p;
                }
            }
        }
    }
    if (gather) {          // `order` is read back through global memory below
        __threadfence();
        T.sync();
    }
    ST_MARK;   // 6 order written

    // ---------------------------------------------------------------- sort 2: segment-major order
    unsigned int nseg_here = n > 0 ? 1u : 0u;
    uint32_t *gs = a.gseg_start + (size_t)g * (a.seg_cap + 1);
    if (resort) {
        if (spatial) {
            if (radix_pass(T, cur, 0, n, share)) cur ^= 1;
            if (radix_pass(T, cur, 8, n, share)) cur ^= 1;
        }
        if (two) {
            const int s0 = (GLOBAL && !spatial) ? 0 : 16;
            if (radix_pass(T, cur, s0, n, share)) cur ^= 1;
            if (id_range > 256ull)
                if (radix_pass(T, cur, s0 + 8, n, share)) cur ^= 1;
        }
        const bool key_hi = !(GLOBAL && !spatial);       // the segment key sits in the upper half of the key
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
                    const uint2 e = T.ld(cur, share, lp);
                    const uint32_t k = key_hi ? (e.x >> 16) : e.x;
                    const uint32_t p = gather ? e.y : (e.x & 0xffffu);
                    a.seg_rank[obase + q] = p;
                    a.seg_order[obase + q] = gather ? __ldcg(a.order + obase + p) : (uint32_t)(obase + e.y);
                    if (q == 0u) {
                        isb = true;
                    } else {
                        const uint32_t pk = lp > 0u ? T.ld(cur, share, lp - 1u).x : T.ld_pos_key(cur, share, q - 1u);
                        isb = (key_hi ? (pk >> 16) : pk) != k;
                    }
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, isb);
                if (isb) bmask |= 1u << i;
                ordl[i] = run + __popc(bal & lt);
                run += __popc(bal);
            }
        }
        unsigned int ctot;
        const unsigned int wbefore = warp_totals_exclusive(run, sm.wsum[0], ctot);
        if (tid == 0) T.publish_bcount(ctot);
        T.sync();
        unsigned int before = 0, total = 0;
        if (!GLOBAL) {
            unsigned int h[kSortMaxCluster];
#pragma unroll
            for (int c = 0; c < kSortMaxCluster; ++c) h[c] = c < T.size ? T.peer_bcount(c) : 0u;
#pragma unroll
            for (int c = 0; c < kSortMaxCluster; ++c) {
                total += h[c];
                before += c < T.rank ? h[c] : 0u;
            }
        } else {
            for (int c = 0; c < T.size; ++c) {
                const unsigned int h = T.peer_bcount(c);
                total += h;
                before += c < T.rank ? h : 0u;
            }
        }
        nseg_here = total;
        if (total <= (unsigned int)a.seg_cap) {
            const bool to_rank0 = fast_plan && total <= (unsigned int)kPlanLocal;
#pragma unroll
            for (int i = 0; i < kItems; ++i) {
                if (i < iu && (bmask >> i) & 1u) {
                    const uint32_t lp = (uint32_t)warp * 32u * iu + (uint32_t)i * 32u + lane;
                    const uint32_t at = before + wbefore + ordl[i];
                    gs[at] = (uint32_t)T.rank * share + lp;
                    if (to_rank0) T.st_gs_rank0(at, (uint32_t)T.rank * share + lp);
                }
            }
        }
    } else if (T.rank == 0 && tid == 0 && n > 0) {
        gs[0] = 0u;
    }
    ST_MARK;   // 7 sort 2 + segment starts
    // ---------------------------------------------------------------- group summary, plan
    unsigned int nseg_ok = nseg_here > (unsigned int)a.seg_cap ? 0u : nseg_here;
    if (T.rank == 0 && tid == 0) {
        unsigned int err = range_bad ? kSortErrRange : 0u;
        if (nseg_here > (unsigned int)a.seg_cap) err |= kSortErrSegments;
        gs[nseg_ok] = n;      // end of the last segment
        if (nseg_ok <= (unsigned int)kPlanLocal) sm.gs_local[nseg_ok] = n;
        a.gseg_count[g] = (int)nseg_ok;
        if (a.ngroups == 1) {
            // the only writer: plain stores, and the launch needs no memset of the header in front of it
            // (a memset node + its launch gap were 9 us of every single-group call)
            a.hdr->done = 0u;
            a.hdr->error = err;
            a.hdr->n_valid = (long long)n;
        } else {
            if (err) atomicOr(&a.hdr->error, err);
            atomicAdd(reinterpret_cast<unsigned long long *>(&a.hdr->n_valid), (unsigned long long)n);
        }
        if (mc) {
            a.hdr->mc_max[0] = ordered_float(R.max_xy);
            a.hdr->mc_max[1] = ordered_float(R.max_wh);
        }
    }
    __threadfence();
    T.sync();       // every CTA of the team has written its share
    ST_MARK;   // 8 summary
    if (T.rank == 0) {
        if (fast_plan && nseg_ok <= (unsigned int)kPlanLocal) {
            make_plan_local(a, sm, (int)nseg_ok);       // no ticket, no global round trips
        } else {
            if (tid == 0) {
                __threadfence();
                sm.last = atomicAdd(&a.hdr->done, 1u) == (unsigned int)a.ngroups - 1u ? 1 : 0;
            }
            __syncthreads();
            if (sm.last) {
                __threadfence();
                make_plan(a);
            }
        }
    }
    ST_MARK;   // 9 plan
    ST_PRINT;
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
    T.base = reinterpret_cast<uint2 *>(dyn);
    T.stride = (uint32_t)a.cta_cap;
    T.stage = T.base + 2 * (size_t)a.cta_cap;
    T.ghist = nullptr; T.gred = nullptr; T.gcount = nullptr;
    cluster.sync();        // every CTA of the cluster has started: its shared memory may be written
    sort_group<false>(a, g, T);
    cluster.sync();        // nobody exits while a peer may still read its shared memory
}

// cooperative grid: the groups the clusters skipped, one after the other
__global__ void __launch_bounds__(kSortThreads, 1)
sort_global_kernel(const SortArgs a, int K_cluster) {
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ SortSmem sm;
    for (int g = 0; g < a.ngroups; ++g) {
        const int64_t n_in = group_start(a, g + 1) - group_start(a, g);
        if (n_in <= (int64_t)K_cluster * kCap) continue;
        Team<true> T;
        T.rank = (int)blockIdx.x;
        T.size = (int)gridDim.x;
        T.sm = &sm;
        T.base = reinterpret_cast<uint2 *>(a.big_buf);
        T.stride = a.big_cap;
        T.stage = reinterpret_cast<uint2 *>(dyn);       // kSortCap entries: the grid has >= n / kSortCap CTAs
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
    l->off_big_buf = off;    off = align_up(off + (size_t)l->big_cap * 16, 256);     // two buffers of uint2
    l->off_big_hist = off;   off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * 256 * 4 : 0), 256);
    l->off_big_red = off;    off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * sizeof(Red) : 0), 256);
    l->off_big_count = off;  off = align_up(off + (size_t)(l->big_cap ? kNmsMaxBlocks * 4 : 0), 256);
    l->total_bytes = off;
}

// CTAs per group: what the capacity needs, and beyond that as many as keep all clusters of the
// launch resident at once (one 1024-thread CTA per SM): a pass is mostly fixed cost, so a single
// image of 10 000 boxes sorts faster on 8 SMs than on one, while a batch of 64 images is spread
// 2 CTAs per image
static int sort_cluster_size(int64_t max_group, int64_t ngroups, int sms) {
    int K = 1;
    while (K < kSortMaxCluster && (int64_t)K * kSortCap < max_group) K *= 2;
    while (K < kSortMaxCluster && (int64_t)2 * K * ngroups <= sms && (int64_t)K * 1024 < max_group) K *= 2;
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
    // several groups: tickets and sums accumulate in the header (one group writes it with plain stores)
    if (a.ngroups > 1 && cudaMemsetAsync(a.hdr, 0, sizeof(SortHeader), stream) != cudaSuccess) return -3;

    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        return -3;
    const int K = sort_cluster_size(max_group, a.ngroups, sms);
    // elements per CTA buffer: the share of the largest group that fits the cluster
    int64_t cap_elems = max_group < (int64_t)K * kSortCap ? (max_group + K - 1) / K : kSortCap;
    cap_elems = (cap_elems + 31) / 32 * 32;
    if (cap_elems < 32) cap_elems = 32;
    a.cta_cap = (int)cap_elems;
    // two buffers + the staging array, which also holds the 32 KB of per-warp match masks while ranking
    size_t stage_bytes = (size_t)cap_elems * sizeof(uint2);
    if (stage_bytes < (size_t)kSortThreads / 32 * 256 * 4) stage_bytes = (size_t)kSortThreads / 32 * 256 * 4;
    const size_t dyn = (size_t)2 * cap_elems * sizeof(uint2) + stage_bytes;
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
        // the oversize groups: cooperative grid, buffers in global memory, staging in shared memory
        int64_t want = (max_group + kSortCap - 1) / kSortCap;
        const int64_t limit = sms < kNmsMaxBlocks ? sms : kNmsMaxBlocks;
        if (want > limit) return -6;              // beyond sms * 8192 elements in one group
        if (want < 8) want = 8;
        const size_t gdyn = (size_t)kSortCap * sizeof(uint2);
        if (cudaFuncSetAttribute(sort_global_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gdyn) != cudaSuccess)
            return -3;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)want);
        cfg.blockDim = dim3(kSortThreads);
        cfg.dynamicSmemBytes = gdyn;
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
