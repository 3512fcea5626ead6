// Segmented greedy rotated NMS, entirely on the device.  sm_100a.
//
// Input: boxes packed in SEGMENT-MAJOR order (segment = class / FPN level /
// image x level ...), score-descending inside each segment, every segment
// starting at a packed slot that is a multiple of 128.  Two kernels:
//
//   nms_mask_kernel     the suppression relation as a bit matrix, block-diagonal and
//                       LOWER-triangular: bit i of row j (i < j in score order) says
//                       "i suppresses j if i is kept" — a row lists the earlier boxes
//                       its own fate depends on.  One work item per 128x128 tile of the
//                       lower triangle of each segment's pair matrix (pairs of different
//                       segments are never touched).  Same machinery as the IoU kernel:
//                       CTAs made persistent by cluster-launch-control work stealing,
//                       cull data (row bin ranges + the column block's inverted bin
//                       index) TMA-staged in a shared-memory ring, tile zeros stored
//                       first, surviving pairs queued per warp, circle-tested 32 at a
//                       time, re-queued and clipped 32 at a time; `IoU >= thr` becomes an
//                       atomicOr into the (already zeroed) mask word plus one into the
//                       row's SUMMARY (one bit per 128-column block that holds a bit).
//   nms_resolve_kernel  the greedy decision as a data-parallel fixed point instead of a
//                       serial scan.  State: two monotone bitmaps, kept and removed.  A
//                       box is removed as soon as one earlier neighbour is kept, and kept
//                       as soon as all earlier neighbours are removed — exactly the
//                       greedy result, whatever the evaluation order.  One cooperative
//                       persistent grid; every 64-box chunk belongs to one warp, which
//                       walks the set summary bits of each undecided row (16 B of mask per
//                       bit), tests them against the bitmaps, settles the chunk's internal
//                       chains by peeling (ballots, usually 1-3 iterations), and repeats
//                       until its chunks are decided.  No rounds and no barriers: segments
//                       of <= 4096 boxes live in one CTA's shared memory, larger ones
//                       exchange their bitmaps through L2 (ld.cg / st.cg); a row's
//                       summary drops the blocks whose boxes are all removed.  The same
//                       grid then writes the per-box flags and compacts them in GLOBAL
//                       score order into keep_out (and, for the fused multiclass call,
//                       the (box, score) rows and labels) — the order
//                       mmcv.ops.nms_rotated returns and every reference caller slices
//                       with [:max_num].  The mask never leaves the GPU.
//
// Replaces mmcv's nms_rotated (64x64 bitmask over the full square + D2H copy of
// the mask + serial host scan; reference call sites
// mmrotate/core/post_processing/bbox_nms_rotated.py:80,122,
// mmrotate/core/patch/merge_results.py:116-118, mmrotate/datasets/dota.py:379).
#include <math.h>
#include <stdlib.h>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTile = 128;      // tile edge in boxes = one inverted-index block = two mask words
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxStages = 4;   // cull-data ring: 4 stages of the x/y index (5 KB), or 3 with the area axis (7.7 KB)
constexpr int kTabWords = 3 * kTableWords;   // 5760 words >= 4 * kTableXYWords
constexpr int kInfoSlots = 8;   // tile descriptors / launch-control responses (> kStages)
constexpr int kQueueCap = 64;   // pairs per warp queue (never more than 31 + 32 queued)
static_assert(kTile == kBlock, "a tile column is one inverted-index block");

struct TileInfo {
    int valid;        // 0: nothing left to steal
    int seg;          // segment index
    int I, J;         // row / column block inside the segment (J <= I)
};

struct QEntry {       // one pair that survived cull level 1
    int prow, pcol;               // packed slots
    unsigned long long word;      // low 48 bits: index of the mask word holding its bit;
                                  // high 16 bits: column block J (its summary bit)
};
constexpr unsigned long long kWordMask = (1ull << 48) - 1ull;

struct __align__(16) MaskSmem {
    uint32_t tab[kTabWords];
    uint32_t rowr[kMaxStages][kTile];
    QEntry queue1[kWarps][kQueueCap];   // survived the bin ranges
    QEntry queue2[kWarps][kQueueCap];   // survived the circle test as well
    uint4 resp[kInfoSlots];
    TileInfo info[kInfoSlots];
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t sched[kInfoSlots];
    uint64_t ready[kInfoSlots];
    uint64_t zeroed[kInfoSlots];        // list mode: all eight warps have zeroed their rows of my k-th tile
    // the scheduler's state (the duty rotates over the warps; nobody carries it in registers)
    struct Sched {
        unsigned next_ask, next_resp;   // launch-control requests issued / answers consumed, from 1
        unsigned grp_next, grp_left;    // rest of the tile group in hand
        int alive;                      // no request has failed yet
        int cur_nb;                     // 128-row blocks of cur's segment
        TileInfo cur;                   // the tile located last
    } st;
};

struct MaskArgs {
    const float4 *p0, *p1;
    const uint32_t *pr, *pt;
    const NmsSegment *segs;
    int nseg;
    const long long *nseg_dev;    // nullable: actual segment count (device-planned launch)
    int zero_in_kernel;           // list mode: the mask kernel zeroes the listed tiles itself (one tile ahead)
    float thr;
    float area_ratio;             // pairs with min(area) < area_ratio * max(area) cannot reach thr
    float inter_need;             // 0.999 thr / (1 + thr): intersection / (areaA + areaB) a pair needs to reach thr
    int area_axis;                // the packed set carries the area axis of the index (launch-uniform)
    unsigned long long *mask;
    unsigned long long *summary;  // per packed slot: sum_words words, zeroed by the pack kernel
    int sum_words;
    const int64_t *total_tiles;   // nullable: device-planned launch, tile ids >= *total_tiles do not exist
    int ahead;                    // 1 .. 3: how many tiles ahead the scheduler thread runs (launch-uniform)
    int group;                    // tiles per grid CTA (= per launch-control request): CTA c owns tiles c*group ...
    int64_t tiles;                // host-side tile count (a bound when total_tiles is set)
    // spatially packed sets: the tiles to visit (seg << 32 | I << 16 | J, any order; *total_tiles of them) and the
    // score rank of every slot; the bit of a pair goes into the row of the LOWER-ranked box
    const unsigned long long *tile_list;
    const uint32_t *rank;
    unsigned long long *tile_next;   // list mode: tiles handed out beyond the first one of every CTA (zeroed by the pack kernel)
    int64_t ahead_min_tiles;      // 8 tiles per resident CTA of this device: smaller grids keep look-ahead 1
};

__device__ __forceinline__ int segment_of_slot(const NmsSegment *__restrict__ segs, int nseg, int64_t slot);

// tile id -> (segment, I, J): tiles of a segment are numbered row block by row block,
// row block I holding the I + 1 tiles J = 0 .. I
__device__ TileInfo locate_tile(const NmsSegment *__restrict__ segs, int nseg, unsigned tile) {
    int lo = 0, hi = nseg - 1;  // last segment with tile_base <= tile
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (segs[mid].tile_base <= (int64_t)tile) lo = mid; else hi = mid - 1;
    }
    const int64_t t = (int64_t)tile - segs[lo].tile_base;
    // tiles before block-row i: i (i + 1) / 2
    int64_t I = (int64_t)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
    if (I < 0) I = 0;
    while ((I + 1) * (I + 2) / 2 <= t) ++I;
    while (I > 0 && I * (I + 1) / 2 > t) --I;
    TileInfo ti;
    ti.valid = 1;
    ti.seg = lo;
    ti.I = (int)I;
    ti.J = (int)(t - I * (I + 1) / 2);
    return ti;
}

// Spatially packed sets: of the C (C + 1) / 2 block pairs of a segment only those whose blocks' bounding
// boxes touch can hold an overlapping pair.  One thread per candidate pair; the survivors are appended to
// `list` (warp-aggregated atomic; the order is irrelevant).  C5: 175 k candidate tiles -> a few thousand.
__global__ void __launch_bounds__(256)
nms_tile_list_kernel(const NmsSegment *__restrict__ segs, int nseg_in, const long long *__restrict__ nseg_dev,
                     const int64_t *__restrict__ total_dev, int64_t tiles_bound, const float4 *__restrict__ aabb,
                     unsigned long long *__restrict__ list, unsigned long long *__restrict__ count) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t limit = total_dev ? min(*total_dev, tiles_bound) : tiles_bound;
    const int nseg = nseg_dev ? (int)*nseg_dev : nseg_in;
    const int lane = threadIdx.x & 31;
    bool keep = false;
    TileInfo ti;
    ti.valid = 0; ti.seg = 0; ti.I = 0; ti.J = 0;
    if (t < limit) {
        ti = locate_tile(segs, nseg, (unsigned)t);
        const int64_t b0 = segs[ti.seg].pstart / kTile;
        const float4 A = aabb[b0 + ti.I], B = aabb[b0 + ti.J];
        keep = ti.I == ti.J || (A.x <= B.z && B.x <= A.z && A.y <= B.w && B.y <= A.w);
    }
    unsigned bal = __ballot_sync(0xffffffffu, keep);
    // one atomic per CTA on the list length
    __shared__ unsigned int s_cnt[8];
    __shared__ unsigned long long s_base;
    const int warp = threadIdx.x >> 5;
    if (lane == 0) s_cnt[warp] = __popc(bal);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tot = 0;
        for (int w = 0; w < 8; ++w) tot += s_cnt[w];
        s_base = tot ? atomicAdd(count, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    if (bal == 0u) return;
    unsigned long long base = s_base;
    for (int w = 0; w < warp; ++w) base += s_cnt[w];
    if (keep)
        list[base + __popc(bal & ((1u << lane) - 1u))] =
            ((unsigned long long)ti.seg << 32) | ((unsigned long long)ti.I << 16) | (unsigned long long)ti.J;
}

#ifdef ROTGEO_ZERO_KERNEL   // probe builds only: the separate zeroing pass of the first list-mode version (A/B against the
                            // in-kernel zeroing one tile ahead: C5 282 -> 244 us of GPU time per call, profiles/r02_c5_zero_in_kernel.txt)
// The mask words of the listed tiles start from zero — BOTH word blocks of a pair of blocks, because a
// pair's bit lands in the row of its lower-ranked box, which may be the column box.  Zeroed here and not
// in the mask kernel: there any warp of a CTA may hit any row of the mirrored block while another warp's
// zero store is still on its way.  One warp per listed tile over a persistent grid (doing it inside the
// listing kernel left the stores to the few warps that held kept tiles: 57 us, 5 % of the issue slots).
__global__ void __launch_bounds__(256)
nms_tile_zero_kernel(const NmsSegment *__restrict__ segs, const unsigned long long *__restrict__ list,
                     const unsigned long long *__restrict__ count, unsigned long long *__restrict__ mask) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t total = (int64_t)*count;
    for (int64_t t = blockIdx.x * (int64_t)(blockDim.x >> 5) + (threadIdx.x >> 5); t < total; t += nwarps) {
        const unsigned long long v = list[t];
        const int I = (int)((v >> 16) & 0xffffull), J = (int)(v & 0xffffull);
        const NmsSegment &sg = segs[(int)(v >> 32)];
        const int64_t L = sg.len, W = sg.words;
        unsigned long long *const M = mask + sg.mask_base;
#pragma unroll
        for (int r = lane; r < kTile; r += 32) {
            const int64_t ri = (int64_t)I * kTile + r, rj = (int64_t)J * kTile + r;
            if (ri < L) *reinterpret_cast<uint4 *>(M + ri * W + 2 * J) = make_uint4(0u, 0u, 0u, 0u);
            if (I != J && rj < L) *reinterpret_cast<uint4 *>(M + rj * W + 2 * I) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
}
#endif

// AREA: the packed set carries the area axis of the index (high thresholds, see
// nms_area_axis_ratio): 3 ring stages of 7.7 KB instead of 4 of 5 KB, one more range lookup
template <bool AREA>
__global__ void __launch_bounds__(kThreads, 4)   // 4 CTAs per SM: at most 64 registers
nms_mask_kernel(const MaskArgs g) {
    __shared__ MaskSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kStages = AREA ? 3 : 4;
    constexpr uint32_t tab_words = AREA ? kTableWords : kTableXYWords;
    const int nseg = g.nseg_dev ? (int)*g.nseg_dev : g.nseg;
    const bool LIST = g.tile_list != nullptr;       // launch-uniform
    // Launches whose tile count is only known on the device (every fused front-end) run a persistent grid that
    // takes tiles from a global counter.  Launch-control stealing needs a grid sized for the host-side BOUND: one C4
    // image is 680 tiles under a bound of 55 945 (now 3 415), and cancelling the empty CTA ids one round trip at a
    // time was most of that kernel (ncu: 65 % of the samples in the response wait, 94 waves per SM).
    const bool COUNTER = g.tile_next != nullptr;
    if (LIST) pdl_wait();                           // the list and its length come from the kernel before this one

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], kWarps);
        }
#pragma unroll
        for (int s = 0; s < kInfoSlots; ++s) {
            mbar_init(&sm.sched[s], 1);
            mbar_init(&sm.ready[s], 1);
            mbar_init(&sm.zeroed[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();

    // thread 0 is the scheduler: my k-th tile's descriptor lives in info[k % kInfoSlots]
    // (published through ready[]), its launch-control response in resp[(k-1) % kInfoSlots]
    auto issue_cull_data = [&](unsigned k, const TileInfo &ti) {
        const int s = k % kStages;
        const NmsSegment &sg = g.segs[ti.seg];
        mbar_expect_tx(&sm.full[s], (kTile + tab_words) * sizeof(uint32_t));
        tma_load_1d(sm.rowr[s], g.pr + sg.pstart + (int64_t)ti.I * kTile, kTile * sizeof(uint32_t),
                    &sm.full[s]);
        tma_load_1d(sm.tab + s * tab_words, g.pt + (sg.pstart / kBlock + ti.J) * kTableWords,
                    tab_words * sizeof(uint32_t), &sm.full[s]);
    };
    // launch-control request number a = 1, 2, ... answers into resp[(a-1) % kInfoSlots]
    auto ask_next = [&](unsigned a) {
        const int slot = (a - 1) % kInfoSlots;
        mbar_expect_tx(&sm.sched[slot], 16);
        clc_try_cancel(&sm.resp[slot], &sm.sched[slot]);
    };
    // The scheduler runs D tiles ahead of the warps: while tile k is culled, the descriptor of
    // tile k + D is published and its cull data are in flight, with D launch-control requests
    // outstanding behind it.  A tile of a sparse scene is a few hundred instructions per warp,
    // far less than one launch-control or TMA round trip: one tile of look-ahead left every CTA
    // waiting for those two latencies in turn (ncu: 37 % of the instructions issued in the
    // barrier spin loops), and a warp 0 that was busy clipping delayed everybody's next tile.
    // Small grids keep D = 1: requests taken early are tiles the other CTAs cannot take.
    int ahead = g.ahead & 15;
    if (!(g.ahead & 16) && g.total_tiles && *g.total_tiles < g.ahead_min_tiles) ahead = 1;   // the real count, not the launch bound
    const int D = min(max(ahead, 1), kStages - 1);
    // One grid CTA (one launch-control request) stands for `group` consecutive tiles: the
    // hardware answers requests at a few ns each for the whole GPU, which is as long as an
    // empty tile takes.
    const unsigned G = (unsigned)g.group;
    const int64_t limit = g.total_tiles ? *g.total_tiles : g.tiles;
    // Inside a group the next tile is one step along the triangle: a few integer instructions
    // instead of a binary search over the segment table and a square root, which made thread 0
    // the bottleneck of sparse scenes.
    auto seg_blocks = [&](int sgi) {
        return sgi < nseg ? (int)((g.segs[sgi].len + kTile - 1) / kTile) : 0x7fffffff;
    };
    auto take_tile = [&](unsigned j) {      // thread 0: find my j-th tile, start its cull data, publish it
        TileInfo nx;
        nx.valid = 0; nx.seg = 0; nx.I = 0; nx.J = 0;
        if (LIST) {
            // spatially packed launches: a persistent grid strides over the tile list.  The grid bound of a
            // device-planned launch can exceed the real tile count 20-fold, and every empty CTA id would cost
            // one launch-control round trip to cancel (C5: 146 us for 9 k tiles); the list is in arbitrary
            // order.  Tiles after a CTA's first come from a global counter (taken D tiles ahead of use, so its
            // latency is hidden): dense scenes have tiles of very different cost, a static stride left SMs idle.
            const int64_t tn = (int64_t)gridDim.x + (int64_t)atomicAdd(g.tile_next, 1ull);
            (void)j;
            if (tn < limit) {
                const unsigned long long v = g.tile_list[tn];
                nx.valid = 1;
                nx.seg = (int)(v >> 32);
                nx.I = (int)((v >> 16) & 0xffffull);
                nx.J = (int)(v & 0xffffull);
                if (j >= (unsigned)kStages) mbar_wait(&sm.empty[j % kStages], ((j / kStages) - 1) & 1);
                issue_cull_data(j, nx);
            }
            sm.info[j % kInfoSlots] = nx;
            mbar_arrive(&sm.ready[j % kInfoSlots]);
            return;
        }
        MaskSmem::Sched st = sm.st;
        for (;;) {
            unsigned tn;
            bool step = false;
            if (st.grp_left > 0u) {
                tn = st.grp_next++;
                --st.grp_left;
                step = true;
            } else if (COUNTER) {
                if (!st.alive) break;
                const unsigned long long cta = (unsigned long long)gridDim.x + atomicAdd(g.tile_next, 1ull);
                if ((int64_t)(cta * G) >= limit) {       // the counter only grows: nothing left for anybody
                    st.alive = 0;
                    break;
                }
                tn = (unsigned)(cta * G);
                st.grp_next = tn + 1u;
                st.grp_left = G - 1u;
            } else {
                if (!st.alive) break;
                const unsigned a = st.next_resp++;
                mbar_wait(&sm.sched[(a - 1) % kInfoSlots], ((a - 1) / kInfoSlots) & 1);
                unsigned cta;
                if (!clc_decode(&sm.resp[(a - 1) % kInfoSlots], cta)) {
                    // nothing left to cancel: later requests fail as well; wait for their answers so
                    // that none is still on its way to this CTA's shared memory when it exits
                    st.alive = 0;
                    while (st.next_resp < st.next_ask) {
                        const unsigned b = st.next_resp++;
                        mbar_wait(&sm.sched[(b - 1) % kInfoSlots], ((b - 1) / kInfoSlots) & 1);
                    }
                    break;
                }
                fence_proxy_async_smem();   // my read of the slot precedes its reuse by a later request
                ask_next(st.next_ask++);
                tn = cta * G;
                st.grp_next = tn + 1u;
                st.grp_left = G - 1u;
            }
            if ((int64_t)tn < limit) {
                if (step) {             // the tile with the next id
                    if (++st.cur.J > st.cur.I) {
                        st.cur.J = 0;
                        ++st.cur.I;
                        while (st.cur.I >= st.cur_nb) {   // next segment (empty ones have no tiles)
                            ++st.cur.seg;
                            st.cur.I = 0;
                            st.cur_nb = seg_blocks(st.cur.seg);
                        }
                    }
                } else {
                    st.cur = locate_tile(g.segs, nseg, tn);
                    st.cur_nb = seg_blocks(st.cur.seg);
                }
                nx = st.cur;
                // ring stage of tile j was last read for tile j-kStages: wait until every warp
                // left it; they then also left descriptor / response slots that old
                if (j >= (unsigned)kStages) mbar_wait(&sm.empty[j % kStages], ((j / kStages) - 1) & 1);
                issue_cull_data(j, nx);
                break;
            }
            // past the (device-side) tile count: no work there, look further
        }
        sm.st = st;
        sm.info[j % kInfoSlots] = nx;
        mbar_arrive(&sm.ready[j % kInfoSlots]);
    };
    if (tid == 0) {
        TileInfo t0;
        t0.valid = 0; t0.seg = 0; t0.I = 0; t0.J = 0;
        MaskSmem::Sched st;
        st.next_ask = st.next_resp = 1u;
        st.grp_next = blockIdx.x * G + 1u;
        st.grp_left = G - 1u;
        st.alive = 1;
        st.cur_nb = 0;
        if ((int64_t)blockIdx.x * G < limit) {
            if (LIST) {
                const unsigned long long v = g.tile_list[blockIdx.x];      // G == 1 in list mode
                t0.valid = 1;
                t0.seg = (int)(v >> 32);
                t0.I = (int)((v >> 16) & 0xffffull);
                t0.J = (int)(v & 0xffffull);
            } else {
                t0 = locate_tile(g.segs, nseg, blockIdx.x * G);
                st.cur_nb = seg_blocks(t0.seg);
            }
        }
        st.cur = t0;
        sm.st = st;
        sm.info[0] = t0;
        mbar_arrive(&sm.ready[0]);
    }

    // Two per-warp queues whose fill counts are warp-uniform registers.  Survivors of the
    // bin ranges are pushed 32 at most per round (ballot ranks, no atomics); whenever 32
    // are queued they are popped for the circle test with every lane busy, and its
    // survivors go to the second queue, popped 32 at a time for the exact clipper.
    QEntry *const q1 = sm.queue1[warp];
    QEntry *const q2 = sm.queue2[warp];
    int n1 = 0, n2 = 0;
    const uint32_t lt = (1u << lane) - 1u;
    // List mode: the mask words of a listed tile — BOTH word blocks of the pair of blocks, a pair's bit lands in the
    // row of its lower-ranked box — are zeroed by the CTA that owns the tile, one tile ahead: at the end of tile k
    // every warp zeroes 16 rows of either block of tile k + 1 and arrives on zeroed[k + 1]; the first bit a warp sets
    // during tile k + 1 waits for that barrier (any warp may hit any row of either block).  Queue entries carried over
    // from earlier tiles were covered by earlier barriers.  A separate zeroing kernel before this one cost 46 us of
    // the 275 us C5 call, latency-bound; here the stores hide under the cull.
    unsigned kcur = 0;
    bool zero_ok = true;
    auto zero_tile = [&](const TileInfo &tz, unsigned kz) {
        const NmsSegment &sz = g.segs[tz.seg];
        const int64_t Lz = sz.len, Wz = sz.words;
        unsigned long long *const Mz = g.mask + sz.mask_base;
        const int rr = warp + kWarps * (lane & 15);
        if (lane < 16) {
            const int64_t ri = (int64_t)tz.I * kTile + rr;
            if (ri < Lz) *reinterpret_cast<uint4 *>(Mz + ri * Wz + 2 * tz.J) = make_uint4(0u, 0u, 0u, 0u);
        } else if (tz.I != tz.J) {
            const int64_t rj = (int64_t)tz.J * kTile + rr;
            if (rj < Lz) *reinterpret_cast<uint4 *>(Mz + rj * Wz + 2 * tz.I) = make_uint4(0u, 0u, 0u, 0u);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.zeroed[kz % kInfoSlots]);
    };
    auto clip32 = [&](int take) {            // pop `take` <= 32 entries of queue 2
        if (!zero_ok) {
            mbar_wait(&sm.zeroed[kcur % kInfoSlots], (kcur / kInfoSlots) & 1);
            zero_ok = true;
        }
        if (lane < take) {
            const QEntry e = q2[n2 - take + lane];
            const float v = rect_overlap<0>(__ldg(g.p0 + e.prow), __ldg(g.p1 + e.prow),
                                            __ldg(g.p0 + e.pcol), __ldg(g.p1 + e.pcol));
            if (v >= g.thr) {
                const int J = (int)(e.word >> 48);
                if (LIST && g.rank[e.prow] < g.rank[e.pcol]) {
                    // the row box is the better one: it suppresses the column box, the bit belongs to the
                    // column box's row (rare path: one segment lookup per suppressing pair)
                    const NmsSegment &sh = g.segs[segment_of_slot(g.segs, nseg, e.prow)];
                    const int64_t rpos = e.pcol - sh.pstart, cpos = e.prow - sh.pstart;
                    const int I = (int)(cpos >> 7);
                    atomicOr(g.mask + sh.mask_base + rpos * sh.words + (cpos >> 6), 1ull << (cpos & 63));
                    atomicOr(g.summary + (int64_t)e.pcol * g.sum_words + (I >> 6), 1ull << (I & 63));
                } else {
                    atomicOr(g.mask + (e.word & kWordMask), 1ull << (e.pcol & 63));
                    atomicOr(g.summary + (int64_t)e.prow * g.sum_words + (J >> 6), 1ull << (J & 63));
                }
            }
        }
        n2 -= take;
        __syncwarp();
    };
    auto circle32 = [&](int take) {          // pop `take` <= 32 entries of queue 1
        bool pass = false;
        QEntry e;
        if (lane < take) {
            e = q1[n1 - take + lane];
            const float4 A0 = __ldg(g.p0 + e.prow), B0 = __ldg(g.p0 + e.pcol);
            // IoU <= min(area) / max(area): with a high threshold (RPN: 0.8) most pairs of
            // different sizes end here, 1e-4 below the threshold to stay clear of rounding
            pass = !circles_disjoint(A0.x, A0.y, A0.z, B0.x, B0.y, B0.z) &&
                   fminf(A0.w, B0.w) >= g.area_ratio * fmaxf(A0.w, B0.w);
            if (AREA && pass) {
                // high thresholds (RPN: 0.8): IoU >= thr needs inter >= thr (areaA + areaB) / (1 + thr);
                // a cheap upper bound of the intersection (the smaller box clipped by the other's
                // bounding box in its frame) sends most circle survivors home before the ~170-instruction
                // clipper.  0.1 % below the requirement: clear of the rounding of either computation.
                const bool a_small = A0.w <= B0.w;
                const float4 S0 = a_small ? A0 : B0, L0 = a_small ? B0 : A0;
                const float4 S1 = __ldg(g.p1 + (a_small ? e.prow : e.pcol)), L1 = __ldg(g.p1 + (a_small ? e.pcol : e.prow));
                pass = intersection_upper_bound(S0, S1, L0, L1) >= g.inter_need * (A0.w + B0.w);
            }
        }
        n1 -= take;
        const uint32_t b = __ballot_sync(0xffffffffu, pass);
        if (pass) q2[n2 + __popc(b & lt)] = e;
        n2 += __popc(b);
        __syncwarp();
        if (n2 >= 32) clip32(32);
    };

    mbar_wait(&sm.ready[0], 0);
    TileInfo ti = sm.info[0];
    if (!ti.valid) return;       // launched past the device-side tile count
    const bool zero_here = LIST && g.zero_in_kernel;
    if (zero_here) zero_tile(ti, 0u);

    for (unsigned k = 0;; ++k) {
        kcur = k;
        zero_ok = !zero_here;
        const int s = k % kStages;
        const NmsSegment sg = g.segs[ti.seg];
        const int L = (int)sg.len, W = sg.words;
        const int row_base = ti.I * kTile, col_base = ti.J * kTile;   // positions inside the segment

        // ---- zeros of this warp's 16 rows of the tile (two mask words a row) ---------------
        if (!LIST && lane < 16) {      // (list mode: the tile-list kernel zeroed both word blocks of the tile)
            const int r = row_base + warp + kWarps * lane;
            if (r < L)
                *reinterpret_cast<uint4 *>(g.mask + sg.mask_base + (int64_t)r * W + 2 * ti.J) =
                    make_uint4(0u, 0u, 0u, 0u);
        }
        if (k == 0) {
            // everything below reads the packed set: as a programmatic dependent of the pack
            // kernel, this CTA's first zeros overlapped it
            pdl_wait();
            if (tid == 0) {
                issue_cull_data(0, ti);
                if (!COUNTER)
                    for (int e = 0; e < D; ++e) ask_next(sm.st.next_ask++);
                for (int j = 1; j <= D; ++j) take_tile((unsigned)j);   // incl. this iteration's duty
            }
        }

        // ---- scheduler: learn tile k+D, prefetch its cull data, keep D requests outstanding ---
        // The duty rotates over the warps (tile j: lane 0 of warp j % 8): one fixed scheduler
        // thread made its warp ~1.5x slower than the others, which then waited for it at every
        // tile (ncu: a third of all stall samples at the ready[] wait).  The state travels in
        // shared memory; ready[j-1] (arrive = release, wait = acquire) orders the hand-over.
        {
            const unsigned j = k + (unsigned)D;
            // (k == 0: thread 0 did it in the prologue; anybody else would race with its requests)
            if (k > 0 && lane == 0 && warp == (int)(j % kWarps)) {
                mbar_wait(&sm.ready[(j - 1) % kInfoSlots], ((j - 1) / kInfoSlots) & 1);
                take_tile(j);
            }
        }

        // ---- cull level 1: per (row, 64-column mask word) two range-OR lookups per axis ---------
        // lane = (row of the warp's sixteen, half of the 128 columns): the whole tile in one pass,
        // 8-byte table loads, and the range arithmetic of a row done by two lanes instead of four
        mbar_wait(&sm.full[s], (k / kStages) & 1);
        const int rl = lane >> 1, ch = lane & 1;
        const int rloc = warp + kWarps * rl;                        // row inside the tile
        unsigned long long m = 0ull;
        {
            const uint32_t rr = sm.rowr[s][rloc];
            const BinRanges br = unpack_ranges(rr);
            const int sx = br.sx, nx = br.nx, sy = br.sy, ny = br.ny;
            if (nx != 0) {
                const int jx = min(31 - __clz(nx), kTableLevels - 1), jy = min(31 - __clz(ny), kTableLevels - 1);
                const uint32_t *const tab = sm.tab + s * tab_words + 2 * ch;
                const uint32_t *tx = tab + jx * 128;
                const uint32_t *ty = tab + kTableAxisWords + jy * 128;
                const uint2 x0 = *reinterpret_cast<const uint2 *>(tx + sx * 4);
                const uint2 x1 = *reinterpret_cast<const uint2 *>(tx + ((sx + nx - (1 << jx)) & 31) * 4);
                const uint2 y0 = *reinterpret_cast<const uint2 *>(ty + sy * 4);
                const uint2 y1 = *reinterpret_cast<const uint2 *>(ty + ((sy + ny - (1 << jy)) & 31) * 4);
                uint32_t lo = (x0.x | x1.x) & (y0.x | y1.x), hi = (x0.y | x1.y) & (y0.y | y1.y);
                if (AREA && br.na != 0) {
                    // third axis: only columns whose area lies within the ratio that can still
                    // reach the threshold (IoU <= min(area) / max(area))
                    const int ja = min(31 - __clz(br.na), kTableLevels - 1);
                    const uint32_t *ta = tab + 2 * kTableAxisWords + ja * 128;
                    const uint2 a0 = *reinterpret_cast<const uint2 *>(ta + br.sa * 4);
                    const uint2 a1 = *reinterpret_cast<const uint2 *>(ta + ((br.sa + br.na - (1 << ja)) & 31) * 4);
                    lo &= a0.x | a1.x;
                    hi &= a0.y | a1.y;
                }
                m = (unsigned long long)lo | ((unsigned long long)hi << 32);
                if (ti.I == ti.J) {                                  // keep columns before the row only
                    const int d = rloc - 64 * ch;                    // row's bit inside this mask word
                    if (d <= 0) m = 0ull;
                    else if (d < 64) m &= (1ull << d) - 1ull;
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);   // this warp is done with the stage

        // (measured and dropped: walking a dense warp-tile row by row — the row's planes one broadcast load, lane l
        // testing column 32 g + l with one coalesced load, only the survivors compacted into queue 2 — instead of
        // queue 1.  A C4 warp-tile keeps 256 .. 512 of its 2048 pairs past the bins and the area axis: walked in
        // place the mask kernel of the 64-image batch took 540 us (threshold 256) and 812 us (96) against 350 us;
        // at 512 and above the path never ran on C4 and saved 3 % on one dense 30 000-box segment.  The compaction
        // of queue 1 is what keeps the lanes busy at these densities.)
        // ---- survivors -> queue 1, at most one per lane and round ---------------------------------
        {
            QEntry e;
            e.prow = (int)sg.pstart + row_base + rloc;
            e.word = (unsigned long long)(sg.mask_base + (int64_t)(row_base + rloc) * W + 2 * ti.J + ch) |
                     ((unsigned long long)ti.J << 48);
            const int pcol0 = (int)sg.pstart + col_base + 64 * ch;
            for (;;) {
                const uint32_t act = __ballot_sync(0xffffffffu, m != 0ull);
                if (act == 0u) break;
                if (m) {
                    e.pcol = pcol0 + __ffsll((long long)m) - 1;
                    m &= m - 1ull;
                    q1[n1 + __popc(act & lt)] = e;
                }
                n1 += __popc(act);
                __syncwarp();
                if (n1 >= 32) circle32(32);
            }
        }

        // ---- next tile (or none: nothing left to steal) ----------------------------------------
        mbar_wait(&sm.ready[(k + 1) % kInfoSlots], ((k + 1) / kInfoSlots) & 1);
        const TileInfo tnext = sm.info[(k + 1) % kInfoSlots];
        const bool more = tnext.valid != 0;
        if (zero_here && more) zero_tile(tnext, k + 1u);

        if (!more) {            // last tile of this CTA: flush both queues
            if (n1 > 0) circle32(n1);
            if (n2 > 0) clip32(n2);
        }
        if (!more) break;
        ti = tnext;
    }
}


// ---- greedy decision: parallel fixed point over (kept, removed) bitmaps -------------
constexpr int kResThreads = 1024;
constexpr int kResWarps = kResThreads / 32;
constexpr int kLocalChunks = 64;                       // segments of <= 4096 boxes are settled by ONE CTA,
                                                       // state in shared memory, __syncthreads between rounds
constexpr int kEmitItems = 8;                          // positions of `order` per thread and tile
constexpr int kEmitTile = kResThreads * kEmitItems;

struct ResolveArgs {
    const NmsSegment *segs;
    int nseg;
    const long long *nseg_dev;    // nullable: actual segment count
    const long long *n_pad_dev;   // nullable: slots actually used
    const unsigned int *error_dev;  // nullable: front-end error bits -> num_keep_out[1]
    int spatial;                  // rows may depend on boxes at LATER positions (spatially packed sets)
    int has_large;                // some segment may have more than kLocalChunks chunks
    const unsigned long long *mask;
    NmsScratch sc;
    int64_t chunks;               // 64-slot words to visit (an upper bound when device-planned)
    const int64_t *order;         // global score order: box indices ...
    const int64_t *seg_order;     // ... and segment-major order
    const uint32_t *order32;      // the same as 32-bit indices (fused front-ends); used when non-null
    const uint32_t *seg_order32;
    int64_t n;
    const int64_t *n_dev;         // nullable: overrides n
    uint8_t *flags;
    int64_t *keep_out, *num_keep_out;
    McOut mc;
};

__device__ __forceinline__ int64_t order_at(const ResolveArgs &a, int64_t i) {
    return a.order32 ? (int64_t)a.order32[i] : a.order[i];
}
__device__ __forceinline__ int64_t seg_order_at(const ResolveArgs &a, int64_t i) {
    return a.seg_order32 ? (int64_t)a.seg_order32[i] : a.seg_order[i];
}

// sync[0] arrival count, sync[1] generation, sync[2..4] "something is still undecided" per round % 3
__device__ __forceinline__ void grid_barrier(unsigned int *sync, unsigned int nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int *gen = sync + 1;
        const unsigned int g = *gen;
        __threadfence();
        if (atomicAdd(sync, 1u) == nblocks - 1u) {
            *(volatile unsigned int *)sync = 0u;
            __threadfence();
            *gen = g + 1u;
        } else {
            while (*gen == g) {}
        }
        __threadfence();
    }
    __syncthreads();
}

__device__ __forceinline__ ulonglong2 ldcg_u64x2(const unsigned long long *p) {
    ulonglong2 v;
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
// (kept, removed) words of one chunk: shared memory (SMEM) or global memory through L2
template <bool SMEM>
__device__ __forceinline__ ulonglong2 kr_load(const unsigned long long *p) {
    if (SMEM) {
        const volatile unsigned long long *v = p;
        return make_ulonglong2(v[0], v[1]);
    }
    return ldcg_u64x2(p);
}
template <bool SMEM>
__device__ __forceinline__ void kr_store(unsigned long long *p, unsigned long long k, unsigned long long r) {
    if (SMEM) {
        volatile unsigned long long *v = p;
        v[0] = k;
        v[1] = r;
    } else {
        asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(k), "l"(r) : "memory");
    }
}
__device__ __forceinline__ unsigned long long ballot64(bool lo, bool hi) {
    return (unsigned long long)__ballot_sync(0xffffffffu, lo) |
           ((unsigned long long)__ballot_sync(0xffffffffu, hi) << 32);
}

// segment that owns packed slot `slot` (last one with pstart <= slot)
__device__ __forceinline__ int segment_of_slot(const NmsSegment *__restrict__ segs, int nseg, int64_t slot) {
    int lo = 0, hi = nseg - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (segs[mid].pstart <= slot) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// One warp, one 64-box chunk (word c of its segment's mask rows), one round.  Lane l owns rows
// l and l + 32.  kr_others: where the state of the segment's OTHER chunks is read (shared
// memory when RD_SMEM: the one-CTA path, or the grid path's per-round snapshot); kr_mine: the
// live (kept, removed) pairs of the segment, read and written for chunk c only.  Returns true
// while the chunk still has an undecided box.
// Both bitmaps only ever gain bits, and a reader that sees an old word merely treats a decided
// box as undecided for one more round: chunks can be settled concurrently, in any order.
template <bool RD_SMEM, bool WR_SMEM>
__device__ bool resolve_chunk(const ResolveArgs &a, const NmsSegment &sg, int c,
                              const unsigned long long *kr_others, unsigned long long *kr_mine, int lane) {
    const int64_t pos0 = (int64_t)c * 64, len = sg.len;
    const unsigned long long valid = len - pos0 >= 64 ? ~0ull : ((1ull << (len - pos0)) - 1ull);
    const ulonglong2 kr0 = kr_load<WR_SMEM>(kr_mine + 2 * c);
    unsigned long long Kc = kr0.x, Rc = kr0.y;
    const unsigned long long und = valid & ~(Kc | Rc);
    if (und == 0ull) return false;

    const int W = sg.words;
    const unsigned long long *const M = a.mask + sg.mask_base;
    const unsigned long long *const S = a.sc.summary + (sg.pstart + pos0) * a.sc.sum_words;
    unsigned long long D[2];
    bool exK[2], exU[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int q = lane + 32 * p;
        D[p] = 0ull;
        exK[p] = exU[p] = false;
        if ((und >> q) & 1ull) {
            const int64_t row = pos0 + q;
            const unsigned long long *mrow = M + row * W;
            const unsigned long long *srow = S + q * a.sc.sum_words;
            // summary words that can hold a bit: column blocks 0 .. row / 128 when packed in score order
            const int kmax = a.spatial ? (sg.words / 2 - 1) >> 6 : (int)(row >> 13);
            for (int k = 0; k <= kmax && !exK[p]; ++k) {
                // (plain loads: this row's summary word is rewritten below)
                const unsigned long long sb0 = __ldcg(srow + k);
                unsigned long long sb = sb0, live = sb0;
                while (sb != 0ull && !exK[p]) {
                    // four column blocks at a time: their mask loads are in flight together
                    int J[4];
                    ulonglong2 m[4];
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        J[t] = sb ? (k << 6) + __ffsll((long long)sb) - 1 : -1;
                        sb &= sb - 1ull;     // 0 stays 0
                    }
#pragma unroll
                    for (int t = 0; t < 4; ++t)
                        m[t] = J[t] >= 0 ? __ldg(reinterpret_cast<const ulonglong2 *>(mrow + 2 * J[t]))
                                         : make_ulonglong2(0ull, 0ull);
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        bool dead = J[t] >= 0;       // every neighbour in this block is removed
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const unsigned long long mv = h ? m[t].y : m[t].x;
                            const int w = 2 * J[t] + h;
                            if (mv == 0ull) continue;
                            if (w != c) {                     // another chunk (an earlier one when packed in score order): its state
                                const ulonglong2 o = kr_load<RD_SMEM>(kr_others + 2 * w);
                                if (mv & o.x) exK[p] = true;
                                if (mv & ~(o.x | o.y)) exU[p] = true;
                                if (mv & ~o.y) dead = false;
                            } else if (w == c) {
                                D[p] = mv;                    // earlier boxes of this chunk
                                dead = false;
                            }
                        }
                        // removed boxes stay removed: this block can never again decide the row,
                        // later rounds need not load it (dense scenes: a row has neighbours in a
                        // hundred blocks and used to re-walk all of them every round)
                        if (!WR_SMEM && dead) live &= ~(1ull << (J[t] & 63));
                    }
                }
                if (!WR_SMEM && live != sb0 && !exK[p]) __stcg(const_cast<unsigned long long *>(srow + k), live);
            }
        }
    }
    const unsigned long long EXK = ballot64(exK[0], exK[1]);
    const unsigned long long EXU = ballot64(exU[0], exU[1]);
    Rc |= und & EXK;
    unsigned long long cand = und & ~EXK;
    // chains inside the chunk: peel until nothing changes
    for (;;) {
        bool rm[2], kp[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int q = lane + 32 * p;
            const bool in = (cand >> q) & 1ull;
            rm[p] = in && (D[p] & Kc) != 0ull;
            kp[p] = in && !rm[p] && (D[p] & cand) == 0ull && !((EXU >> q) & 1ull);
        }
        const unsigned long long REM = ballot64(rm[0], rm[1]);
        const unsigned long long KEEP = ballot64(kp[0], kp[1]);
        if ((REM | KEEP) == 0ull) break;
        Kc |= KEEP;
        Rc |= REM;
        cand &= ~(REM | KEEP);
    }
    if (lane == 0 && (Kc != kr0.x || Rc != kr0.y)) kr_store<WR_SMEM>(kr_mine + 2 * c, Kc, Rc);
    return cand != 0ull;
}

// ---- the same evaluation by a TEAM of 8 warps (the grid path of large segments) -------------------
// One warp per chunk walks its rows' column blocks one lane = two rows, block after block: on a dense
// 30 000-box segment the first evaluation of a chunk took 12 .. 31 us and there are ten more before
// the last chunk is decided, while nine warps in ten of the grid have no chunk at all
// (profiles/r02_resolve_team.txt).  Here a chunk belongs to 256 threads: four lanes per row, each
// taking four of the row's next sixteen column blocks, so sixteen mask loads of a row are in flight at
// once and a lane has one row, not two.  The lanes of a row agree through shuffles inside their group of
// four; the rows of the eight warps meet in shared memory, and every warp of the team then finishes
// the chunk redundantly (same inputs, same result; warp 0 stores).
constexpr int kTeamWarps = 8;
constexpr int kTeams = kResWarps / kTeamWarps;
struct TeamSmem {
    unsigned long long D[64];        // in-chunk neighbour word per row
    uint32_t exk[kTeamWarps], exu[kTeamWarps];
};
__device__ __forceinline__ void team_bar(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(kTeamWarps * 32) : "memory");
}
// bits 0, 4, 8 .. 28 -> bits 0 .. 7
__device__ __forceinline__ uint32_t every_fourth_bit(uint32_t b) {
    uint32_t x = b & 0x11111111u;
    x = (x | (x >> 3)) & 0x03030303u;
    x = (x | (x >> 6)) & 0x000f000fu;
    return (x | (x >> 12)) & 0xffu;
}
__device__ bool resolve_chunk_team(const ResolveArgs &a, const NmsSegment &sg, int c, unsigned long long *kr,
                                   TeamSmem &ts, int team, int tw, int lane) {
    const int64_t pos0 = (int64_t)c * 64, len = sg.len;
    const unsigned long long valid = len - pos0 >= 64 ? ~0ull : ((1ull << (len - pos0)) - 1ull);
    // the chunk's own words are written by this team only, and every store is followed by a team barrier:
    // all eight warps read the same value here, so the early return is taken by all or none
    const ulonglong2 kr0 = ldcg_u64x2(kr + 2 * c);
    unsigned long long Kc = kr0.x, Rc = kr0.y;
    const unsigned long long und = valid & ~(Kc | Rc);
    if (und == 0ull) return false;

    const int W = sg.words;
    const unsigned long long *const M = a.mask + sg.mask_base;
    const unsigned long long *const S = a.sc.summary + (sg.pstart + pos0) * a.sc.sum_words;
    const int q = tw * 8 + (lane >> 2), sub = lane & 3;
    const unsigned gmask = 0xfu << (lane & 28);
    // (measured and dropped: one coalesced copy of the segment's (kept, removed) words into shared memory per
    // evaluation instead of an L2 lookup per non-empty mask word — evaluations got no shorter, the staler state cost
    // one or two more of them: 85 -> 100 us on the dense 30 000-box segment.  The mask loads are the latency.)
    unsigned long long Dq = 0ull;
    int exK = 0, exU = 0;
    if ((und >> q) & 1ull) {              // the same for the four lanes of the row
        const int64_t row = pos0 + q;
        const unsigned long long *mrow = M + row * W;
        unsigned long long *srow = const_cast<unsigned long long *>(S) + q * a.sc.sum_words;
        const int kmax = a.spatial ? (sg.words / 2 - 1) >> 6 : (int)(row >> 13);
        for (int k = 0; k <= kmax; ++k) {
            const unsigned long long sb0 = __ldcg(srow + k);
            unsigned long long sb = sb0, live = sb0;
            while (sb != 0ull) {          // sixteen set bits per round: sb is the same in the four lanes at this point
                for (int i = 0; i < 4 * sub; ++i) sb &= sb - 1ull;
                int J[4];
                ulonglong2 m[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    J[t] = sb ? (k << 6) + __ffsll((long long)sb) - 1 : -1;
                    sb &= sb - 1ull;
                }
                for (int i = 0; i < 4 * (3 - sub); ++i) sb &= sb - 1ull;
#pragma unroll
                for (int t = 0; t < 4; ++t)
                    m[t] = J[t] >= 0 ? __ldg(reinterpret_cast<const ulonglong2 *>(mrow + 2 * J[t]))
                                     : make_ulonglong2(0ull, 0ull);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    bool dead = J[t] >= 0;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const unsigned long long mv = h ? m[t].y : m[t].x;
                        const int w = 2 * J[t] + h;
                        if (mv == 0ull) continue;
                        if (w != c) {
                            const ulonglong2 o = ldcg_u64x2(kr + 2 * w);
                            if (mv & o.x) exK = 1;
                            if (mv & ~(o.x | o.y)) exU = 1;
                            if (mv & ~o.y) dead = false;
                        } else {
                            Dq = mv;
                            dead = false;
                        }
                    }
                    if (dead) live &= ~(1ull << (J[t] & 63));
                }
            }
            live &= __shfl_xor_sync(gmask, live, 1);
            live &= __shfl_xor_sync(gmask, live, 2);
            exK |= __shfl_xor_sync(gmask, exK, 1);
            exK |= __shfl_xor_sync(gmask, exK, 2);
            if (sub == 0 && live != sb0 && !exK) __stcg(srow + k, live);
            if (exK) break;               // a kept neighbour: the row is removed, nothing else matters
        }
        Dq |= __shfl_xor_sync(gmask, Dq, 1);
        Dq |= __shfl_xor_sync(gmask, Dq, 2);
        exU |= __shfl_xor_sync(gmask, exU, 1);
        exU |= __shfl_xor_sync(gmask, exU, 2);
    }
    __syncwarp();
    const uint32_t bk = __ballot_sync(0xffffffffu, exK != 0), bu = __ballot_sync(0xffffffffu, exU != 0);
    if (lane == 0) {
        ts.exk[tw] = every_fourth_bit(bk);
        ts.exu[tw] = every_fourth_bit(bu);
    }
    if (sub == 0) ts.D[q] = Dq;
    team_bar(team);
    unsigned long long EXK = 0ull, EXU = 0ull;
#pragma unroll
    for (int w = 0; w < kTeamWarps; ++w) {
        EXK |= (unsigned long long)ts.exk[w] << (8 * w);
        EXU |= (unsigned long long)ts.exu[w] << (8 * w);
    }
    const unsigned long long D[2] = {ts.D[lane], ts.D[lane + 32]};
    Rc |= und & EXK;
    unsigned long long cand = und & ~EXK;
    for (;;) {
        bool rm[2], kp[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int r = lane + 32 * p;
            const bool in = (cand >> r) & 1ull;
            rm[p] = in && (D[p] & Kc) != 0ull;
            kp[p] = in && !rm[p] && (D[p] & cand) == 0ull && !((EXU >> r) & 1ull);
        }
        const unsigned long long REM = ballot64(rm[0], rm[1]);
        const unsigned long long KEEP = ballot64(kp[0], kp[1]);
        if ((REM | KEEP) == 0ull) break;
        Kc |= KEEP;
        Rc |= REM;
        cand &= ~(REM | KEEP);
    }
    if (tw == 0 && lane == 0 && (Kc != kr0.x || Rc != kr0.y)) kr_store<false>(kr + 2 * c, Kc, Rc);
    team_bar(team);    // the store precedes every later read of these words by the team; ts is free again
    return cand != 0ull;
}

// kept bits of one chunk -> per-box flags (indexed like the caller's boxes)
__device__ __forceinline__ void write_flags(const ResolveArgs &a, const NmsSegment &sg, int c,
                                            unsigned long long K, int lane) {
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int q = lane + 32 * p;
        const int64_t pos = (int64_t)c * 64 + q;
        if (pos < sg.len) a.flags[seg_order_at(a, sg.start + pos)] = (uint8_t)((K >> q) & 1ull);
    }
}

__device__ __forceinline__ int block_sum(int v, int *s_red /* kResWarps */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < kResWarps; ++w) t += s_red[w];
    return t;
}

#ifdef ROTGEO_RESOLVE_TIMING   // probe builds: phase timestamps of CTA 0 (ns, %globaltimer), printed at the end
#define RT_DECL unsigned long long rt_[12]; int rtn_ = 0;
#define RT_MARK do { if (blockIdx.x == 0 && threadIdx.x == 0 && rtn_ < 12) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(rt_[rtn_])); ++rtn_; } } while (0)
#define RT_PRINT do { if (blockIdx.x == 0 && threadIdx.x == 0) { printf("resolve phases (us):"); for (int i_ = 1; i_ < rtn_; ++i_) printf(" %.1f", (double)(rt_[i_] - rt_[i_ - 1]) * 1e-3); printf("\n"); } } while (0)
#else
#define RT_DECL
#define RT_MARK do {} while (0)
#define RT_PRINT do {} while (0)
#endif

template <bool MC>
__global__ void __launch_bounds__(kResThreads, 1)
nms_resolve_kernel(const ResolveArgs a) {
    // device-planned launches: the real counts (a copy of the argument struct with them patched in put the
    // whole struct into local memory)
    const int nseg = a.nseg_dev ? (int)*a.nseg_dev : a.nseg;
    const int64_t chunks = a.n_pad_dev ? min(a.chunks, (int64_t)(*a.n_pad_dev >> 6)) : a.chunks;
    RT_DECL
    RT_MARK;   // 0 start
    // dynamic shared memory: the (kept, removed) pairs of one small segment
    extern __shared__ __align__(16) unsigned long long s_kr[];
    __shared__ int s_red[kResWarps];
    __shared__ int s_warp_cnt[kResWarps];
    __shared__ uint32_t s_list[kEmitTile];   // kept indices of one emit tile
    __shared__ TeamSmem s_team[kTeams];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int G = gridDim.x;

    // ---- small segments: one CTA each, state in shared memory ------------------------------
    for (int s = blockIdx.x; s < nseg; s += G) {
        const NmsSegment sg = a.segs[s];
        const int nch = sg.row_blocks;
        if (sg.len <= 0 || nch > kLocalChunks) continue;
        for (int i = tid; i < 2 * nch; i += kResThreads) s_kr[i] = 0ull;
        __syncthreads();
        // No barrier between the passes: both bitmaps only gain bits and any evaluation order
        // reaches the greedy result, so every warp simply re-evaluates its own chunks until they
        // are decided; a decision is visible to the warps behind it a shared-memory round trip
        // later instead of one __syncthreads round later (a chain over c chunks needed c rounds).
        for (;;) {
            bool left = false;
            for (int c = warp; c < nch; c += kResWarps)
                left |= resolve_chunk<true, true>(a, sg, c, s_kr, s_kr, lane);
            if (!left) break;
        }
        __syncthreads();
        for (int c = warp; c < nch; c += kResWarps) write_flags(a, sg, c, s_kr[2 * c], lane);
        __syncthreads();
    }

    RT_MARK;   // 1 small segments decided + flags
    // ---- large segments: their chunks spread over the whole grid -----------------------------
    if (a.has_large) {
        // consecutive chunks go to different CTAs: a mid-size segment is spread over every SM
        const int64_t gwarp = (int64_t)blockIdx.x + (int64_t)G * warp, nwarps = (int64_t)G * kResWarps;
        const int team = warp / kTeamWarps, tw = warp % kTeamWarps;
        const int64_t gteam = (int64_t)blockIdx.x + (int64_t)G * team, nteams = (int64_t)G * kTeams;
        // No rounds and no grid barrier here either: every team re-evaluates its own chunks until
        // they are decided, reading the state of the other chunks through L2 (ld.cg / st.cg).  All
        // CTAs are co-resident (cooperative launch), the bitmaps only gain bits, so this ends with
        // the greedy result; a decision travels one L2 round trip instead of one grid-barrier
        // round (~10 us with 148 CTAs) to the chunks that wait for it.
#ifdef ROTGEO_RESOLVE_TIMING
        unsigned long long pc_t0, pc_first = 0, pc_last = 0;
        int pc_evals = 0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pc_t0));
#endif
        // Teams pay when there are fewer chunks than warps and the rows are long (one dense 30 000-box segment:
        // 470 chunks, 155 -> 85 us); with thousands of short-row chunks (C5: 3125, C4 batch: 10 000) a warp per
        // chunk keeps eight times as many chunks in flight and is the faster of the two (C5 35 vs 74 us).
        const bool teams = chunks <= 2 * nteams;
        for (; !teams;) {
            bool left = false;
            for (int64_t gw = gwarp; gw < chunks; gw += nwarps) {
                const NmsSegment &sg = a.segs[segment_of_slot(a.segs, nseg, gw * 64)];
                const int64_t c = gw - (sg.pstart >> 6), w0 = 2 * (sg.pstart >> 6);
                if (sg.row_blocks <= kLocalChunks || c >= sg.row_blocks) continue;
                left |= resolve_chunk<false, false>(a, sg, (int)c, a.sc.kr + w0, a.sc.kr + w0, lane);
            }
            if (!left) break;
        }
        for (; teams;) {
            bool left = false;           // the same in the eight warps of a team
            for (int64_t gw = gteam; gw < chunks; gw += nteams) {
                const NmsSegment &sg = a.segs[segment_of_slot(a.segs, nseg, gw * 64)];
                const int64_t c = gw - (sg.pstart >> 6), w0 = 2 * (sg.pstart >> 6);
                if (sg.row_blocks <= kLocalChunks || c >= sg.row_blocks) continue;
                left |= resolve_chunk_team(a, sg, (int)c, a.sc.kr + w0, s_team[team], team, tw, lane);
#ifdef ROTGEO_RESOLVE_TIMING
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pc_last));
                if (pc_evals++ == 0) pc_first = pc_last;
#endif
            }
            if (!left) break;
        }
#ifdef ROTGEO_RESOLVE_TIMING   // per-chunk: evaluations until decided, duration of the first one, time of the decision
        if (lane == 0 && tw == 0 && gteam < chunks && gteam % 41 == 0)
            printf("  chunk %lld: %d evaluations, first %.1f us, decided at %.1f us\n", (long long)gteam, pc_evals,
                   (double)(pc_first - pc_t0) * 1e-3, (double)(pc_last - pc_t0) * 1e-3);
#endif
        __syncthreads();                 // every team of the CTA is done before the flags of its chunks are written
        for (int64_t gw = gwarp; gw < chunks; gw += nwarps) {
            const NmsSegment &sg = a.segs[segment_of_slot(a.segs, nseg, gw * 64)];
            const int64_t c = gw - (sg.pstart >> 6);
            if (sg.row_blocks <= kLocalChunks || c >= sg.row_blocks) continue;
            write_flags(a, sg, (int)c, ldcg_u64x2(a.sc.kr + 2 * gw).x, lane);
        }
    }
    RT_MARK;   // 2 large segments
    grid_barrier(a.sc.sync, G);
    RT_MARK;   // 3 barrier

    // ---- ordered compaction in global score order: count, barrier, write ----------------------
    const int64_t n = a.n_dev ? *a.n_dev : a.n;
    const int64_t cap = MC ? a.mc.cap : n;
    // equal shares of the order for all CTAs: the flag gathers are random 32-byte sectors, and one
    // SM's L2 bandwidth made a single CTA with a full 8192-position tile the slowest part of the
    // small calls (count 4.6 us, emit 12 us on C2 with everything in CTA 0)
    // (measured again in round 2 and dropped: for jobs of at most two emit tiles, let the LAST CTA to finish its
    // segments compact alone — no grid barrier, no counting pass.  Resolve kernel 24.3 -> 26.8 us on C2 sparse and
    // 19.7 -> 53.8 us on a C4 image: one SM gathering 10 000 flags and 2 000 rows is slower than both barriers.)
    const int64_t per = (n + G - 1) / G;
    const int64_t lo = min(n, (int64_t)blockIdx.x * per), hi = min(n, lo + per);
    {
        // (kEmitItems independent index loads, then as many independent flag loads per thread)
        int cnt = 0;
        for (int64_t start = lo; start < hi; start += kEmitTile) {
            const int64_t i0 = start + (int64_t)tid * kEmitItems;
            int64_t idx[kEmitItems];
#pragma unroll
            for (int k = 0; k < kEmitItems; ++k) idx[k] = i0 + k < hi ? order_at(a, i0 + k) : -1;
#pragma unroll
            for (int k = 0; k < kEmitItems; ++k)
                if (idx[k] >= 0) cnt += __ldcg(a.flags + idx[k]) != 0;
        }
        cnt = block_sum(cnt, s_red);
        if (tid == 0) a.sc.block_counts[blockIdx.x] = cnt;
    }
    RT_MARK;   // 4 count
    grid_barrier(a.sc.sync, G);
    RT_MARK;   // 5 barrier
    int before = 0, total = 0;
    for (unsigned int b = tid; b < G; b += kResThreads) {
        const int cb = __ldcg(a.sc.block_counts + b);
        total += cb;
        if (b < blockIdx.x) before += cb;
    }
    before = block_sum(before, s_red);
    total = block_sum(total, s_red);
    if (blockIdx.x == 0 && tid == 0) {
        a.num_keep_out[0] = (int64_t)total < cap ? (int64_t)total : cap;
        if (a.error_dev) a.num_keep_out[1] = (int64_t)*a.error_dev;
    }

    RT_MARK;   // 6 offsets
    int64_t base = before;
    for (int64_t start = lo; start < hi && base < cap; start += kEmitTile) {
        const int64_t i0 = start + (int64_t)tid * kEmitItems;
        int64_t idx[kEmitItems];
        unsigned f = 0;
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k) idx[k] = i0 + k < hi ? order_at(a, i0 + k) : -1;
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k)
            if (idx[k] >= 0 && __ldcg(a.flags + idx[k]) != 0) f |= 1u << k;
        const int cnt = __popc(f);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        __syncthreads();
        if (lane == 31) s_warp_cnt[warp] = incl;
        __syncthreads();
        int wbefore = 0, wtotal = 0;
#pragma unroll
        for (int w = 0; w < kResWarps; ++w) {
            const int cw = s_warp_cnt[w];
            wbefore += (w < warp) ? cw : 0;
            wtotal += cw;
        }
        // the tile's kept indices, compacted in shared memory; then ONE output row per thread and
        // step: the gathers of a row (box, score) are independent loads of different threads
        // instead of eight dependent chains inside one thread (C2: emit 16 -> 4 us), and the
        // stores are coalesced
        int local = wbefore + incl - cnt;
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k)
            if (f & (1u << k)) s_list[local++] = (uint32_t)idx[k];
        __syncthreads();
        for (int j = tid; j < wtotal; j += kResThreads) {
            const int64_t pos = base + j;
            if (pos >= cap) break;
            const uint32_t id = s_list[j];
            a.keep_out[pos] = (int64_t)id;
            if (MC) {
                float *d = a.mc.dets_out + pos * 6;
                if (a.mc.classes == 0) {
                    // plain rows: cat(boxes[keep], scores[keep]) of mmcv's nms_rotated / batched_nms
                    const float *bx = a.mc.boxes + (int64_t)id * a.mc.box_stride;
                    d[0] = bx[0]; d[1] = bx[1]; d[2] = bx[2]; d[3] = bx[3]; d[4] = bx[4];
                    d[5] = a.mc.scores[id];
                } else {
                    // the output row of a kept candidate: its box, its (factored) score, its class
                    const uint32_t row = id / (uint32_t)a.mc.classes;
                    const int cls = (int)(id - row * (uint32_t)a.mc.classes);
                    const float *bx = a.mc.boxes + (a.mc.shared ? (int64_t)row : (int64_t)id) * 5;
                    float sc = a.mc.scores[(int64_t)row * a.mc.score_stride + cls];
                    if (a.mc.factors) sc *= a.mc.factors[row];
                    d[0] = bx[0]; d[1] = bx[1]; d[2] = bx[2]; d[3] = bx[3]; d[4] = bx[4]; d[5] = sc;
                    a.mc.labels_out[pos] = cls;
                }
            }
        }
        base += wtotal;
    }
    RT_MARK;   // 7 emit
    RT_PRINT;
}

// ---- host-side plan + launch -------------------------------------------------------
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

void nms_plan(int64_t n, const int64_t *seg_offsets_host, int64_t num_segments, NmsPlan *plan,
              NmsSegment *segments_out) {
    plan->n = n;
    int64_t one[2] = {0, n};
    if (seg_offsets_host == nullptr || num_segments <= 0) {
        seg_offsets_host = one;
        num_segments = 1;
    }
    plan->num_segments = num_segments;
    int64_t tiles = 0, words = 0, slots = 0, max_blocks = 0;
    for (int64_t s = 0; s < num_segments; ++s) {
        const int64_t start = seg_offsets_host[s], len = seg_offsets_host[s + 1] - start;
        const int64_t C = (len + kTile - 1) / kTile;
        if (segments_out) {
            NmsSegment &g = segments_out[s];
            g.start = start;
            g.pstart = slots;
            g.len = len;
            g.mask_base = words;
            g.tile_base = tiles;
            g.words = (int32_t)(2 * C);
            g.row_blocks = (int32_t)((len + 63) / 64);
        }
        slots += C * kTile;
        tiles += C * (C + 1) / 2;
        words += len * 2 * C;
        max_blocks = C > max_blocks ? C : max_blocks;
    }
    plan->n_pad = slots > 0 ? slots : kTile;   // every segment padded to whole 128-slot blocks
    plan->total_tiles = tiles;
    plan->mask_words = words;
    plan->sum_words = nms_sum_words(max_blocks);
    plan->max_chunks = 2 * max_blocks;
    size_t off = 0;
    plan->off_packed = off;   off = align_up(off + (size_t)plan->n_pad * kPackedBytesPerBox, 256);
    plan->off_segments = off; off = align_up(off + (size_t)num_segments * sizeof(NmsSegment), 256);
    plan->off_mask = off;     off = align_up(off + (size_t)words * sizeof(uint64_t), 256);
    plan->off_flags = off;    off = align_up(off + (size_t)n, 256);
    plan->off_kr = off;       off = align_up(off + (size_t)(plan->n_pad / 64) * 16, 256);
    plan->off_summary = off;  off = align_up(off + (size_t)plan->n_pad * plan->sum_words * 8, 256);
    plan->off_sync = off;     off = align_up(off + kNmsSyncWords * sizeof(unsigned int), 256);
    plan->off_counts = off;   off = align_up(off + kNmsMaxBlocks * sizeof(int), 256);
    plan->total_bytes = off;
}

// Per-device launch limits, cached per device (not per process): SM count and the number of
// resident CTAs the mask kernel's grid-size heuristics are expressed in.
struct DeviceLimits {
    int sms;          // multiprocessors
    int mask_slots;   // resident mask-kernel CTAs: 4 per SM (__launch_bounds__(kThreads, 4))
};
static int device_limits(DeviceLimits *out) {
    constexpr int kMaxDevices = 64;
    static int cached_sms[kMaxDevices];   // 0 = not queried yet; benign race: every writer stores the same value
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    int sms = (dev >= 0 && dev < kMaxDevices) ? cached_sms[dev] : 0;
    if (sms == 0) {
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) return -1;
        if (dev >= 0 && dev < kMaxDevices) cached_sms[dev] = sms;
    }
    out->sms = sms;
    out->mask_slots = sms * 4;
    return 0;
}

// look-ahead of the mask kernel's scheduler thread (MaskArgs::ahead): grids of fewer than 8 tiles
// per resident CTA keep 1 (requests taken early are tiles the other CTAs cannot take).
constexpr int kAheadDefault = 3;
#ifdef ROTGEO_PROBES   // probe builds only: ROTGEO_MASK_AHEAD=1..3 / ROTGEO_MASK_GROUP=1..64 override the heuristics
static int probe_env(const char *name, int lo, int hi) {
    const char *e = getenv(name);
    const int v = e ? atoi(e) : 0;
    return (v < lo || v > hi) ? 0 : v;
}
#endif

// tiles per launch-control request (MaskArgs::group): 4 for very large grids (measured on C5,
// 175k mostly empty tiles: 0.88 -> 0.80 ms), else 1 (C4, 43k dense tiles: groups cost balance)
static int mask_group(int64_t tiles, const DeviceLimits &lim) {
#ifdef ROTGEO_PROBES
    if (const int forced = probe_env("ROTGEO_MASK_GROUP", 1, 64)) return forced;
#endif
    return tiles >= (int64_t)lim.mask_slots * 128 ? 4 : 1;
}

static int launch_mask(const MaskArgs &g_in, int64_t tiles, cudaStream_t stream) {
    if (tiles > 0x7fffffffLL) return -1;
    if (tiles <= 0) return 0;
    MaskArgs g = g_in;
    DeviceLimits lim;
    if (device_limits(&lim)) return -3;
    g.ahead_min_tiles = (int64_t)lim.mask_slots * 8;
    g.ahead = tiles >= g.ahead_min_tiles ? kAheadDefault : 1;
#ifdef ROTGEO_PROBES
    if (const int forced = probe_env("ROTGEO_MASK_AHEAD", 1, 3)) g.ahead = forced + 16;   // + 16: also for small grids
#endif
    g.tiles = tiles;
    g.group = g.tile_list ? 1 : mask_group(tiles, lim);
    int64_t ctas = (tiles + g.group - 1) / g.group;
    if (g.tile_next && ctas > lim.mask_slots) ctas = lim.mask_slots;      // persistent: one CTA per resident slot
    // one CTA per tile; running CTAs steal the not-yet-launched ones (launch control)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;   // overlap with the pack kernel
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = g.area_axis ? cudaLaunchKernelEx(&cfg, nms_mask_kernel<true>, g)
                                      : cudaLaunchKernelEx(&cfg, nms_mask_kernel<false>, g);
    return e == cudaSuccess ? 0 : -3;
}

static int launch_resolve(const ResolveArgs &a_in, cudaStream_t stream) {
    const bool mc = a_in.mc.dets_out != nullptr;
    DeviceLimits lim;
    if (device_limits(&lim)) return -3;
    // all CTAs co-resident (cooperative launch): one 1024-thread CTA per SM
    const int limit = lim.sms < kNmsMaxBlocks ? lim.sms : kNmsMaxBlocks;
    // one small segment per CTA; with large segments, their chunks spread over every SM (the row
    // tests are latency-bound: more SMs, more loads in flight)
    // (measured and dropped: sending mid-size segments — a C2-dense class, 32 chunks — to the grid's 8-warp teams as
    // well instead of one CTA each changed nothing: resolve 48.3 vs 47.2 us on C2 dense, 19.9 vs 20.0 on a C4 image)
    int64_t want = a_in.has_large ? a_in.chunks : a_in.nseg;
#ifndef ROTGEO_RESOLVE_SHARE
#define ROTGEO_RESOLVE_SHARE 512
#endif
    // the compaction shares the score order out over the grid: with one CTA per small segment only (15 for the C2
    // classes) every CTA gathered 2000 flags and rows; at least one CTA per ROTGEO_RESOLVE_SHARE positions
    if (ROTGEO_RESOLVE_SHARE > 0 && want < a_in.n / ROTGEO_RESOLVE_SHARE) want = a_in.n / ROTGEO_RESOLVE_SHARE;
    if (want < 1) want = 1;
    ResolveArgs a = a_in;
    // dynamic shared memory: one small segment's state
    const size_t dyn = 2 * kLocalChunks * sizeof(unsigned long long);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(want < limit ? want : limit));
    cfg.blockDim = dim3(kResThreads);
    cfg.dynamicSmemBytes = dyn;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;              // the grid barrier needs every CTA resident
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = mc ? cudaLaunchKernelEx(&cfg, nms_resolve_kernel<true>, a)
                             : cudaLaunchKernelEx(&cfg, nms_resolve_kernel<false>, a);
    return e == cudaSuccess ? 0 : -3;
}

int launch_nms(const NmsPlan &plan, const void *packed, const NmsSegment *segments_dev,
               const int64_t *order, const int64_t *seg_order, const uint32_t *order32,
               const uint32_t *seg_order32, float thr, uint64_t *mask, uint8_t *flags,
               const NmsScratch &scratch, int64_t *keep_out, int64_t *num_keep_out,
               cudaStream_t stream) {
    const PackedView P = packed_view(packed, plan.n_pad);
    MaskArgs g;
    g.p0 = P.p0; g.p1 = P.p1; g.pr = P.pr; g.pt = P.pt;
    g.segs = segments_dev;
    g.nseg = (int)plan.num_segments;
    g.thr = thr;
    g.area_ratio = thr - 1e-4f;
    g.inter_need = 0.999f * thr / (1.f + thr);
    g.area_axis = nms_area_axis_ratio(thr) > 0.f;
    g.mask = reinterpret_cast<unsigned long long *>(mask);
    g.summary = scratch.summary;
    g.sum_words = scratch.sum_words;
    g.total_tiles = nullptr;
    g.nseg_dev = nullptr;
    g.tile_list = nullptr;
    g.zero_in_kernel = 0;
    g.rank = nullptr;
    g.tile_next = nullptr;
    int rc = launch_mask(g, plan.total_tiles, stream);
    if (rc) return rc;
    ResolveArgs a = {};
    a.segs = segments_dev;
    a.nseg = (int)plan.num_segments;
    a.mask = g.mask;
    a.sc = scratch;
    a.has_large = plan.max_chunks > kLocalChunks;
    a.chunks = plan.n_pad / 64;
    a.order = order;
    a.seg_order = seg_order;
    a.order32 = order32;
    a.seg_order32 = seg_order32;
    a.n = plan.n;
    a.n_dev = nullptr;
    a.flags = flags;
    a.keep_out = keep_out;
    a.num_keep_out = num_keep_out;
    return launch_resolve(a, stream);
}

int launch_nms_device_planned(const NmsDevLaunch &d, cudaStream_t stream) {
    const PackedView P = packed_view(d.packed, d.n_pad_bound);
    MaskArgs g;
    g.p0 = P.p0; g.p1 = P.p1; g.pr = P.pr; g.pt = P.pt;
    g.segs = d.segs_dev;
    g.nseg = d.nseg;
    g.thr = d.thr;
    g.area_ratio = d.thr - 1e-4f;
    g.inter_need = 0.999f * d.thr / (1.f + d.thr);
    g.area_axis = nms_area_axis_ratio(d.thr) > 0.f;
    g.mask = reinterpret_cast<unsigned long long *>(d.mask);
    g.summary = d.scratch.summary;
    g.sum_words = d.scratch.sum_words;
    g.total_tiles = d.total_tiles_dev;
    g.nseg_dev = d.nseg_dev;
    g.tile_list = nullptr;
    g.zero_in_kernel = 0;
    g.rank = nullptr;
    // the tile count lives on the device: persistent grid + global tile counter (zeroed by the pack kernel)
    g.tile_next = reinterpret_cast<unsigned long long *>(d.scratch.sync + 10);
    if (d.tile_list) {
        // spatially packed: list the block pairs whose bounding boxes touch, then visit only those
        if (d.tiles_bound > 0x7fffffffLL) return -1;
        unsigned long long *count = reinterpret_cast<unsigned long long *>(d.scratch.sync + 8);   // zeroed by the pack kernel
        if (d.tiles_bound > 0) {
            nms_tile_list_kernel<<<(unsigned)((d.tiles_bound + 255) / 256), 256, 0, stream>>>(
                d.segs_dev, d.nseg, d.nseg_dev, d.total_tiles_dev, d.tiles_bound, d.aabb, d.tile_list, count);
#ifdef ROTGEO_ZERO_KERNEL   // probe builds: the separate zeroing pass of the first list-mode version
            DeviceLimits zl;
            if (device_limits(&zl)) return -3;
            nms_tile_zero_kernel<<<(unsigned)(zl.sms * 8), 256, 0, stream>>>(
                d.segs_dev, d.tile_list, count, reinterpret_cast<unsigned long long *>(d.mask));
#endif
            if (cudaGetLastError() != cudaSuccess) return -3;
        }
        g.tile_list = d.tile_list;
#ifndef ROTGEO_ZERO_KERNEL
        g.zero_in_kernel = 1;
#endif
        g.rank = d.rank;
        g.tile_next = reinterpret_cast<unsigned long long *>(d.scratch.sync + 10);
        g.total_tiles = reinterpret_cast<const int64_t *>(count);
    }
    int rc = launch_mask(g, d.tiles_bound, stream);
    if (rc) return rc;
    ResolveArgs a = {};
    a.segs = d.segs_dev;
    a.nseg = d.nseg;
    a.nseg_dev = d.nseg_dev;
    a.n_pad_dev = d.n_pad_dev;
    a.error_dev = d.error_dev;
    a.spatial = d.tile_list != nullptr;
    a.mask = g.mask;
    a.sc = d.scratch;
    a.has_large = d.max_chunks_bound > kLocalChunks;
    a.chunks = d.n_pad_bound / 64;
    a.order32 = d.order32;
    a.seg_order32 = d.seg_order32;
    a.n = d.n_bound;
    a.n_dev = d.n_dev;
    a.flags = d.flags;
    a.keep_out = d.keep_out;
    a.num_keep_out = d.num_keep_out;
    a.mc = d.mc;
    return launch_resolve(a, stream);
}

}  // namespace rotgeo
