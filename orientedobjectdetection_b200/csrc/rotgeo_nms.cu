// Segmented greedy rotated NMS, entirely on the device.  sm_100a.
//
// Input: boxes packed in SEGMENT-MAJOR order (segment = class / FPN level /
// image x level ...), score-descending inside each segment, every segment
// starting at a packed slot that is a multiple of 128.  Three kernels:
//
//   nms_mask_kernel  the suppression bit matrix, block-diagonal and upper-triangular:
//                    one work item per 128x128 tile of the upper triangle of each
//                    segment's pair matrix (pairs from different segments are never
//                    touched).  Same machinery as the IoU kernel: CTAs made persistent
//                    by cluster-launch-control work stealing, cull data (row bin
//                    ranges + the column block's inverted bin index) TMA-staged in a
//                    shared-memory ring, tile zeros stored first, surviving pairs
//                    queued per warp, circle-tested 32 at a time, re-queued and clipped
//                    32 at a time; `IoU >= thr` becomes an atomicOr into the (already
//                    zeroed) mask word.
//   nms_scan_kernel  one CTA per segment, blocked right-looking greedy scan.  A panel
//                    of 512 rows x 8 words around the diagonal is staged in shared
//                    memory (cp.async, double-buffered); its eight 64-box chunks are
//                    resolved there — inside a chunk the kept set is found by peeling
//                    (warp redux.or rounds, usually 1-3) instead of a 64-step serial
//                    loop — and only then are the kept rows OR-ed into the `removed`
//                    bitmap right of the panel in one bulk pass over global memory.
//                    The mask never leaves the GPU.
//   nms_emit_kernel  (several segments only) compacts the kept flags in GLOBAL score
//                    order into keep_out — the order mmcv.ops.nms_rotated returns and
//                    every reference caller slices with [:max_num].
//
// Replaces mmcv's nms_rotated (64x64 bitmask over the full square + D2H copy of
// the mask + serial host scan; reference call sites
// mmrotate/core/post_processing/bbox_nms_rotated.py:80,122,
// mmrotate/core/patch/merge_results.py:116-118, mmrotate/datasets/dota.py:379).
#include <math.h>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

constexpr int kTile = 128;      // tile edge in boxes = one inverted-index block = two mask words
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kStages = 4;      // cull-data ring
constexpr int kInfoSlots = 8;   // tile descriptors / launch-control responses (> kStages)
constexpr int kQueueCap = 64;   // pairs per warp queue (never more than 31 + 32 queued)
static_assert(kTile == kBlock, "a tile column is one inverted-index block");

struct TileInfo {
    int valid;        // 0: nothing left to steal
    int seg;          // segment index
    int I, J;         // row / column block inside the segment (I <= J)
};

struct QEntry {       // one pair that survived cull level 1
    int prow, pcol;               // packed slots
    unsigned long long word;      // index of the mask word holding its bit
};

struct __align__(16) MaskSmem {
    uint32_t tab[kStages][kTableWords];
    uint32_t rowr[kStages][kTile];
    QEntry queue1[kWarps][kQueueCap];   // survived the bin ranges
    QEntry queue2[kWarps][kQueueCap];   // survived the circle test as well
    uint4 resp[kInfoSlots];
    TileInfo info[kInfoSlots];
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t sched[kInfoSlots];
    uint64_t ready[kInfoSlots];
};

struct MaskArgs {
    const float4 *p0, *p1;
    const uint32_t *pr, *pt;
    const NmsSegment *segs;
    int nseg;
    float thr;
    unsigned long long *mask;
    const int64_t *total_tiles;   // nullable: device-planned launch, tile ids >= *total_tiles do not exist
};

// tile id -> (segment, I, J): tiles of a segment are numbered row block by row block
__device__ TileInfo locate_tile(const NmsSegment *__restrict__ segs, int nseg, unsigned tile) {
    int lo = 0, hi = nseg - 1;  // last segment with tile_base <= tile
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (segs[mid].tile_base <= (int64_t)tile) lo = mid; else hi = mid - 1;
    }
    const int64_t t = (int64_t)tile - segs[lo].tile_base;
    const int C = segs[lo].words >> 1;  // 128-box column blocks
    // tiles before block-row i: i*C - i*(i-1)/2
    int I = (int)(((2.0 * C + 1.0) - sqrt((2.0 * C + 1.0) * (2.0 * C + 1.0) - 8.0 * (double)t)) * 0.5);
    I = max(0, min(I, C - 1));
    while (I + 1 < C && (int64_t)(I + 1) * C - (int64_t)(I + 1) * I / 2 <= t) ++I;
    while (I > 0 && (int64_t)I * C - (int64_t)I * (I - 1) / 2 > t) --I;
    TileInfo ti;
    ti.valid = 1;
    ti.seg = lo;
    ti.I = I;
    ti.J = I + (int)(t - ((int64_t)I * C - (int64_t)I * (I - 1) / 2));
    return ti;
}

__global__ void __launch_bounds__(kThreads)
nms_mask_kernel(const MaskArgs g) {
    __shared__ MaskSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], kWarps);
        }
#pragma unroll
        for (int s = 0; s < kInfoSlots; ++s) {
            mbar_init(&sm.sched[s], 1);
            mbar_init(&sm.ready[s], 1);
        }
        fence_mbar_init();
    }
    __syncthreads();

    // thread 0 is the scheduler: my k-th tile's descriptor lives in info[k % kInfoSlots]
    // (published through ready[]), its launch-control response in resp[(k-1) % kInfoSlots]
    auto issue_cull_data = [&](unsigned k, const TileInfo &ti) {
        const int s = k % kStages;
        const NmsSegment &sg = g.segs[ti.seg];
        mbar_expect_tx(&sm.full[s], (kTile + kTableWords) * sizeof(uint32_t));
        tma_load_1d(sm.rowr[s], g.pr + sg.pstart + (int64_t)ti.I * kTile, kTile * sizeof(uint32_t),
                    &sm.full[s]);
        tma_load_1d(sm.tab[s], g.pt + (sg.pstart / kBlock + ti.J) * kTableWords,
                    kTableWords * sizeof(uint32_t), &sm.full[s]);
    };
    auto ask_next = [&](unsigned k) {
        const int slot = (k - 1) % kInfoSlots;
        mbar_expect_tx(&sm.sched[slot], 16);
        clc_try_cancel(&sm.resp[slot], &sm.sched[slot]);
    };
    if (tid == 0) {
        TileInfo t0;
        t0.valid = 0; t0.seg = 0; t0.I = 0; t0.J = 0;
        if (!g.total_tiles || (int64_t)blockIdx.x < *g.total_tiles)
            t0 = locate_tile(g.segs, g.nseg, blockIdx.x);
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
    auto clip32 = [&](int take) {            // pop `take` <= 32 entries of queue 2
        if (lane < take) {
            const QEntry e = q2[n2 - take + lane];
            const float v = rect_overlap<0>(__ldg(g.p0 + e.prow), __ldg(g.p1 + e.prow),
                                            __ldg(g.p0 + e.pcol), __ldg(g.p1 + e.pcol));
            if (v >= g.thr) atomicOr(g.mask + e.word, 1ull << (e.pcol & 63));
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
            pass = !circles_disjoint(A0.x, A0.y, A0.z, B0.x, B0.y, B0.z);
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

    for (unsigned k = 0;; ++k) {
        const int s = k % kStages;
        const NmsSegment sg = g.segs[ti.seg];
        const int L = (int)sg.len, W = sg.words;
        const int row_base = ti.I * kTile, col_base = ti.J * kTile;   // positions inside the segment

        // ---- zeros of this warp's 16 rows of the tile (two mask words a row) ---------------
        if (lane < 16) {
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
                ask_next(1);
            }
        }

        // ---- scheduler: learn tile k+1, prefetch its cull data, ask for k+2, publish ---------
        if (tid == 0) {
            const unsigned kn = k + 1;
            mbar_wait(&sm.sched[k % kInfoSlots], (k / kInfoSlots) & 1);
            unsigned tn;
            TileInfo nx;
            nx.valid = 0; nx.seg = 0; nx.I = 0; nx.J = 0;
            if (clc_decode(&sm.resp[k % kInfoSlots], tn) &&
                (!g.total_tiles || (int64_t)tn < *g.total_tiles)) {
                nx = locate_tile(g.segs, g.nseg, tn);
                // ring stage of tile k+1 was last read for tile k+1-kStages: wait until every
                // warp left it; they then also left descriptor / response slots that old
                if (kn >= kStages) mbar_wait(&sm.empty[kn % kStages], ((kn / kStages) - 1) & 1);
                issue_cull_data(kn, nx);
                fence_proxy_async_smem();
                ask_next(k + 2);
            }
            sm.info[kn % kInfoSlots] = nx;
            mbar_arrive(&sm.ready[kn % kInfoSlots]);
        }

        // ---- cull level 1: per (row, 32-column word) two range-OR lookups per axis -----------
        mbar_wait(&sm.full[s], (k / kStages) & 1);
        uint32_t hits[2];
        const int rl = lane >> 2, cw = lane & 3;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int rloc = warp + kWarps * (rl + 8 * p);          // row inside the tile
            const uint32_t rr = sm.rowr[s][rloc];
            const int sx = rr & 31, nx = (rr >> 8) & 63, sy = (rr >> 16) & 31, ny = rr >> 24;
            uint32_t h = 0;
            if (nx != 0) {
                const int jx = min(31 - __clz(nx), kTableLevels - 1), jy = min(31 - __clz(ny), kTableLevels - 1);
                const uint32_t *tx = sm.tab[s] + jx * 128 + cw;
                const uint32_t *ty = sm.tab[s] + kTableAxisWords + jy * 128 + cw;
                const uint32_t X = tx[sx * 4] | tx[((sx + nx - (1 << jx)) & 31) * 4];
                const uint32_t Y = ty[sy * 4] | ty[((sy + ny - (1 << jy)) & 31) * 4];
                h = X & Y;
                if (ti.I == ti.J) {                                  // keep columns after the row only
                    const int d = rloc - 32 * cw;                    // row's bit inside this word
                    if (d >= 31) h = 0u;
                    else if (d >= 0) h &= ~((2u << d) - 1u);
                }
            }
            hits[p] = h;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);   // this warp is done with the stage

        // ---- survivors -> queue 1, at most one per lane and round ---------------------------------
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            uint32_t m = hits[p];
            const int rloc = warp + kWarps * (rl + 8 * p);
            QEntry e;
            e.prow = (int)sg.pstart + row_base + rloc;
            e.word = (unsigned long long)(sg.mask_base + (int64_t)(row_base + rloc) * W + 2 * ti.J + (cw >> 1));
            const int pcol0 = (int)sg.pstart + col_base + 32 * cw;
            for (;;) {
                const uint32_t act = __ballot_sync(0xffffffffu, m != 0u);
                if (act == 0u) break;
                if (m) {
                    e.pcol = pcol0 + __ffs(m) - 1;
                    m &= m - 1;
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

        if (!more) {            // last tile of this CTA: flush both queues
            if (n1 > 0) circle32(n1);
            if (n2 > 0) clip32(n2);
        }
        if (!more) break;
        ti = tnext;
    }
}

// ---- greedy scan -----------------------------------------------------------------
constexpr int kScanThreads = 512;
constexpr int kPanelChunks = 8;                       // 64-box chunks per panel
constexpr int kPanelRows = 64 * kPanelChunks;         // 512
constexpr int kPanelWords = kPanelRows * kPanelChunks;  // u64 words per panel buffer (32 KB)

__device__ __forceinline__ unsigned long long warp_or64(unsigned long long v) {
    const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
    const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// dynamic shared memory: two panel buffers, then the `removed` bitmap (seg.words entries)
template <bool EMIT>
__global__ void __launch_bounds__(kScanThreads)
nms_scan_kernel(const NmsSegment *__restrict__ segs, const unsigned long long *__restrict__ mask,
                const int64_t *__restrict__ seg_order, int64_t idx_mask, uint8_t *__restrict__ flags,
                int64_t *__restrict__ keep_out, int64_t *__restrict__ num_keep_out) {
    extern __shared__ __align__(16) unsigned long long dyn[];
    unsigned long long *const panel0 = dyn;
    unsigned long long *const removed = dyn + 2 * kPanelWords;
    __shared__ int s_rows[kPanelRows];   // kept rows of the current panel (segment positions)
    __shared__ int s_nk_chunk, s_nk_panel;

    const NmsSegment seg = segs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = seg.words;
    const int L = (int)seg.len;
    const unsigned long long *__restrict__ M = mask + seg.mask_base;
    const int panels = (seg.row_blocks + kPanelChunks - 1) / kPanelChunks;

    // panel P: rows [512 P, 512 P + 512) x words [8 P, 8 P + 8), 64 B per row
    auto load_panel = [&](int P, unsigned long long *buf) {
        const int r0 = P * kPanelRows, w0 = P * kPanelChunks;
        for (int idx = tid; idx < kPanelRows * 4; idx += kScanThreads) {
            const int r = idx >> 2, q = idx & 3;           // row, 16-byte quarter
            if (r0 + r < L && w0 + 2 * q < W)
                cp_async16(buf + r * kPanelChunks + 2 * q, M + (int64_t)(r0 + r) * W + w0 + 2 * q);
        }
        cp_async_commit();
    };
    load_panel(0, panel0);

    for (int w = tid; w < W; w += kScanThreads) {
        // positions >= L do not exist: mark them removed
        const int first = w * 64;
        unsigned long long v = 0ull;
        if (first >= L) v = ~0ull;
        else if (first + 64 > L) v = ~0ull << (L - first);
        removed[w] = v;
    }

    int64_t nkeep = 0;
    for (int P = 0; P < panels; ++P) {
        unsigned long long *const buf = panel0 + (P & 1) * kPanelWords;
        if (P + 1 < panels) {
            load_panel(P + 1, panel0 + ((P + 1) & 1) * kPanelWords);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        if (tid == 0) s_nk_panel = 0;
        __syncthreads();

        const int r0 = P * kPanelRows, w0 = P * kPanelChunks;
        const int chunks_here = min(kPanelChunks, seg.row_blocks - w0);
        for (int cl = 0; cl < chunks_here; ++cl) {
            const int c = w0 + cl;
            if (warp == 0) {
                // lane holds the diagonal words of rows c*64+lane and c*64+32+lane
                const int q0 = cl * 64 + lane, q1 = q0 + 32;       // rows inside the panel
                const unsigned long long d0 = r0 + q0 < L ? buf[q0 * kPanelChunks + cl] : 0ull;
                const unsigned long long d1 = r0 + q1 < L ? buf[q1 * kPanelChunks + cl] : 0ull;
                unsigned long long cand = ~removed[c];
                unsigned long long kept = 0ull;
                while (cand) {
                    const bool in0 = (cand >> lane) & 1ull, in1 = (cand >> (lane + 32)) & 1ull;
                    // everything some undecided box could still suppress
                    const unsigned long long threat = warp_or64((in0 ? d0 : 0ull) | (in1 ? d1 : 0ull));
                    const unsigned long long safe = cand & ~threat;   // nobody left can suppress these
                    kept |= safe;
                    const bool s0 = (safe >> lane) & 1ull, s1 = (safe >> (lane + 32)) & 1ull;
                    const unsigned long long killed = warp_or64((s0 ? d0 : 0ull) | (s1 ? d1 : 0ull));
                    cand &= ~(safe | killed);
                }
                const bool k0 = (kept >> lane) & 1ull, k1 = (kept >> (lane + 32)) & 1ull;
                const int rk0 = __popcll(kept & ((1ull << lane) - 1ull));
                const int rk1 = __popcll(kept & ((1ull << (lane + 32)) - 1ull));
                const int base = s_nk_panel;
                if (k0) s_rows[base + rk0] = r0 + q0;
                if (k1) s_rows[base + rk1] = r0 + q1;
                if (EMIT) {
                    if (k0) keep_out[nkeep + rk0] = seg_order[seg.start + r0 + q0] & idx_mask;
                    if (k1) keep_out[nkeep + rk1] = seg_order[seg.start + r0 + q1] & idx_mask;
                } else {
                    if (r0 + q0 < L) flags[seg_order[seg.start + r0 + q0] & idx_mask] = k0 ? 1 : 0;
                    if (r0 + q1 < L) flags[seg_order[seg.start + r0 + q1] & idx_mask] = k1 ? 1 : 0;
                }
                __syncwarp();
                if (lane == 0) {
                    s_nk_chunk = __popcll(kept);
                    s_nk_panel = base + __popcll(kept);
                }
            }
            __syncthreads();
            const int nk = s_nk_chunk;
            nkeep += nk;
            // kept rows of this chunk -> removed bits of the panel's remaining words (shared memory only)
            const int wrem = chunks_here - (cl + 1);
            if (wrem > 0 && nk > 0) {
                const int base = s_nk_panel - nk;
                for (int it = tid; it < nk * wrem; it += kScanThreads) {
                    const int kk = it / wrem, j = cl + 1 + (it - kk * wrem);
                    const unsigned long long v = buf[(s_rows[base + kk] - r0) * kPanelChunks + j];
                    if (v) atomicOr(&removed[w0 + j], v);
                }
            }
            __syncthreads();
        }
        // kept rows of the whole panel -> removed bits right of the panel, from global memory:
        // every thread takes (kept row, word) items, a row's words by consecutive threads
        const int wfar0 = w0 + kPanelChunks, wfar = W - wfar0;
        const int nkp = s_nk_panel;
        if (wfar > 0 && nkp > 0) {
            const int items = nkp * wfar;
#pragma unroll 4
            for (int it = tid; it < items; it += kScanThreads) {
                const int kk = it / wfar, wi = it - kk * wfar;
                const unsigned long long v = M[(int64_t)s_rows[kk] * W + wfar0 + wi];
                if (v) atomicOr(&removed[wfar0 + wi], v);
            }
        }
        __syncthreads();
    }
    if (EMIT && tid == 0) *num_keep_out = nkeep;
}

// ---- ordered compaction of the kept flags (several segments) ---------------------
// One CTA; each thread takes 8 consecutive positions of `order` per round, so a
// round covers 8192 boxes with two block-wide barriers.
constexpr int kEmitThreads = 1024;
constexpr int kEmitItems = 8;

template <bool MC>
__global__ void __launch_bounds__(kEmitThreads)
nms_emit_kernel(const int64_t *__restrict__ order, const uint8_t *__restrict__ flags, int64_t n,
                const int64_t *__restrict__ n_dev, int64_t idx_mask, int64_t *__restrict__ keep_out,
                int64_t *__restrict__ num_keep_out, const McOut mc) {
    __shared__ int warp_cnt[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (n_dev) n = *n_dev;
    const int64_t cap = MC ? mc.cap : n;
    int64_t base = 0;
    for (int64_t start = 0; start < n && base < cap; start += kEmitThreads * kEmitItems) {
        const int64_t i0 = start + (int64_t)tid * kEmitItems;
        int64_t idx[kEmitItems];
        unsigned f = 0;
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k) {
            idx[k] = i0 + k < n ? (order[i0 + k] & idx_mask) : -1;
        }
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k)
            if (idx[k] >= 0 && flags[idx[k]] != 0) f |= 1u << k;
        const int cnt = __popc(f);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) warp_cnt[warp] = incl;
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 32; ++w) {
            const int c = warp_cnt[w];
            before += (w < warp) ? c : 0;
            total += c;
        }
        int64_t pos = base + before + incl - cnt;
#pragma unroll
        for (int k = 0; k < kEmitItems; ++k) {
            if ((f & (1u << k)) && pos < cap) {
                keep_out[pos] = idx[k];
                if (MC) {
                    // the output row of a kept candidate: its box, its (factored) score, its class
                    const int64_t row = idx[k] / mc.classes;
                    const int cls = (int)(idx[k] - row * mc.classes);
                    const float *bx = mc.boxes + (mc.shared ? row : idx[k]) * 5;
                    float sc = mc.scores[row * mc.score_stride + cls];
                    if (mc.factors) sc *= mc.factors[row];
                    float *d = mc.dets_out + pos * 6;
                    d[0] = bx[0]; d[1] = bx[1]; d[2] = bx[2]; d[3] = bx[3]; d[4] = bx[4]; d[5] = sc;
                    mc.labels_out[pos] = cls;
                }
            }
            if (f & (1u << k)) ++pos;
        }
        base += total;
        __syncthreads();
    }
    if (tid == 0) *num_keep_out = base < cap ? base : cap;
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
    int64_t tiles = 0, words = 0, slots = 0;
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
    }
    plan->n_pad = slots > 0 ? slots : kTile;   // every segment padded to whole 128-slot blocks
    plan->total_tiles = tiles;
    plan->mask_words = words;
    size_t off = 0;
    plan->off_packed = off;   off = align_up(off + (size_t)plan->n_pad * kPackedBytesPerBox, 256);
    plan->off_segments = off; off = align_up(off + (size_t)num_segments * sizeof(NmsSegment), 256);
    plan->off_mask = off;     off = align_up(off + (size_t)words * sizeof(uint64_t), 256);
    plan->off_flags = off;    off = align_up(off + (size_t)n, 256);
    plan->off_seg_keep = off;
    plan->total_bytes = off;
}

size_t nms_scan_smem_bytes(int max_words) {
    return (size_t)(2 * kPanelWords + max_words) * sizeof(unsigned long long);
}

int launch_nms(const NmsPlan &plan, const void *packed, const NmsSegment *segments_dev,
               int max_words, const int64_t *order, const int64_t *seg_order, float thr,
               uint64_t *mask, uint8_t *flags, int64_t *keep_out, int64_t *num_keep_out,
               cudaStream_t stream) {
    const PackedView P = packed_view(packed, plan.n_pad);
    if (plan.total_tiles > 0x7fffffffLL) return -1;
    if (plan.total_tiles > 0) {
        MaskArgs g;
        g.p0 = P.p0; g.p1 = P.p1; g.pr = P.pr; g.pt = P.pt;
        g.segs = segments_dev;
        g.nseg = (int)plan.num_segments;
        g.thr = thr;
        g.mask = reinterpret_cast<unsigned long long *>(mask);
        g.total_tiles = nullptr;
        // one CTA per tile; running CTAs steal the not-yet-launched ones (launch control)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)plan.total_tiles);
        cfg.blockDim = dim3(kThreads);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;   // overlap with the pack kernel
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, nms_mask_kernel, g) != cudaSuccess) return -3;
    }
    if (cudaGetLastError() != cudaSuccess) return -3;
    const size_t dyn = nms_scan_smem_bytes(max_words);
    const bool single = plan.num_segments == 1;
    if (single) {
        if (cudaFuncSetAttribute(nms_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)dyn) != cudaSuccess)
            return -3;
        nms_scan_kernel<true><<<1, kScanThreads, dyn, stream>>>(
            segments_dev, reinterpret_cast<const unsigned long long *>(mask), seg_order, ~0ll, flags,
            keep_out, num_keep_out);
    } else {
        if (cudaFuncSetAttribute(nms_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)dyn) != cudaSuccess)
            return -3;
        nms_scan_kernel<false><<<(unsigned)plan.num_segments, kScanThreads, dyn, stream>>>(
            segments_dev, reinterpret_cast<const unsigned long long *>(mask), seg_order, ~0ll, flags,
            keep_out, num_keep_out);
        if (cudaGetLastError() != cudaSuccess) return -3;
        McOut none = {};
        nms_emit_kernel<false><<<1, kEmitThreads, 0, stream>>>(order, flags, plan.n, nullptr, ~0ll,
                                                              keep_out, num_keep_out, none);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int launch_nms_device_planned(const NmsDevLaunch &a, cudaStream_t stream) {
    const PackedView P = packed_view(a.packed, a.n_pad_bound);
    if (a.tiles_bound > 0x7fffffffLL) return -1;
    if (a.tiles_bound > 0) {
        MaskArgs g;
        g.p0 = P.p0; g.p1 = P.p1; g.pr = P.pr; g.pt = P.pt;
        g.segs = a.segs_dev;
        g.nseg = a.nseg;
        g.thr = a.thr;
        g.mask = reinterpret_cast<unsigned long long *>(a.mask);
        g.total_tiles = a.total_tiles_dev;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)a.tiles_bound);
        cfg.blockDim = dim3(kThreads);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;   // overlap with the pack kernel
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, nms_mask_kernel, g) != cudaSuccess) return -3;
    }
    const size_t dyn = nms_scan_smem_bytes(a.max_words_bound);
    if (cudaFuncSetAttribute(nms_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)dyn) != cudaSuccess)
        return -3;
    nms_scan_kernel<false><<<(unsigned)a.nseg, kScanThreads, dyn, stream>>>(
        a.segs_dev, reinterpret_cast<const unsigned long long *>(a.mask), a.seg_order, a.idx_mask,
        a.flags, a.keep_out, a.num_keep_out);
    if (cudaGetLastError() != cudaSuccess) return -3;
    if (a.mc.dets_out)
        nms_emit_kernel<true><<<1, kEmitThreads, 0, stream>>>(a.order, a.flags, a.n_bound, a.n_dev,
                                                             a.idx_mask, a.keep_out, a.num_keep_out, a.mc);
    else
        nms_emit_kernel<false><<<1, kEmitThreads, 0, stream>>>(a.order, a.flags, a.n_bound, a.n_dev,
                                                              a.idx_mask, a.keep_out, a.num_keep_out, a.mc);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
