// Pairwise rotated IoU / IoF matrix on packed boxes.  sm_100a.
//
// The grid has one CTA (8 warps) per 64-row x 128-column output tile, but a CTA
// that is running keeps going: through cluster launch control it cancels the
// launch of the next not-yet-started CTA and takes over its tile, so the CTAs
// are persistent (barriers initialised once, masks prefetched, warp queues kept
// across tiles) while the hardware scheduler still balances the load.  Per tile:
//   0. the bin ranges of the tile's 64 row boxes (256 B) and the inverted bin index
//      of its 128-column block (5 KB sparse table, built once at pack time) arrive
//      in a 4-stage shared-memory ring filled by TMA bulk copies one tile ahead
//      (full/empty mbarriers; no __syncthreads anywhere in the loop);
//   1. warp w owns rows 8w .. 8w+7: lane l streams the zeros of its four
//      columns 4l..4l+3 to HBM (float4, 512 B per warp instruction — most of the
//      matrix is zero), always one tile ahead of the cull so that the write stream
//      never waits for a warp that is busy clipping;
//   2. cull level 1: lane l takes row l/4 of the warp's eight and column word l%4:
//      the 32 columns of that word the row can touch = (two table lookups OR-ed
//      for x) AND (two for y) — ~1 instruction a pair instead of a test a pair;
//   3. surviving pairs go to the warp's first queue (ballot ranks, at most one per
//      lane and round; fill counts are warp-uniform registers); whenever 32 are
//      queued the warp pops them for the circle test with every lane busy, its
//      survivors go to a second queue that is popped 32 at a time for the exact
//      clipper, which reads the geometry planes through L1/L2 and overwrites the
//      zeros with the non-zero values.
// A value always lands after its zero: both stores come from the same warp with
// a __syncwarp between.
// Disjoint pairs cost ~1 instruction and 4 B of HBM write; the kernel is bound by
// the HBM write stream (DESIGN.md "Kernels").
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

#ifndef ROTGEO_ABLATE
#define ROTGEO_ABLATE 0   // probe builds (tools/ablate_iou.py): 1 no zero stores, 2 no cull, 4 no exact pass, 8 no final drain,
                          // 16 no value stores, 32 cull data staged by plain loads of warp 0 instead of TMA bulk copies, 64 streaming zero stores
#endif

namespace rotgeo {

constexpr int kTileRows = 64;
constexpr int kTileCols = 128;
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kRowsPerWarp = kTileRows / kWarps;  // 8
constexpr int kStages = 4;        // mask ring
constexpr int kRespSlots = 8;     // launch-control responses (> kStages, see the reuse argument below)
constexpr int kQueueCap = 64;     // pairs per warp queue (never more than 31 + 32 queued)
constexpr int kPreMax = 3;        // tiles stolen and zero-filled before waiting for the pack kernel (< kRespSlots)

static_assert(kTileCols == kBlock, "a column tile is one inverted-index block");

struct __align__(16) IouSmem {
    uint32_t tab[kStages][kTableXYWords];   // x and y axes of the column block's index
    uint32_t rowr[kStages][kTileRows];
    int2 queue1[kWarps][kQueueCap];   // survived the bin ranges
    int2 queue2[kWarps][kQueueCap];   // survived the circle test as well
    uint4 resp[kRespSlots];
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t sched[kRespSlots];
};

struct IouArgs {
    const float4 *a0, *a1;
    const uint32_t *ar;      // row boxes: bin ranges
    const float4 *b0, *b1;
    const uint32_t *bt;      // column boxes: inverted-index tables, one per 128-block
    int64_t n1, n2;
    float *out;
    int64_t ld_out;
    unsigned tiles_x;
    // SINK 1: running maxima instead of the matrix.  key = float bits << 32 | ~index, so that
    // atomicMax keeps the largest value and, among equal values, the smallest index; 0 = no
    // positive value yet
    unsigned long long *row_key, *col_key;
    // SINK 2: which columns hold each row's maximum (low-quality matching of MaxIoUAssigner)
    const float *row_max;
    float min_pos_iou;
    int *col_lowq;           // atomicMax of row + 1
    int pre;                 // 0 .. kPreMax: tiles to steal and zero-fill before griddepcontrol.wait
};
enum { kSinkMatrix = 0, kSinkMax = 1, kSinkRowMaxHits = 2 };

// exact value of a pair that passed both culls; non-zero values overwrite the zero
template <int MODE, int SINK>
__device__ __forceinline__ void emit_pair(const IouArgs &g, int row, int col) {
    const float v = rect_overlap<MODE>(__ldg(g.a0 + row), __ldg(g.a1 + row), __ldg(g.b0 + col),
                                       __ldg(g.b1 + col));
    if (SINK == kSinkMatrix) {
#if ROTGEO_ABLATE & 16   // probe: no value stores (v == 12345 never happens)
        if (v == 12345.f) g.out[(int64_t)row * g.ld_out + col] = v;
#else
        if (v != 0.f) g.out[(int64_t)row * g.ld_out + col] = v;
#endif
    } else if (SINK == kSinkMax) {
        if (v > 0.f) {
            const unsigned long long hi = (unsigned long long)__float_as_uint(v) << 32;
            atomicMax(g.row_key + row, hi | (0xffffffffu - (unsigned)col));
            atomicMax(g.col_key + col, hi | (0xffffffffu - (unsigned)row));
        }
    } else {
        // the same instructions on the same inputs as in the kSinkMax pass: bitwise equal values
        const float m = g.row_max[row];
        if (v > 0.f && v == m && m >= g.min_pos_iou) atomicMax(g.col_lowq + col, row + 1);
    }
}

// SINK == kSinkMatrix: the (n1, n2) matrix; otherwise nothing is stored per pair, only the
// reductions of emit_pair (fused assigner path, rotgeo_assign.cu)
template <int MODE, bool VEC, int SINK>
__global__ void __launch_bounds__(kThreads)
iou_tile_kernel(const IouArgs g) {
    __shared__ IouSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], kWarps);
        }
#pragma unroll
        for (int s = 0; s < kRespSlots; ++s) mbar_init(&sm.sched[s], 1);
        fence_mbar_init();
    }
    __syncthreads();

    auto issue_masks = [&](unsigned k, unsigned t) {   // thread 0: cull data of tile t, my k-th
        const int s = k % kStages;
        const int64_t rb = (int64_t)(t / g.tiles_x) * kTileRows;
        const int64_t cblk = (int64_t)(t % g.tiles_x);
        mbar_expect_tx(&sm.full[s], (kTileRows + kTableXYWords) * sizeof(uint32_t));
        tma_load_1d(sm.rowr[s], g.ar + rb, kTileRows * sizeof(uint32_t), &sm.full[s]);
        tma_load_1d(sm.tab[s], g.bt + cblk * kTableWords, kTableXYWords * sizeof(uint32_t), &sm.full[s]);
    };
#if ROTGEO_ABLATE & 32
    // A/B probe: the same bytes through registers (ld.global + st.shared by the 32 lanes of
    // warp 0), then one plain arrival on the full barrier
    auto copy_masks_plain = [&](unsigned k, unsigned t) {   // warp 0, all lanes
        const int s = k % kStages;
        const int64_t rb = (int64_t)(t / g.tiles_x) * kTileRows;
        const int64_t cblk = (int64_t)(t % g.tiles_x);
        const uint4 *src_t = reinterpret_cast<const uint4 *>(g.bt + cblk * kTableWords);
        uint4 *dst_t = reinterpret_cast<uint4 *>(sm.tab[s]);
        for (int i = lane; i < kTableXYWords / 4; i += 32) dst_t[i] = __ldg(src_t + i);
        const uint4 *src_r = reinterpret_cast<const uint4 *>(g.ar + rb);
        uint4 *dst_r = reinterpret_cast<uint4 *>(sm.rowr[s]);
        for (int i = lane; i < kTileRows / 4; i += 32) dst_r[i] = __ldg(src_r + i);
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.full[s]);
    };
#endif
    // the answer for my k-th tile (k >= 1) lives in slot (k-1) % kRespSlots, use (k-1) / kRespSlots
    auto ask_next = [&](unsigned k) {                  // thread 0: who is my k-th tile?
        const int slot = (k - 1) % kRespSlots;
        mbar_expect_tx(&sm.sched[slot], 16);
        clc_try_cancel(&sm.resp[slot], &sm.sched[slot]);
    };
    unsigned t = blockIdx.x;       // my tile k = 0 is my own launch

    int2 *const q1 = sm.queue1[warp];
    int2 *const q2 = sm.queue2[warp];
    int n1 = 0, n2 = 0;                       // warp-uniform fill counts
    const uint32_t lt = (1u << lane) - 1u;
    auto clip32 = [&](int take) {             // pop `take` <= 32 entries of queue 2
        if (lane < take && !(ROTGEO_ABLATE & 4)) {
            const int2 e = q2[n2 - take + lane];
            emit_pair<MODE, SINK>(g, e.x, e.y);
        }
        n2 -= take;
        __syncwarp();
    };
    auto circle32 = [&](int take) {           // pop `take` <= 32 entries of queue 1
        bool pass = false;
        int2 e = make_int2(0, 0);
        if (lane < take) {
            e = q1[n1 - take + lane];
            const float4 A0 = __ldg(g.a0 + e.x), B0 = __ldg(g.b0 + e.y);
            pass = !circles_disjoint(A0.x, A0.y, A0.z, B0.x, B0.y, B0.z);
        }
        n1 -= take;
        const uint32_t b = __ballot_sync(0xffffffffu, pass);
        if (pass) q2[n2 + __popc(b & lt)] = e;
        n2 += __popc(b);
        __syncwarp();
        if (n2 >= 32) clip32(32);
    };

    // zeros of this warp's 8 x 128 strip of tile tt (independent of the cull data)
    auto store_zeros = [&](unsigned tt) {
        if ((ROTGEO_ABLATE & 1) || SINK != kSinkMatrix) return;
        const int64_t row_base = (int64_t)(tt / g.tiles_x) * kTileRows;
        const int64_t col_base = (int64_t)(tt % g.tiles_x) * kTileCols;
        const int64_t gcol = col_base + 4 * lane;
        const int cols_here = (int)min((int64_t)4, g.n2 - gcol);  // <= 0: lane is past the edge
        const int rows_here = (int)min((int64_t)kTileRows, g.n1 - row_base);
        // warp w owns the 8 CONSECUTIVE rows 8w .. 8w+7 of the tile (measured faster for the write
        // stream than rows w, w+8, ...: profiles/r01_hbm_write_probe.txt)
        float *dst = g.out + (row_base + kRowsPerWarp * warp) * g.ld_out + gcol;
        const int64_t step = g.ld_out;
        if (VEC) {
            if (rows_here == kTileRows && cols_here == 4) {   // interior tile: 8 plain stores
#pragma unroll
                for (int it = 0; it < kRowsPerWarp; ++it) {
#if ROTGEO_ABLATE & 64   // probe: streaming (evict-first) stores.  Same time on sparse scenes, but a value
                         // that lands after its zero line has left L2 costs a DRAM read-modify-write:
                         // dense C3 (extent 1024) 60 -> 82 us (profiles/r01_iou_ablation_exact.txt)
                    __stcs(reinterpret_cast<float4 *>(dst), make_float4(0.f, 0.f, 0.f, 0.f));
#else
                    *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
#endif
                    dst += step;
                }
            } else if (cols_here > 0) {
                for (int r = kRowsPerWarp * warp; r < min(rows_here, kRowsPerWarp * (warp + 1)); ++r) {
                    *reinterpret_cast<float4 *>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
                    dst += step;
                }
            }
        } else {
            for (int r = kRowsPerWarp * warp; r < min(rows_here, kRowsPerWarp * (warp + 1)); ++r) {
                for (int k = 0; k < cols_here; ++k) dst[k] = 0.f;
                dst += step;
            }
        }
    };
    store_zeros(t);   // the zeros always run one tile ahead of the cull
    // Launched as a programmatic dependent of the pack kernel, everything that reads the packed
    // sets has to wait for it (~7 us, latency-bound).  Neither the zeros nor the launch-control
    // requests depend on it: up to kPre further tiles are stolen and zero-filled first, so the
    // write stream is already running at full rate when the cull data become available.
    const int kPre = g.pre;
    if (kPre > 0) {
        if (tid == 0) {
            for (int j = 1; j <= kPre; ++j) ask_next(j);
        }
        for (int j = 0; j < kPre; ++j) {
            unsigned tj;
            mbar_wait(&sm.sched[j], 0);
            if (!clc_decode(&sm.resp[j], tj)) break;
            store_zeros(tj);
        }
    }
    pdl_wait();
#if ROTGEO_ABLATE & 32
    if (warp == 0) {
        copy_masks_plain(0, t);
        if (lane == 0 && kPre == 0) ask_next(1);
    }
#else
    if (tid == 0) {
        issue_masks(0, t);
        if (kPre == 0) ask_next(1);
    }
#endif

    for (unsigned k = 0;; ++k) {
        const int s = k % kStages;
        const int64_t row_base = (int64_t)(t / g.tiles_x) * kTileRows;
        const int64_t col_base = (int64_t)(t % g.tiles_x) * kTileCols;

        // ---- scheduler (thread 0): learn tile k+1, prefetch its masks, ask for k+2 ---
        // (the answer was requested a whole tile ago, so this wait is short)
#if ROTGEO_ABLATE & 32
        if (warp == 0) {
            const unsigned kn = k + 1;
            unsigned tn = 0;
            int ok = 0;
            if (lane == 0) {
                mbar_wait(&sm.sched[k % kRespSlots], (k / kRespSlots) & 1);
                ok = clc_decode(&sm.resp[k % kRespSlots], tn) ? 1 : 0;
                if (ok && kn >= kStages) mbar_wait(&sm.empty[kn % kStages], ((kn / kStages) - 1) & 1);
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            tn = __shfl_sync(0xffffffffu, tn, 0);
            if (ok) {
                copy_masks_plain(kn, tn);
                if (lane == 0) {
                    fence_proxy_async_smem();
                    if (k + 2 > (unsigned)kPre) ask_next(k + 2);
                }
            }
        }
#else
        if (tid == 0) {
            const unsigned kn = k + 1;
            mbar_wait(&sm.sched[k % kRespSlots], (k / kRespSlots) & 1);
            unsigned tn;
            if (clc_decode(&sm.resp[k % kRespSlots], tn)) {
                // ring stage of tile k+1 was last read for tile k+1-kStages: wait until every
                // warp left it.  That also proves every warp has read the responses up to
                // that tile, so the slot of tile k+2 (last used kRespSlots tiles ago) is free.
                if (kn >= kStages) mbar_wait(&sm.empty[kn % kStages], ((kn / kStages) - 1) & 1);
                issue_masks(kn, tn);
                fence_proxy_async_smem();
                if (k + 2 > (unsigned)kPre) ask_next(k + 2);   // the first kPre requests went out in the prologue
            }
        }
#endif

        // ---- cull level 1: which of my 32 columns can my row touch? ------------------
        // (padding rows have empty ranges, padding columns are in no table entry)
        mbar_wait(&sm.full[s], (k / kStages) & 1);
        uint32_t hits = 0;
        const int rl = lane >> 2, cw = lane & 3;        // row of the warp's eight, column word
        if (!(ROTGEO_ABLATE & 2)) {
            const uint32_t rr = sm.rowr[s][kRowsPerWarp * warp + rl];
            const BinRanges br = unpack_ranges(rr);
            const int sx = br.sx, nx = br.nx, sy = br.sy, ny = br.ny;
            if (nx != 0) {
                const int jx = min(31 - __clz(nx), kTableLevels - 1), jy = min(31 - __clz(ny), kTableLevels - 1);
                const uint32_t *tx = sm.tab[s] + jx * 128 + cw;
                const uint32_t *ty = sm.tab[s] + kTableAxisWords + jy * 128 + cw;
                const uint32_t X = tx[sx * 4] | tx[((sx + nx - (1 << jx)) & 31) * 4];
                const uint32_t Y = ty[sy * 4] | ty[((sy + ny - (1 << jy)) & 31) * 4];
                hits = X & Y;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);   // this warp is done with the stage
        // (measured and dropped: prefetch.global.L1 of the geometry lines a tile's candidates can touch, issued here —
        // the loads of the queue pops wait for first touches, a third of the stall samples of the dense C3 launch —
        // made the kernel SLOWER, 47.1 -> 51.2 us sparse and 59.4 -> 66.6 us dense: profiles/r02_iou_dense_ncu.txt)

        // ---- survivors -> queue 1, at most one per lane and round ----------------------
        {
            const int row = (int)row_base + kRowsPerWarp * warp + rl, col0 = (int)col_base + 32 * cw;
            uint32_t m = hits;
            for (;;) {
                const uint32_t act = __ballot_sync(0xffffffffu, m != 0u);
                if (act == 0u) break;
                if (m) {
                    q1[n1 + __popc(act & lt)] = make_int2(row, col0 + __ffs(m) - 1);
                    m &= m - 1;
                }
                n1 += __popc(act);
                __syncwarp();
                if (n1 >= 32) circle32(32);
            }
        }

        // ---- next tile (or none: nothing left to steal) ---------------------------------
        unsigned tn;
        mbar_wait(&sm.sched[k % kRespSlots], (k / kRespSlots) & 1);
        const bool more = clc_decode(&sm.resp[k % kRespSlots], tn);
        fence_proxy_async_smem();   // my read of the slot precedes its reuse by a later request
        // next tile's zeros are in flight while this warp clips (the first kPre were stored up front)
        if (more && k + 1 > (unsigned)kPre) store_zeros(tn);

        if (!more) {            // last tile of this CTA: flush both queues
            if (!(ROTGEO_ABLATE & 8)) {
                if (n1 > 0) circle32(n1);
                if (n2 > 0) clip32(n2);
            }
        }
        if (!more) break;
        t = tn;
    }
}

static int launch_tiles(void (*kern)(IouArgs), IouArgs &g, const void *packed1, int64_t n1,
                        const void *packed2, int64_t n2, bool after_pack, cudaStream_t stream) {
    if (n1 > 0x7fffffffLL || n2 > 0x7fffffffLL) return -1;
    const PackedView A = packed_view(packed1, padded_count(n1));
    const PackedView B = packed_view(packed2, padded_count(n2));
    const int64_t tiles_x = (n2 + kTileCols - 1) / kTileCols;
    const int64_t tiles = tiles_x * ((n1 + kTileRows - 1) / kTileRows);
    if (tiles > 0x7fffffffLL) return -1;
    g.a0 = A.p0; g.a1 = A.p1; g.ar = A.pr;
    g.b0 = B.p0; g.b1 = B.p1; g.bt = B.pt;
    g.n1 = n1; g.n2 = n2;
    g.tiles_x = (unsigned)tiles_x;
    // one CTA per tile; running CTAs steal the not-yet-launched ones (launch control)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)tiles);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = after_pack ? 1 : 0;   // overlap with the pack kernel launched just before
    g.pre = after_pack ? 2 : 0;          // measured: 1 .. 3 all hide the pack kernel (step 49.8 -> 48.3 us)
    if (cudaLaunchKernelEx(&cfg, kern, g) != cudaSuccess) return -3;
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int launch_iou_pairwise(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                        float *out, int64_t ld_out, bool after_pack, cudaStream_t stream) {
    if (n1 == 0 || n2 == 0) return 0;
    const bool vec = (ld_out % 4 == 0) && (n2 % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    void (*kern)(IouArgs) = mode == 0 ? (vec ? iou_tile_kernel<0, true, kSinkMatrix> : iou_tile_kernel<0, false, kSinkMatrix>)
                                      : (vec ? iou_tile_kernel<1, true, kSinkMatrix> : iou_tile_kernel<1, false, kSinkMatrix>);
    IouArgs g = {};
    g.out = out; g.ld_out = ld_out;
    return launch_tiles(kern, g, packed1, n1, packed2, n2, after_pack, stream);
}

int launch_iou_max(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                   unsigned long long *row_key, unsigned long long *col_key, bool after_pack,
                   cudaStream_t stream) {
    if (n1 == 0 || n2 == 0) return 0;
    IouArgs g = {};
    g.row_key = row_key; g.col_key = col_key;
    return launch_tiles(mode == 0 ? iou_tile_kernel<0, false, kSinkMax> : iou_tile_kernel<1, false, kSinkMax>,
                        g, packed1, n1, packed2, n2, after_pack, stream);
}

int launch_iou_row_max_hits(const void *packed1, int64_t n1, const void *packed2, int64_t n2, int mode,
                            const float *row_max, float min_pos_iou, int *col_lowq, cudaStream_t stream) {
    if (n1 == 0 || n2 == 0) return 0;
    IouArgs g = {};
    g.row_max = row_max; g.min_pos_iou = min_pos_iou; g.col_lowq = col_lowq;
    return launch_tiles(mode == 0 ? iou_tile_kernel<0, false, kSinkRowMaxHits> : iou_tile_kernel<1, false, kSinkRowMaxHits>,
                        g, packed1, n1, packed2, n2, false, stream);
}

// ---- aligned variant: out[i] = overlap(boxes1[i], boxes2[i]) -------------------
__device__ __forceinline__ void load_box(const float *b, float min_wh, float4 &p0, float4 &p1) {
    float w = b[2], h = b[3];
    if (min_wh > 0.f) {
        w = fmaxf(w, min_wh);
        h = fmaxf(h, min_wh);
    }
    double sd, cd;
    sincos((double)b[4], &sd, &cd);
    p0 = make_float4(b[0], b[1], 0.f, w * h);
    p1 = make_float4((float)cd, (float)sd, 0.5f * w, 0.5f * h);
}

template <int MODE>
__global__ void __launch_bounds__(128)
iou_aligned_kernel(const float *__restrict__ b1, const float *__restrict__ b2, int64_t n, int s1,
                   int s2, float min_wh, float *__restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 A0, A1, B0, B1;
    load_box(b1 + i * s1, min_wh, A0, A1);
    load_box(b2 + i * s2, min_wh, B0, B1);
    out[i] = rect_overlap<MODE>(A0, A1, B0, B1);
}

int launch_iou_aligned(const float *boxes1, const float *boxes2, int64_t n, int stride1,
                       int stride2, int mode, float min_wh, float *out, cudaStream_t stream) {
    if (n == 0) return 0;
    const int threads = 128;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (mode == 0)
        iou_aligned_kernel<0><<<blocks, threads, 0, stream>>>(boxes1, boxes2, n, stride1, stride2, min_wh, out);
    else
        iou_aligned_kernel<1><<<blocks, threads, 0, stream>>>(boxes1, boxes2, n, stride1, stride2, min_wh, out);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
