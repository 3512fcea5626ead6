// Multi-GPU exchange of the sharded hot path over peer memory (NVLink / NVSwitch), one process
// per GPU.  sm_100a.
//
// The reference has no multi-GPU version of this path (python loops inside one rank:
// mmrotate/models/dense_heads/rotated_rpn_head.py:414-426,515 over images,
// mmrotate/core/patch/merge_results.py:90-119 over classes).  Here every rank computes its
// shard (IoU rows, images, classes) and the results reach ALL ranks by direct stores into the
// peers' memory — no staging copy, no library collective in the data path:
//
//   window        one symmetric buffer per rank, allocated and exchanged by the host layer
//                 (torch symmetric memory: plumbing); the library only sees the table of peer
//                 base addresses.  The first kPeerPadBytes of a window are its signal pad.
//   barrier       every rank stores a monotonically increasing epoch into slot [rank] of every
//                 peer's pad (st.release.sys) and spins on its own pad (ld.acquire.sys): a few us
//                 over NVSwitch, no host involvement, stream-ordered with the kernels around it.
//   IoU matrix    all-gather of a row-sharded matrix WITHOUT moving zeros: every rank zero-fills
//                 the rows it does not own in its own window at HBM speed, computes its rows in
//                 place, and after a barrier pushes only the NON-ZERO 16-byte groups of its block
//                 to the peers (0.06 % of the pairs of BASELINE C3 overlap at all; a dense matrix
//                 degrades to the plain P2P all-gather).  The 235 MB per GPU that a library
//                 all-gather moves over NVLink (0.45 of 0.51 ms at 8 GPUs in round 1) never travel.
//   maxima        row / column (max, argmax) of the sharded matrix: packed 64-bit keys pushed to
//                 every peer, reduced locally after the barrier (an all-reduce of 64 KB).
//   keep lists    per-image / per-class results of the segmented NMS written into fixed slots of
//                 every peer's window by the rank that computed them.
#include <stdio.h>

#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

namespace {

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ char *peer_ptr(const PeerTable &t, int r, size_t off) {
    return reinterpret_cast<char *>(t.base[r]) + off;
}

// One CTA, thread r talks to peer r.  A peer that never arrives (crashed process) must not hang
// the GPU: after kPeerTimeoutNs the wait gives up and raises the error word of the local pad.
__global__ void __launch_bounds__(kPeerMax)
peer_barrier_kernel(const PeerTable t, unsigned long long epoch) {
    const int r = threadIdx.x;
    if (r >= t.world) return;
    __threadfence_system();      // my earlier writes (remote ones included) before my signal
    st_release_sys(reinterpret_cast<unsigned long long *>(t.base[r]) + t.rank, epoch);
    const unsigned long long *mine = reinterpret_cast<const unsigned long long *>(t.base[t.rank]) + r;
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(mine) < epoch) {
        if (global_timer_ns() - t0 > kPeerTimeoutNs) {
            reinterpret_cast<unsigned long long *>(t.base[t.rank])[kPeerErrorSlot] = epoch;
            break;
        }
    }
    __threadfence_system();
}

// non-zero 16-byte groups of elements [first, first + count) of the local window -> all peers
__global__ void __launch_bounds__(256)
push_nonzero_vec_kernel(const PeerTable t, size_t off, int64_t first4, int64_t count4) {
    const float4 *src = reinterpret_cast<const float4 *>(peer_ptr(t, t.rank, off)) + first4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count4; i += 4 * stride) {
        // four independent 16-byte loads in flight per thread: the scan is an HBM/L2 read stream
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int64_t j = i + k * stride;
            v[k] = j < count4 ? __ldcs(src + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (v[k].x != 0.f || v[k].y != 0.f || v[k].z != 0.f || v[k].w != 0.f) {
                const int64_t j = i + k * stride;
                for (int r = 0; r < t.world; ++r)
                    if (r != t.rank) reinterpret_cast<float4 *>(peer_ptr(t, r, off))[first4 + j] = v[k];
            }
        }
    }
}
__global__ void __launch_bounds__(256)
push_nonzero_kernel(const PeerTable t, size_t off, int64_t first, int64_t count) {
    const float *src = reinterpret_cast<const float *>(peer_ptr(t, t.rank, off)) + first;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = __ldcs(src + i);
        if (v != 0.f)
            for (int r = 0; r < t.world; ++r)
                if (r != t.rank) reinterpret_cast<float *>(peer_ptr(t, r, off))[first + i] = v;
    }
}

// row keys of my rows -> slot [row_lo + i] of every rank's row array; column keys (partial maxima
// over my rows, row indices made global) -> stage [rank][j] of every rank
__global__ void __launch_bounds__(256)
push_keys_kernel(const PeerTable t, size_t row_off, size_t col_off, const unsigned long long *__restrict__ row_key,
                 int64_t row_lo, int64_t n1_local, const unsigned long long *__restrict__ col_key, int64_t n2) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n1_local) {
        unsigned long long k = row_key ? row_key[i] : 0ull;
        for (int r = 0; r < t.world; ++r)
            reinterpret_cast<unsigned long long *>(peer_ptr(t, r, row_off))[row_lo + i] = k;
    }
    if (i < n2) {
        unsigned long long k = col_key ? col_key[i] : 0ull;
        if (k) {      // low word = ~row: local row -> global row
            const unsigned int row = 0xffffffffu - (unsigned int)k;
            k = (k & 0xffffffff00000000ull) | (0xffffffffu - (row + (unsigned int)row_lo));
        }
        for (int r = 0; r < t.world; ++r)
            reinterpret_cast<unsigned long long *>(peer_ptr(t, r, col_off))[(int64_t)t.rank * n2 + i] = k;
    }
}

__global__ void __launch_bounds__(256)
reduce_keys_kernel(const PeerTable t, size_t row_off, size_t col_off, int64_t n1, int64_t n2,
                   float *__restrict__ row_max, int64_t *__restrict__ row_arg, float *__restrict__ col_max,
                   int64_t *__restrict__ col_arg) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n1) {
        const unsigned long long k = __ldcg(reinterpret_cast<const unsigned long long *>(peer_ptr(t, t.rank, row_off)) + i);
        if (row_max) row_max[i] = __uint_as_float((unsigned)(k >> 32));
        if (row_arg) row_arg[i] = k ? (int64_t)(0xffffffffu - (unsigned)k) : 0;
    }
    if (i < n2) {
        unsigned long long k = 0ull;
        const unsigned long long *st = reinterpret_cast<const unsigned long long *>(peer_ptr(t, t.rank, col_off));
        for (int r = 0; r < t.world; ++r) {
            const unsigned long long v = __ldcg(st + (int64_t)r * n2 + i);
            k = v > k ? v : k;       // larger value first, then the smaller (global) row
        }
        if (col_max) col_max[i] = __uint_as_float((unsigned)(k >> 32));
        if (col_arg) col_arg[i] = k ? (int64_t)(0xffffffffu - (unsigned)k) : 0;
    }
}

// kept rows of my groups -> fixed slots of every rank's window at `off`:
//   int64 counts[ngroups_total] | int64 errors[world] | int64 keep[ngroups_total][cap] |
//   float rows[ngroups_total][cap][row_floats]   (optional payload)
struct KeepPush {
    const int64_t *keep;          // group after group (rows inside the group), as rotgeo_nms_run leaves them
    const int64_t *tail;          // [kept in all, error bits, kept per local group ...] (device)
    int nlocal;
    int64_t cap;                  // slots per group (keep lists are cut to this)
    int64_t ngroups_total;
    int gid[kSortGroupsInline];   // global id of every local group
    const float *rows_src;        // nullable: payload rows, row_floats floats each ...
    int row_floats, row_stride;
    int64_t group_base[kSortGroupsInline];   // ... row of rows_src = group_base[g] + keep value
};
__global__ void __launch_bounds__(256)
push_keeps_kernel(const PeerTable t, size_t off, const KeepPush a) {
    __shared__ int64_t s_start[kSortGroupsInline + 1];
    if (threadIdx.x == 0) {
        int64_t s = 0;
        for (int g = 0; g < a.nlocal; ++g) {
            s_start[g] = s;
            s += a.tail[2 + g];
        }
        s_start[a.nlocal] = s;
    }
    __syncthreads();
    const int g = blockIdx.y;
    const int64_t cnt = min(s_start[g + 1] - s_start[g], a.cap);
    const int64_t gid = a.gid[g];
    const size_t err_off = off + (size_t)a.ngroups_total * 8;
    const size_t keep_off = err_off + (size_t)t.world * 8;
    const size_t rows_off = keep_off + (size_t)a.ngroups_total * a.cap * 8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cnt; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = a.keep[s_start[g] + i];
        float row[8];
        if (a.rows_src)
            for (int k = 0; k < a.row_floats; ++k) row[k] = a.rows_src[(a.group_base[g] + v) * a.row_stride + k];
        for (int r = 0; r < t.world; ++r) {
            reinterpret_cast<int64_t *>(peer_ptr(t, r, keep_off))[gid * a.cap + i] = v;
            if (a.rows_src) {
                float *dst = reinterpret_cast<float *>(peer_ptr(t, r, rows_off)) + (gid * a.cap + i) * a.row_floats;
                for (int k = 0; k < a.row_floats; ++k) dst[k] = row[k];
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int r = 0; r < t.world; ++r) {
            reinterpret_cast<int64_t *>(peer_ptr(t, r, off))[gid] = cnt;
            if (g == 0) reinterpret_cast<int64_t *>(peer_ptr(t, r, err_off))[t.rank] = a.tail[1];
        }
    }
}
__global__ void push_error_only_kernel(const PeerTable t, size_t err_off, int64_t value) {
    if (threadIdx.x < t.world) reinterpret_cast<int64_t *>(peer_ptr(t, threadIdx.x, err_off))[t.rank] = value;
}

}  // namespace

int launch_peer_barrier(const PeerTable &t, unsigned long long epoch, cudaStream_t stream) {
    peer_barrier_kernel<<<1, kPeerMax, 0, stream>>>(t, epoch);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

static int launch_push_nonzero(const PeerTable &t, size_t off, int64_t first, int64_t count, cudaStream_t stream) {
    if (count <= 0 || t.world <= 1) return 0;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        return -3;
    const bool vec = (off % 16 == 0) && (first % 4 == 0) && (count % 4 == 0);
    const int64_t items = vec ? count / 4 : count;
    int64_t blocks = (items + 1023) / 1024;
    if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
    if (blocks < 1) blocks = 1;
    if (vec)
        push_nonzero_vec_kernel<<<(unsigned)blocks, 256, 0, stream>>>(t, off, first / 4, count / 4);
    else
        push_nonzero_kernel<<<(unsigned)blocks, 256, 0, stream>>>(t, off, first, count);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int launch_dist_iou_allgather(const PeerTable &t, unsigned long long epoch, size_t matrix_off,
                              const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                              int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                              void *workspace, cudaStream_t stream) {
    float *out = reinterpret_cast<float *>(reinterpret_cast<char *>(t.base[t.rank]) + matrix_off);
    // the rows of the other ranks: zeros now, their non-zero values arrive after the first barrier
    if (row_lo > 0 && cudaMemsetAsync(out, 0, (size_t)row_lo * n2 * 4, stream) != cudaSuccess) return -3;
    const int64_t row_hi = row_lo + n1_local;
    if (row_hi < n1_total &&
        cudaMemsetAsync(out + row_hi * n2, 0, (size_t)(n1_total - row_hi) * n2 * 4, stream) != cudaSuccess)
        return -3;
    if (n1_local > 0) {
        char *ws = reinterpret_cast<char *>(workspace);
        void *pk1 = ws;
        void *pk2 = ws + ((size_t)padded_count(n1_local) * kPackedBytesPerBox + 255) / 256 * 256;
        const PackSet s1 = make_pack_set(boxes1_local, n1_local, stride1, nullptr, nullptr, pk1, padded_count(n1_local));
        const PackSet s2 = make_pack_set(boxes2, n2, stride2, nullptr, nullptr, pk2, padded_count(n2));
        if (launch_pack(s1, s2, min_wh, 0.f, 0.f, stream)) return -3;
        const int rc = launch_iou_pairwise(pk1, n1_local, pk2, n2, mode, out + row_lo * n2, n2, true, stream);
        if (rc) return rc;
    }
    if (t.world <= 1) return 0;
    if (launch_peer_barrier(t, epoch, stream)) return -3;            // everybody's zeros are in place
    if (launch_push_nonzero(t, matrix_off, row_lo * n2, n1_local * n2, stream)) return -3;
    return launch_peer_barrier(t, epoch + 1, stream);                // everybody's values have landed
}

int launch_dist_iou_rowcol_max(const PeerTable &t, unsigned long long epoch, size_t stage_off,
                               const float *boxes1_local, int64_t row_lo, int64_t n1_local, int64_t n1_total,
                               int stride1, const float *boxes2, int64_t n2, int stride2, int mode, float min_wh,
                               float *row_max, int64_t *row_argmax, float *col_max, int64_t *col_argmax,
                               void *workspace, cudaStream_t stream) {
    unsigned long long *row_key = nullptr, *col_key = nullptr;
    const size_t row_off = stage_off, col_off = stage_off + ((size_t)n1_total * 8 + 255) / 256 * 256;
    if (n1_local > 0) {
        const int rc = launch_iou_max_keys(boxes1_local, n1_local, stride1, boxes2, n2, stride2, mode, min_wh,
                                           workspace, &row_key, &col_key, stream);
        if (rc) return rc;
    }
    const int64_t span = n1_local > n2 ? n1_local : n2;
    push_keys_kernel<<<(unsigned)((span + 255) / 256), 256, 0, stream>>>(t, row_off, col_off, row_key, row_lo,
                                                                        n1_local, col_key, n2);
    if (cudaGetLastError() != cudaSuccess) return -3;
    if (launch_peer_barrier(t, epoch, stream)) return -3;
    const int64_t span2 = n1_total > n2 ? n1_total : n2;
    reduce_keys_kernel<<<(unsigned)((span2 + 255) / 256), 256, 0, stream>>>(t, row_off, col_off, n1_total, n2,
                                                                           row_max, row_argmax, col_max, col_argmax);
    if (cudaGetLastError() != cudaSuccess) return -3;
    // nobody may overwrite the stage of a peer that is still reducing: the next exchange starts behind this
    return launch_peer_barrier(t, epoch + 1, stream);
}

int launch_dist_push_keeps(const PeerTable &t, unsigned long long epoch, size_t out_off, const int64_t *keep,
                           const int64_t *tail, int nlocal, const int64_t *gid_host, int64_t cap,
                           int64_t ngroups_total, const float *rows_src, int row_floats, int row_stride,
                           const int64_t *group_base_host, cudaStream_t stream) {
    if (nlocal > kSortGroupsInline || row_floats > 8) return -1;
    if (nlocal > 0) {
        KeepPush a = {};
        a.keep = keep;
        a.tail = tail;
        a.nlocal = nlocal;
        a.cap = cap;
        a.ngroups_total = ngroups_total;
        a.rows_src = rows_src;
        a.row_floats = row_floats;
        a.row_stride = row_stride;
        for (int g = 0; g < nlocal; ++g) {
            a.gid[g] = (int)gid_host[g];
            a.group_base[g] = group_base_host ? group_base_host[g] : 0;
        }
        int64_t bx = (cap + 255) / 256;
        if (bx > 32) bx = 32;
        if (bx < 1) bx = 1;
        push_keeps_kernel<<<dim3((unsigned)bx, (unsigned)nlocal), 256, 0, stream>>>(t, out_off, a);
    } else {
        push_error_only_kernel<<<1, kPeerMax, 0, stream>>>(t, out_off + (size_t)ngroups_total * 8, 0);
    }
    if (cudaGetLastError() != cudaSuccess) return -3;
    return launch_peer_barrier(t, epoch, stream);
}

}  // namespace rotgeo
