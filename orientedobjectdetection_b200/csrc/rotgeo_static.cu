// rotgeo_iou_pairwise_static: the pairwise IoU entry point for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  sm_100a.
//
// No counterpart in the reference.  Same two
// kernels as rotgeo_iou_pairwise (pack_kernel + iou_tile_kernel, PDL-chained) — the only
// difference is where the packed box sets live: not in a caller-owned workspace but in a
// file-scope __device__ array of ROTGEO_STATIC_SCRATCH_BYTES (1.5 MiB = two packed sets of
// 8192 boxes, the BASELINE C3 shape), which every GPU that loads the library holds until exit.
//
// The array's NAME is a host-side stand-in in host code, not a device address; the address is
// taken once per device with cudaGetSymbolAddress (that device current) and cached.
// Nothing has to start from zero: the pack kernel rewrites every slot and every table block the
// tile kernel reads (padding included), so the previous call's contents are never observed.
// All calls share the array: two calls that overlap on the GPU (different streams, or graph
// branches without a dependency) RACE.  Calls on one stream are ordered and safe.
#include <stdarg.h>
#include <stdio.h>

#include "../../include/rotgeo.h"
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[ROTGEO_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void *g_scratch_ptr[kMaxDevices];   // per device; benign race: every writer stores the same address

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// device address of the scratch array on the CURRENT device (queried once per device)
int scratch_address(void **out, int *dev_out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    *dev_out = dev;
    void *p = (dev >= 0 && dev < kMaxDevices) ? g_scratch_ptr[dev] : nullptr;
    if (!p) {
        if (cudaGetSymbolAddress(&p, g_static_scratch) != cudaSuccess || !p) return -1;
        if (dev >= 0 && dev < kMaxDevices) g_scratch_ptr[dev] = p;
    }
    *out = p;
    return 0;
}

}  // namespace

namespace rotgeo {
int abi_fail(int code, const char *fmt, ...);   // rotgeo_abi.cu: sets the thread-local error text
}

extern "C" int rotgeo_iou_pairwise_static(const float *boxes1, int64_t n1, int stride1, const float *boxes2,
                                          int64_t n2, int stride2, int mode, float min_wh, float *out,
                                          int64_t ld_out, void *stream) {
    using rotgeo::abi_fail;
    if (n1 < 0 || n2 < 0 || stride1 < 5 || stride2 < 5 || (mode != 0 && mode != 1))
        return abi_fail(ROTGEO_ERR_ARG, "iou_static: n1=%lld n2=%lld stride1=%d stride2=%d mode=%d", (long long)n1,
                        (long long)n2, stride1, stride2, mode);
    if (n1 == 0 || n2 == 0) return ROTGEO_OK;
    if (!boxes1 || !boxes2 || !out || ld_out < n2)
        return abi_fail(ROTGEO_ERR_ARG, "iou_static: null pointer or ld_out=%lld < n2", (long long)ld_out);
    const size_t bytes1 = align_up((size_t)rotgeo::padded_count(n1) * rotgeo::kPackedBytesPerBox, 256);
    const size_t bytes2 = align_up((size_t)rotgeo::padded_count(n2) * rotgeo::kPackedBytesPerBox, 256);
    if (bytes1 + bytes2 > (size_t)ROTGEO_STATIC_SCRATCH_BYTES)
        return abi_fail(ROTGEO_ERR_WORKSPACE,
                        "iou_static: %lld + %lld boxes need %zu bytes of packed sets, the static scratch holds %d "
                        "(padded box counts must add up to at most %d); use rotgeo_iou_pairwise with a workspace",
                        (long long)n1, (long long)n2, bytes1 + bytes2, ROTGEO_STATIC_SCRATCH_BYTES,
                        ROTGEO_STATIC_SCRATCH_BYTES / rotgeo::kPackedBytesPerBox);
    void *scratch = nullptr;
    int dev = 0;
    if (scratch_address(&scratch, &dev))
        return abi_fail(ROTGEO_ERR_CUDA, "iou_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                        cudaGetErrorString(cudaGetLastError()));
    char *pk1 = reinterpret_cast<char *>(scratch);
    char *pk2 = pk1 + bytes1;
    cudaStream_t st = (cudaStream_t)stream;
    // both sets in ONE pack launch, so they share the automatically chosen bin width
    const rotgeo::PackSet s1 = rotgeo::make_pack_set(boxes1, n1, stride1, nullptr, nullptr, pk1, rotgeo::padded_count(n1));
    const rotgeo::PackSet s2 = rotgeo::make_pack_set(boxes2, n2, stride2, nullptr, nullptr, pk2, rotgeo::padded_count(n2));
    if (rotgeo::launch_pack(s1, s2, min_wh, 0.f, 0.f, st))
        return abi_fail(ROTGEO_ERR_CUDA, "iou_static: pack launch: %s", cudaGetErrorString(cudaGetLastError()));
    const int rc = rotgeo::launch_iou_pairwise(pk1, n1, pk2, n2, mode, out, ld_out, true, st);
    if (rc == -1) return abi_fail(ROTGEO_ERR_ARG, "iou_static: too many tiles");
    if (rc) return abi_fail(ROTGEO_ERR_CUDA, "iou_static: launch: %s", cudaGetErrorString(cudaGetLastError()));
    return ROTGEO_OK;
}
