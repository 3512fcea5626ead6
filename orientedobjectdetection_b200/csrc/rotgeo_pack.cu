// AoS (cx, cy, w, h, theta) rows -> packed SoA float4 planes (see rotgeo_device.cuh).
// One thread per box; the optional gather / centre offset / size clamp are the
// pieces of the reference's python front-end that would otherwise be separate
// torch kernels (clone + clamp_ at rotate_iou2d_calculator.py:84-87, the offset
// add at bbox_nms_rotated.py:73-77, index_select by score order in mmcv's
// nms_rotated wrapper).
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

__global__ void __launch_bounds__(256)
pack_kernel(const float *__restrict__ boxes, int64_t n, int stride,
            const int64_t *__restrict__ gather, const float *__restrict__ offsets, float min_wh,
            float4 *__restrict__ p0, float4 *__restrict__ p1, int64_t n_pad) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    if (i >= n) {
        // padding: a zero-area box very far away; every pair with it is rejected
        p0[i] = make_float4(1e30f, 1e30f, 0.f, 0.f);
        p1[i] = make_float4(1.f, 0.f, 0.f, 0.f);
        return;
    }
    const int64_t src = gather ? gather[i] : i;
    const float *b = boxes + src * stride;
    float cx = b[0], cy = b[1], w = b[2], h = b[3];
    const float th = b[4];
    if (offsets) {
        const float o = offsets[src];
        cx += o;
        cy += o;
    }
    if (min_wh > 0.f) {
        w = fmaxf(w, min_wh);
        h = fmaxf(h, min_wh);
    }
    // the reference takes cos/sin of theta in double and narrows (Spec A step 3)
    double sd, cd;
    sincos((double)th, &sd, &cd);
    const float a = 0.5f * w, hb = 0.5f * h;
    // circumscribed radius, inflated so the reject test stays conservative
    const float r = sqrtf(fmaf(a, a, hb * hb)) * 1.000001f + 1e-30f;
    p0[i] = make_float4(cx, cy, r, w * h);
    p1[i] = make_float4((float)cd, (float)sd, a, hb);
}

int launch_pack(const float *boxes, int64_t n, int stride, const int64_t *gather,
                const float *offsets, float min_wh, void *packed, int64_t n_pad,
                cudaStream_t stream) {
    if (n_pad == 0) return 0;
    float4 *p0 = reinterpret_cast<float4 *>(packed);
    float4 *p1 = p0 + n_pad;
    const int threads = 256;
    const int64_t blocks = (n_pad + threads - 1) / threads;
    pack_kernel<<<(unsigned)blocks, threads, 0, stream>>>(boxes, n, stride, gather, offsets, min_wh,
                                                          p0, p1, n_pad);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
