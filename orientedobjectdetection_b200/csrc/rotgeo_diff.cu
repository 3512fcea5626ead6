// Differentiable aligned rotated IoU: value and gradient with respect to all ten box
// parameters in one pass.  sm_100a.
//
// Replaces mmcv.ops.diff_iou_rotated_2d as used by the reference's RotatedIoULoss
// (mmrotate/models/losses/rotated_iou_loss.py:10,47; the fork's prototype configs train
// with it, configs/jy/prototype4.py:73).  mmcv builds the intersection polygon with two
// dozen differentiable torch ops (corner tests, edge crossings, a vertex sort kernel, a
// shoelace sum) and lets autograd differentiate through them.  Here the gradient is
// analytic: by the transport theorem the derivative of the intersection area with respect
// to a motion of box A is the integral of the normal velocity of A's boundary over the part
// of that boundary lying inside B — and the clipper of rotgeo_device.cuh already produces
// exactly those pieces, the clipped parameter interval [lo, hi] of each of the eight edges:
//
//   translation along an edge normal n:   +- (hi - lo)                 (the two sides with normal +-n)
//   growing w (h):                        1/2 (hi - lo) on both sides whose normal is the w (h) axis
//   rotation about the box centre:        +- (hi^2 - lo^2) / 2         (first moment of the piece)
//
// so one pass over the eight edges yields the area and d(area)/d(cx, cy, w, h, theta) of both
// boxes; IoU = I / (Sa + Sb - I) is then differentiated by the quotient rule.  One thread per
// pair, registers only.
#include "rotgeo_device.cuh"
#include "rotgeo_internal.h"

namespace rotgeo {

struct EdgeSums {
    // per box: lengths and first moments of the clipped pieces of its four sides,
    // [0] = the side on the positive axis, [1] = the negative one
    float Lw[2], Mw[2];   // sides whose normal is the box's w axis (parameter runs along the h axis)
    float Lh[2], Mh[2];   // sides whose normal is the h axis (parameter runs along the w axis)
};

__device__ __forceinline__ void piece(float lo, float hi, float &L, float &M) {
    const bool ok = hi > lo;
    L = ok ? hi - lo : 0.f;
    M = ok ? 0.5f * (hi - lo) * (hi + lo) : 0.f;
}

// the clipper of rect_intersection_area, keeping every clipped interval; frame and symmetries
// as described there (B's own frame; A rotated by pi if c < 0: sig_pi; frame mirrored about
// B's x axis if s < 0: sig_mir)
__device__ float intersection_with_pieces(float4 A0, float4 A1, float4 B0, float4 B1, EdgeSums &ea,
                                          EdgeSums &eb, float &sig_pi, float &sig_mir) {
    const float aA = A1.z, bA = A1.w, aB = B1.z, bB = B1.w;
    const float dx = A0.x - B0.x, dy = A0.y - B0.y;
    float c = fmaf(A1.x, B1.x, A1.y * B1.y);
    float s = fmaf(A1.y, B1.x, -A1.x * B1.y);
    const float X = fmaf(dx, B1.x, dy * B1.y);
    float Y = fmaf(dy, B1.x, -dx * B1.y);
    sig_pi = 1.f;
    sig_mir = 1.f;
    if (c < 0.f) { c = -c; s = -s; sig_pi = -1.f; }
    if (s < 0.f) { s = -s; Y = -Y; sig_mir = -1.f; }
    c = fmaxf(c, 1e-12f);
    s = fmaxf(s, 1e-12f);
    const float rc = 1.f / c, rs = 1.f / s;
    const float n2 = fmaf(c, c, s * s);
    const bool a_small = A0.w < B0.w;
    const float kE = a_small ? 0.f : fmaf(c, Y, -s * X);
    const float kF = a_small ? 0.f : fmaf(c, X, s * Y);
    const float lX = a_small ? X : 0.f;
    const float lY = a_small ? Y : 0.f;
    const float bAs = bA * s, bAc = bA * c, aAc = aA * c, aAs = aA * s;
    const float wE = bA * n2, wF = aA * n2;

    float acc = 0.f;
    float yE[2][2], xE[2][2], yF[2][2], xF[2][2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float g = k ? -1.f : 1.f;
        {   // A's side at local h-axis coordinate g*bA, parameter t along the w axis
            const float qx = fmaf(-g, bAs, X), qy = fmaf(g, bAc, Y);
            const float txm = (-aB - qx) * rc, txp = (aB - qx) * rc;
            const float tym = (-bB - qy) * rs, typ = (bB - qy) * rs;
            const float lo = fmax3(-aA, txm, tym), hi = fmin3(aA, txp, typ);
            piece(lo, hi, ea.Lh[k], ea.Mh[k]);
            acc = fmaf(ea.Lh[k], fmaf(g, kE, wE), acc);
            yE[k][1] = fmaf(txm, s, qy);
            yE[k][0] = fmaf(txp, s, qy);
            xE[k][1] = fmaf(tym, c, qx);
            xE[k][0] = fmaf(typ, c, qx);
        }
        {   // A's side at local w-axis coordinate g*aA, parameter t along the h axis
            const float px = fmaf(g, aAc, X), py = fmaf(g, aAs, Y);
            const float txp = (px - aB) * rs, txm = (px + aB) * rs;
            const float tym = (-bB - py) * rc, typ = (bB - py) * rc;
            const float lo = fmax3(-bA, txp, tym), hi = fmin3(bA, txm, typ);
            piece(lo, hi, ea.Lw[k], ea.Mw[k]);
            acc = fmaf(ea.Lw[k], fmaf(g, kF, wF), acc);
            yF[k][0] = fmaf(txp, c, py);
            yF[k][1] = fmaf(txm, c, py);
            xF[k][1] = fmaf(-tym, s, px);
            xF[k][0] = fmaf(-typ, s, px);
        }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const float g = r ? -1.f : 1.f;
        {   // B's side x = g*aB, parameter y
            const float lo = fmax3(-bB, yE[1][r], yF[1][r]);
            const float hi = fmin3(bB, yE[0][r], yF[0][r]);
            piece(lo, hi, eb.Lw[r], eb.Mw[r]);
            acc = fmaf(eb.Lw[r], fmaf(-g, lX, aB), acc);
        }
        {   // B's side y = g*bB, parameter x
            const float lo = fmax3(-aB, xE[0][r], xF[1][r]);
            const float hi = fmin3(aB, xE[1][r], xF[0][r]);
            piece(lo, hi, eb.Lh[r], eb.Mh[r]);
            acc = fmaf(eb.Lh[r], fmaf(-g, lY, bB), acc);
        }
    }
    return fmaxf(0.5f * acc, 0.f);
}

// d(area)/d(local u, local v, w, h, theta) of one box from its clipped pieces
// (u = along the w axis, v = along the h axis; outward normals +-u on the Lw sides, +-v on the Lh sides)
__device__ __forceinline__ void area_grad_local(const EdgeSums &e, float g[5]) {
    g[0] = e.Lw[0] - e.Lw[1];
    g[1] = e.Lh[0] - e.Lh[1];
    g[2] = 0.5f * (e.Lw[0] + e.Lw[1]);
    g[3] = 0.5f * (e.Lh[0] + e.Lh[1]);
    // rotation: boundary velocity (-v, u); on the side v = +-b it is +-u = +-t, on the side u = +-a it is -+t
    g[4] = (e.Mh[0] - e.Mh[1]) - (e.Mw[0] - e.Mw[1]);
}

__global__ void __launch_bounds__(128)
diff_iou_kernel(const float *__restrict__ b1, const float *__restrict__ b2, int64_t n, int s1, int s2,
                float *__restrict__ iou_out, float *__restrict__ grad1, float *__restrict__ grad2) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *pa = b1 + i * s1, *pb = b2 + i * s2;
    const float wa = pa[2], ha = pa[3], wb = pb[2], hb = pb[3];
    double sda, cda, sdb, cdb;
    sincos((double)pa[4], &sda, &cda);
    sincos((double)pb[4], &sdb, &cdb);
    const float ca = (float)cda, sa = (float)sda, cb = (float)cdb, sb = (float)sdb;
    const float Sa = wa * ha, Sb = wb * hb;
    const float4 A0 = make_float4(pa[0], pa[1], 0.f, Sa), A1 = make_float4(ca, sa, 0.5f * wa, 0.5f * ha);
    const float4 B0 = make_float4(pb[0], pb[1], 0.f, Sb), B1 = make_float4(cb, sb, 0.5f * wb, 0.5f * hb);
    EdgeSums ea, eb;
    float sig_pi, sig_mir;
    float I = intersection_with_pieces(A0, A1, B0, B1, ea, eb, sig_pi, sig_mir);
    const float cap = fminf(Sa, Sb);
    const bool degenerate = Sa < 1e-14f || Sb < 1e-14f;
    const bool capped = I > cap;          // rounding only: the area cannot exceed the smaller box
    I = fminf(I, cap);
    const float U = Sa + Sb - I;
    const float iou = degenerate ? 0.f : I / U;
    iou_out[i] = iou;
    if (!grad1 && !grad2) return;

    float ga[5], gb[5];
    area_grad_local(ea, ga);
    area_grad_local(eb, gb);
    // undo the two symmetries of the clipper's frame: A rotated by pi flips A's u, v axes;
    // the mirror flips every v axis and every sense of rotation
    ga[0] *= sig_pi;
    ga[1] *= sig_pi * sig_mir;
    ga[4] *= sig_mir;
    gb[1] *= sig_mir;
    gb[4] *= sig_mir;
    const float zero = (degenerate || capped || !(I > 0.f)) ? 0.f : 1.f;
    const float inv_u2 = zero / (U * U);
    const float sum = Sa + Sb;
    if (grad1) {
        float *o = grad1 + i * 5;
        // dIoU = (dI (Sa + Sb) - I dSa) / U^2
        const float du = ga[0] * sum * inv_u2, dv = ga[1] * sum * inv_u2;
        o[0] = du * ca - dv * sa;
        o[1] = du * sa + dv * ca;
        o[2] = (ga[2] * sum - I * ha) * inv_u2;
        o[3] = (ga[3] * sum - I * wa) * inv_u2;
        o[4] = ga[4] * sum * inv_u2;
    }
    if (grad2) {
        float *o = grad2 + i * 5;
        const float du = gb[0] * sum * inv_u2, dv = gb[1] * sum * inv_u2;
        o[0] = du * cb - dv * sb;
        o[1] = du * sb + dv * cb;
        o[2] = (gb[2] * sum - I * hb) * inv_u2;
        o[3] = (gb[3] * sum - I * wb) * inv_u2;
        o[4] = gb[4] * sum * inv_u2;
    }
}

int launch_diff_iou(const float *boxes1, const float *boxes2, int64_t n, int stride1, int stride2,
                    float *iou_out, float *grad1, float *grad2, cudaStream_t stream) {
    if (n == 0) return 0;
    diff_iou_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(boxes1, boxes2, n, stride1, stride2,
                                                                    iou_out, grad1, grad2);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace rotgeo
