// ys_math.cuh -- scalar math of the detection-head hot path, shared by every kernel.
// Everything is __host__ __device__ so the same code is unit-tested on the CPU
// (csrc/ys_hostmath.cu) without a GPU.
//
// Bit-exact section: the IoU forms below reproduce the reference's rounding.  The
// reference evaluates each +,-,*,/ as a separate ATen kernel, so every intermediate is
// rounded to fp32 and nothing is fused; here that is enforced with the *_rn intrinsics
// (device) and -ffp-contract=off (host).  Citations: /root/reference/yolo_scratch_cuda/.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define YS_HD __host__ __device__ __forceinline__

#ifdef __CUDA_ARCH__
#define YS_MUL(a, b) __fmul_rn((a), (b))
#define YS_ADD(a, b) __fadd_rn((a), (b))
#define YS_SUB(a, b) __fsub_rn((a), (b))
#define YS_DIV(a, b) __fdiv_rn((a), (b))
#else
#define YS_MUL(a, b) ((a) * (b))
#define YS_ADD(a, b) ((a) + (b))
#define YS_SUB(a, b) ((a) - (b))
#define YS_DIV(a, b) ((a) / (b))
#endif

namespace ys {

constexpr float kEpsIou = 1e-7f;     // losses/yolo_v8_loss.py:12
constexpr float kEpsInBox = 1e-9f;   // losses/yolo_v8_loss.py:210
constexpr float kEpsNms = 1e-6f;     // train_yolo11_cuda.py:435, utils/metrics.py:79

YS_HD float fminf2(float a, float b) { return a < b ? a : b; }   // torch.minimum on non-NaN input
YS_HD float fmaxf2(float a, float b) { return a > b ? a : b; }

// w*h of an xyxy box with the reference's "+eps on heights only" (losses/yolo_v8_loss.py:33-34)
YS_HD float area_eps_h(float x1, float y1, float x2, float y2) {
    return YS_MUL(YS_SUB(x2, x1), YS_ADD(YS_SUB(y2, y1), kEpsIou));
}

// Intersection area (losses/yolo_v8_loss.py:37-38 == train_yolo11_cuda.py:418-423)
YS_HD float inter_area(float ax1, float ay1, float ax2, float ay2, float bx1, float by1, float bx2, float by2) {
    float iw = YS_SUB(fminf2(ax2, bx2), fmaxf2(ax1, bx1));
    float ih = YS_SUB(fminf2(ay2, by2), fmaxf2(ay1, by1));
    iw = iw < 0.f ? 0.f : iw;
    ih = ih < 0.f ? 0.f : ih;
    return YS_MUL(iw, ih);
}

// Assigner overlap: bbox_iou(xywh=False) then clamp_(0)  (losses/yolo_v8_loss.py:31-44, :199).
// area_p / area_q are area_eps_h of the two boxes (each product is rounded on its own in the reference too).
YS_HD float iou_assign(float4 p, float area_p, float4 q, float area_q) {
    float inter = inter_area(p.x, p.y, p.z, p.w, q.x, q.y, q.z, q.w);
    float uni = YS_ADD(YS_SUB(YS_ADD(area_p, area_q), inter), kEpsIou);
    float iou = YS_DIV(inter, uni);
    return iou < 0.f ? 0.f : iou;
}

// bbox_iou(xywh=False, CIoU=True) in the reference's operation order (losses/yolo_v8_loss.py:31-56), unclamped: the
// pair metric of the non-default ciou_assign mode of the assigner.  atanf is the only non-exact step.
YS_HD float ciou_pair(float4 p, float4 q) {
    const float eps = kEpsIou;
    const float w1 = YS_SUB(p.z, p.x), h1 = YS_ADD(YS_SUB(p.w, p.y), eps);
    const float w2 = YS_SUB(q.z, q.x), h2 = YS_ADD(YS_SUB(q.w, q.y), eps);
    const float inter = inter_area(p.x, p.y, p.z, p.w, q.x, q.y, q.z, q.w);
    const float uni = YS_ADD(YS_SUB(YS_ADD(YS_MUL(w1, h1), YS_MUL(w2, h2)), inter), eps);
    const float iou = YS_DIV(inter, uni);
    const float cw = YS_SUB(fmaxf2(p.z, q.z), fminf2(p.x, q.x));
    const float ch = YS_SUB(fmaxf2(p.w, q.w), fminf2(p.y, q.y));
    const float c2 = YS_ADD(YS_ADD(YS_MUL(cw, cw), YS_MUL(ch, ch)), eps);
    const float sx = YS_SUB(YS_SUB(YS_ADD(q.x, q.z), p.x), p.z), sy = YS_SUB(YS_SUB(YS_ADD(q.y, q.w), p.y), p.w);
    const float rho2 = YS_DIV(YS_ADD(YS_MUL(sx, sx), YS_MUL(sy, sy)), 4.f);
    const float dat = YS_SUB(atanf(YS_DIV(w2, h2)), atanf(YS_DIV(w1, h1)));
    const float v = YS_MUL(0.40528473456935108578f, YS_MUL(dat, dat));
    const float alpha = YS_DIV(v, YS_ADD(YS_SUB(v, iou), YS_ADD(1.f, eps)));
    return YS_SUB(iou, YS_ADD(YS_DIV(rho2, c2), YS_MUL(v, alpha)));
}

// Anchor centre strictly inside the GT (losses/yolo_v8_loss.py:220-224)
YS_HD bool centre_in_box(float ax, float ay, float4 q) {
    float d = fminf2(fminf2(YS_SUB(ax, q.x), YS_SUB(ay, q.y)), fminf2(YS_SUB(q.z, ax), YS_SUB(q.w, ay)));
    return d > kEpsInBox;
}

// NMS / metrics IoU: inter / (a1 + a2 - inter + 1e-6)  (train_yolo11_cuda.py:418-435, utils/metrics.py:65-79)
YS_HD float area_plain(float4 b) { return YS_MUL(YS_SUB(b.z, b.x), YS_SUB(b.w, b.y)); }
YS_HD float iou_nms(float4 a, float area_a, float4 b, float area_b) {
    float inter = inter_area(a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w);
    float uni = YS_SUB(YS_ADD(area_a, area_b), inter);
    return YS_DIV(inter, YS_ADD(uni, kEpsNms));
}

// align = score^alpha * ov^beta ; norm = align*ov/(align+eps)  (losses/yolo_v8_loss.py:206, :173-178)
YS_HD float pow_special(float x, float e) {
    if (e == 0.5f) return sqrtf(x);
    if (e == 1.0f) return x;
    if (e == 2.0f) return x * x;
    if (e == 4.0f) { float x2 = x * x; return x2 * x2; }
    if (e == 6.0f) { float x2 = x * x; return x2 * x2 * x2; }
    return powf(x, e);
}
YS_HD float align_norm(float score, float ov, float alpha, float beta, float eps) {
    float am = pow_special(score, alpha) * pow_special(ov, beta);
    return am * ov / (am + eps);
}

YS_HD float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }
// Order-preserving key for "max value, lowest index wins": (value bits << 32) | ~index.  value >= 0.
YS_HD unsigned long long argmax_key(float v, unsigned idx) {
#ifdef __CUDA_ARCH__
    unsigned bits = __float_as_uint(v);
#else
    union { float f; unsigned u; } c; c.f = v; unsigned bits = c.u;
#endif
    return ((unsigned long long)bits << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
}
YS_HD unsigned argmax_key_index(unsigned long long k) { return 0xFFFFFFFFu - (unsigned)(k & 0xFFFFFFFFull); }

// ---------------------------------------------------------------------------
// Tolerance section (1e-5): CIoU value + gradient wrt the predicted box, alpha held constant
// (losses/yolo_v8_loss.py:46-56).  Tie rules follow torch autograd: min/max split 0.5/0.5,
// clamp(min=0) passes the gradient where the argument is >= 0.
// p = predicted xyxy, t = target xyxy.  Returns ciou; g[0..3] = d ciou / d (x1,y1,x2,y2) of p.
// ---------------------------------------------------------------------------
YS_HD float sel_gt(float a, float b) { return a > b ? 1.f : (a == b ? 0.5f : 0.f); }   // d max(a,b)/da
YS_HD float sel_lt(float a, float b) { return a < b ? 1.f : (a == b ? 0.5f : 0.f); }   // d min(a,b)/da

YS_HD float ciou_with_grad(float4 p, float4 t, float g[4]) {
    const float eps = kEpsIou;
    const float k4pi2 = 0.40528473456935108578f;   // 4/pi^2
    float w1 = p.z - p.x, h1 = p.w - p.y + eps;
    float w2 = t.z - t.x, h2 = t.w - t.y + eps;
    float iwr = fminf2(p.z, t.z) - fmaxf2(p.x, t.x);
    float ihr = fminf2(p.w, t.w) - fmaxf2(p.y, t.y);
    float iw = iwr < 0.f ? 0.f : iwr, ih = ihr < 0.f ? 0.f : ihr;
    float mw = iwr >= 0.f ? 1.f : 0.f, mh = ihr >= 0.f ? 1.f : 0.f;
    float inter = iw * ih;
    float uni = w1 * h1 + w2 * h2 - inter + eps;
    float iou = inter / uni;
    float cw = fmaxf2(p.z, t.z) - fminf2(p.x, t.x);
    float ch = fmaxf2(p.w, t.w) - fminf2(p.y, t.y);
    float c2 = cw * cw + ch * ch + eps;
    float sx = t.x + t.z - p.x - p.z, sy = t.y + t.w - p.y - p.w;
    float rho2 = (sx * sx + sy * sy) * 0.25f;
    float dat = atanf(w2 / h2) - atanf(w1 / h1);
    float v = k4pi2 * dat * dat;
    float alpha = v / (v - iou + (1.f + eps));
    float ciou = iou - (rho2 / c2 + v * alpha);

    // d inter
    float di_x1 = -ih * mw * sel_gt(p.x, t.x);
    float di_x2 = ih * mw * sel_lt(p.z, t.z);
    float di_y1 = -iw * mh * sel_gt(p.y, t.y);
    float di_y2 = iw * mh * sel_lt(p.w, t.w);
    // d union = d(w1*h1) - d inter
    float du_x1 = -h1 - di_x1, du_x2 = h1 - di_x2, du_y1 = -w1 - di_y1, du_y2 = w1 - di_y2;
    float inv_u = 1.f / uni;
    float dq_x1 = (di_x1 - iou * du_x1) * inv_u;
    float dq_x2 = (di_x2 - iou * du_x2) * inv_u;
    float dq_y1 = (di_y1 - iou * du_y1) * inv_u;
    float dq_y2 = (di_y2 - iou * du_y2) * inv_u;
    // d (rho2/c2)
    float inv_c2 = 1.f / c2;
    float r = rho2 * inv_c2;
    float dc_x1 = -2.f * cw * sel_lt(p.x, t.x), dc_x2 = 2.f * cw * sel_gt(p.z, t.z);
    float dc_y1 = -2.f * ch * sel_lt(p.y, t.y), dc_y2 = 2.f * ch * sel_gt(p.w, t.w);
    float dr_x1 = (-0.5f * sx - r * dc_x1) * inv_c2;
    float dr_x2 = (-0.5f * sx - r * dc_x2) * inv_c2;
    float dr_y1 = (-0.5f * sy - r * dc_y1) * inv_c2;
    float dr_y2 = (-0.5f * sy - r * dc_y2) * inv_c2;
    // d v : v = k*(A2 - A1)^2, A1 = atan(w1/h1); dA1/dw1 = h1/(w1^2+h1^2), dA1/dh1 = -w1/(w1^2+h1^2)
    float den = 1.f / (w1 * w1 + h1 * h1);
    float dv_a1 = -2.f * k4pi2 * dat;
    float dv_w1 = dv_a1 * h1 * den, dv_h1 = -dv_a1 * w1 * den;
    g[0] = dq_x1 - dr_x1 - alpha * (-dv_w1);
    g[2] = dq_x2 - dr_x2 - alpha * (dv_w1);
    g[1] = dq_y1 - dr_y1 - alpha * (-dv_h1);
    g[3] = dq_y2 - dr_y2 - alpha * (dv_h1);
    return ciou;
}

// ---------------------------------------------------------------------------
// Fast transcendental building blocks of the tolerance section.  Device: one MUFU each
// (ex2.approx / rcp.approx, ~2^-22 relative error); host: libm, for the CPU unit tests.
// ---------------------------------------------------------------------------
constexpr float kLog2e = 1.4426950408889634f;

YS_HD float ex2_fast(float a) {
#ifdef __CUDA_ARCH__
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
#else
    return exp2f(a);
#endif
}
YS_HD float rcp_fast(float a) {
#ifdef __CUDA_ARCH__
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
#else
    return 1.0f / a;
#endif
}

// sigmoid by one ex2 and one rcp (+ a Newton step): relative error ~1e-7, tolerance section only
YS_HD float sigmoid_fast(float x) {
    const float e = ex2_fast(-x * kLog2e);
    const float s = 1.0f + e;
    const float r = rcp_fast(s);
    return e < 3.0e38f ? r * fmaf(-s, r, 2.0f) : 0.0f;      // exp overflow: 1/inf = 0 (the Newton step would give NaN)
}

// log1p(e)/e on e in [0,1]: degree-7 minimax polynomial (Lawson fit; max relative error of e*P(e) vs log1p(e)
// 3.7e-7 in fp32 Horner form, -1.6e-7 as e -> 0).  No MUFU, no branch.
YS_HD float log1p_over_e(float e) {
    float p = -8.783297754e-03f;
    p = fmaf(p, e, 4.494444638e-02f);
    p = fmaf(p, e, -1.088579088e-01f);
    p = fmaf(p, e, 1.782550513e-01f);
    p = fmaf(p, e, -2.452358495e-01f);
    p = fmaf(p, e, 3.328017316e-01f);
    p = fmaf(p, e, -4.999769926e-01f);
    p = fmaf(p, e, 9.999998357e-01f);
    return p;
}

// Numerically stable BCE-with-logits and sigmoid sharing one exponential.
// loss = max(x,0) - x*t + log1p(exp(-|x|)) ; sig = sigmoid(x)
YS_HD void bce_logits(float x, float t, float &loss, float &sig) {
    const float e = ex2_fast(fabsf(x) * -kLog2e);     // exp(-|x|) in (0,1]
    const float inv = rcp_fast(1.0f + e);              // sigmoid(|x|)
    sig = x >= 0.f ? inv : e * inv;
    loss = fmaf(e, log1p_over_e(e), fmaf(-x, t, fmaxf(x, 0.f)));
}

// Softmax statistics of 16 logits in the exp2 domain: returns the expectation d = sum_j j*softmax(z)_j
// (losses/yolo_v8_loss.py:534-536); m2 = max(z)*log2e and inv_s = 1/sum_j 2^(z_j*log2e - m2), so that
// softmax(z)_j = ex2(z_j*log2e - m2) * inv_s and logsumexp(z) = (m2 + log2(sum)) * ln2.
YS_HD float dfl_stats16(const float (&z)[16], float &m2, float &inv_s) {
    float m = z[0];
#pragma unroll
    for (int j = 1; j < 16; ++j) m = fmaxf(m, z[j]);
    const float mm = m * kLog2e;
    float s = 0.f, d = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const float e = ex2_fast(fmaf(z[j], kLog2e, -mm));
        s += e;
        d = fmaf(e, (float)j, d);
    }
    m2 = mm;
    // 1/s: one MUFU + one Newton step (the IEEE division was the longest dependency chain of the decode; s is in [1, 16])
    const float r = rcp_fast(s);
    inv_s = r * fmaf(-s, r, 2.0f);
    return d * inv_s;
}

}  // namespace ys
