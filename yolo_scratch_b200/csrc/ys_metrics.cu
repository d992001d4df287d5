// ys_metrics.cu -- greedy detection-to-GT matching for mAP, on the device.
// Replaces the matching loops of evaluate_detections (utils/metrics.py:147-205 and :211-256 of the reference): per image,
// predictions in score-descending order; each takes the best-IoU GT that is still unmatched (first maximum, class
// agnostic as in the reference); it is a true positive iff that IoU >= threshold, and only then is the GT consumed.
// The reference runs this loop once per IoU threshold with one .item() per prediction; here one warp per
// (image, threshold) walks the predictions, its lanes covering the GT.
//
// IoU form: calculate_iou_batch (utils/metrics.py:49-81) = inter / (a1 + a2 - inter + 1e-6), fp32, same operation order
// (iou_nms in ys_math.cuh).  The comparison `best_iou >= thr` happens in DOUBLE like the reference's
// (a Python float from .item() against a numpy float64 from np.arange).
#include "ys_common.cuh"

namespace ys {

constexpr int kMatchMaxThr = 16;
constexpr int kMatchMaxGt = 1024;        // GT per image held in shared memory / lane bitmasks (32 chunks of 32)

struct MatchThresholds {
    double v[kMatchMaxThr];
    int n;
};

__global__ void __launch_bounds__(32 * kMatchMaxThr)
k_match(const float4 *__restrict__ pred_boxes, const float *__restrict__ pred_scores, const int *__restrict__ pred_count,
        int pred_stride, const float4 *__restrict__ gt_boxes, const int *__restrict__ gt_off, float conf_thr,
        MatchThresholds T, unsigned char *__restrict__ tp, int *__restrict__ n_valid) {
    __shared__ float4 s_gt[kMatchMaxGt];
    __shared__ float s_area[kMatchMaxGt];
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, t = tid >> 5;
    const int g0 = gt_off[b], G = gt_off[b + 1] - g0;
    const int np = pred_count[b];
    for (int j = tid; j < G; j += blockDim.x) {
        const float4 q = gt_boxes[g0 + j];
        s_gt[j] = q;
        s_area[j] = area_plain(q);
    }
    __syncthreads();
    if (t >= T.n) return;
    const float4 *pb = pred_boxes + (size_t)b * pred_stride;
    const float *ps = pred_scores + (size_t)b * pred_stride;
    unsigned char *out = tp + ((size_t)b * pred_stride) * T.n + t;
    const double thr = T.v[t];
    unsigned matched = 0u;                  // bit c: GT lane + 32c is consumed
    int left = G;                           // unmatched GT
    int i = 0;
    for (; i < np; ++i) {
        const float sc = ps[i];
        if (!(sc >= conf_thr)) break;       // sorted descending: the valid predictions are a prefix (:116)
        bool is_tp = false;
        if (left > 0) {
            const float4 p = pb[i];
            const float area_p = area_plain(p);
            unsigned long long best = 0ull;
            bool any = false;
            for (int c = 0; c * 32 < G; ++c) {
                const int j = c * 32 + lane;
                if (j < G && !((matched >> c) & 1u)) {
                    const float iou = iou_nms(p, area_p, s_gt[j], s_area[j]);
                    // order-preserving key of a float that may be negative or NaN-free: IoU of sane boxes is >= 0
                    const unsigned long long key = argmax_key(iou < 0.f ? 0.f : iou, (unsigned)j);
                    if (!any || key > best) { best = key; any = true; }
                }
            }
            best = warp_max(any ? best : 0ull);
            const unsigned jb = argmax_key_index(best);
            const float biou = __uint_as_float((unsigned)(best >> 32));
            if ((double)biou >= thr) {
                is_tp = true;
                --left;
                if ((jb & 31u) == (unsigned)lane) matched |= 1u << (jb >> 5);
            }
        }
        if (lane == 0) out[(size_t)i * T.n] = is_tp ? 1 : 0;
    }
    if (t == 0 && lane == 0) n_valid[b] = i;
}

}  // namespace ys

using namespace ys;

extern "C" int ys_match_detections(const float *pred_boxes, const float *pred_scores, const int32_t *pred_count, int B,
                                   int pred_stride, const float *gt_boxes, const int32_t *gt_off, int max_gt,
                                   float conf_thr, const double *iou_thresholds_host, int n_thresholds, uint8_t *tp,
                                   int32_t *n_valid, void *stream_v) {
    YS_REQUIRE(B >= 1 && pred_stride >= 1, YS_ERR_BAD_ARG, "B and pred_stride must be >= 1");
    YS_REQUIRE(pred_boxes && pred_scores && pred_count && gt_off && tp && n_valid && iou_thresholds_host, YS_ERR_BAD_ARG,
               "null pointer");
    YS_REQUIRE(n_thresholds >= 1 && n_thresholds <= kMatchMaxThr, YS_ERR_UNSUPPORTED, "1..%d IoU thresholds", kMatchMaxThr);
    YS_REQUIRE(max_gt >= 0 && max_gt <= kMatchMaxGt, YS_ERR_UNSUPPORTED, "at most %d GT per image", kMatchMaxGt);
    YS_REQUIRE((((uintptr_t)pred_boxes | (uintptr_t)gt_boxes) & 15) == 0, YS_ERR_BAD_ARG, "boxes must be 16-byte aligned");
    MatchThresholds T;
    T.n = n_thresholds;
    for (int i = 0; i < kMatchMaxThr; ++i) T.v[i] = i < n_thresholds ? iou_thresholds_host[i] : 2.0;
    k_match<<<B, 32 * n_thresholds, 0, (cudaStream_t)stream_v>>>((const float4 *)pred_boxes, pred_scores, pred_count,
                                                                  pred_stride, (const float4 *)gt_boxes, gt_off, conf_thr, T,
                                                                  tp, n_valid);
    YS_CHECK_LAUNCH("k_match");
    return YS_OK;
}

// ---------------------------------------------------------------------------------------------------
// cda_nms: greedy suppression with the Gaussian-angle distance dGA between the ellipses inscribed in
// the boxes, at most max_det survivors (reference utils/metrics.py:440-620: _calcYmat, _calc_dGA, cda_nms).
// Candidates arrive score-descending; a candidate is dropped iff dGA(candidate, kept) < dga_thresh for some kept one.
// fp64 like the reference (numpy).  One warp per image: lane k holds kept detection k (max_det <= 32), the candidates
// are walked in order until max_det are kept.  Parity with numpy's LAPACK-based det/inv is at tolerance level
// (closed-form 2x2 determinant and inverse here), which only matters for a pair sitting exactly on the threshold.
// ---------------------------------------------------------------------------------------------------
namespace ys {

struct Ellipse {
    double y11, y12, y21, y22, det, xc, yc;
};

__device__ __forceinline__ Ellipse ellipse_of(float4 bx, float phi_f) {
    // fp32 box arithmetic exactly as the torch code (:555-566), then double
    const float w = bx.z - bx.x, h = bx.w - bx.y;
    const float hw = w / 2.0f, hh = h / 2.0f;
    const double a = (double)fmaxf(hw, hh), b = (double)fminf(hw, hh);
    Ellipse e;
    e.xc = (double)((bx.x + bx.z) / 2.0f);
    e.yc = (double)((bx.y + bx.w) / 2.0f);
    const double phi = (double)phi_f;
    const double c = cos(phi), s = sin(phi);
    const double ia = 1.0 / (a * a), ib = 1.0 / (b * b);
    // (R diag(ia, ib)) R^T, products in the reference's association order (:451-454)
    const double m11 = c * ia, m12 = -s * ib, m21 = s * ia, m22 = c * ib;
    e.y11 = m11 * c + m12 * -s;
    e.y12 = m11 * s + m12 * c;
    e.y21 = m21 * c + m22 * -s;
    e.y22 = m21 * s + m22 * c;
    e.det = e.y11 * e.y22 - e.y12 * e.y21;
    return e;
}

__device__ __forceinline__ double dga_of(const Ellipse &i, const Ellipse &j) {
    const double kHalfPi = 1.5707963267948966;
    const double s11 = i.y11 + j.y11, s12 = i.y12 + j.y12, s21 = i.y21 + j.y21, s22 = i.y22 + j.y22;
    const double det_s = s11 * s22 - s12 * s21;
    if (det_s <= 0.0 || i.det <= 0.0 || j.det <= 0.0) return kHalfPi;                 // :470
    const double mult = 4.0 * sqrt(i.det * j.det) / det_s;
    const double dx = i.xc - j.xc, dy = i.yc - j.yc;
    // v = Yj d ; u = inv(Ysum) v ; t = Yi u ; e = -0.5 d.t   (:478)
    const double v1 = j.y11 * dx + j.y12 * dy, v2 = j.y21 * dx + j.y22 * dy;
    const double u1 = (s22 * v1 - s12 * v2) / det_s, u2 = (-s21 * v1 + s11 * v2) / det_s;
    const double t1 = i.y11 * u1 + i.y12 * u2, t2 = i.y21 * u1 + i.y22 * u2;
    const double ex = -0.5 * (dx * t1 + dy * t2);
    double ct = mult * exp(ex);
    ct = ct > 1.0 ? 1.0 : (ct < -1.0 ? -1.0 : ct);
    return acos(ct);
}

__global__ void __launch_bounds__(128)
k_cda_nms(const float4 *__restrict__ boxes, const float *__restrict__ angles, const int *__restrict__ count, int B,
          int stride, int max_det, double thr, int *__restrict__ keep_idx, int *__restrict__ keep_count) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    const int n = count[b];
    const float4 *bx = boxes + (size_t)b * stride;
    const float *an = angles ? angles + (size_t)b * stride : nullptr;
    Ellipse mine = {};       // lane k: kept detection k
    int nk = 0;
    for (int c = 0; c < n && nk < max_det; ++c) {
        const Ellipse cand = ellipse_of(bx[c], an ? an[c] : 0.f);
        const bool dup = lane < nk && dga_of(cand, mine) < thr;
        if (!__any_sync(0xffffffffu, dup)) {
            if (lane == nk) mine = cand;
            if (lane == 0) keep_idx[(size_t)b * max_det + nk] = c;
            ++nk;
        }
    }
    if (lane == 0) keep_count[b] = nk;
}

}  // namespace ys

extern "C" int ys_cda_nms(const float *boxes, const float *angles, const int32_t *count, int B, int stride, int max_det,
                          double dga_thresh, int32_t *keep_idx, int32_t *keep_count, void *stream_v) {
    YS_REQUIRE(B >= 1 && stride >= 1, YS_ERR_BAD_ARG, "B and stride must be >= 1");
    YS_REQUIRE(boxes && count && keep_idx && keep_count, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE(max_det >= 1 && max_det <= 32, YS_ERR_UNSUPPORTED, "max_det must be in [1, 32]");
    YS_REQUIRE(((uintptr_t)boxes & 15) == 0, YS_ERR_BAD_ARG, "boxes must be 16-byte aligned");
    k_cda_nms<<<(B + 3) / 4, 128, 0, (cudaStream_t)stream_v>>>((const float4 *)boxes, angles, count, B, stride, max_det,
                                                               dga_thresh, keep_idx, keep_count);
    YS_CHECK_LAUNCH("k_cda_nms");
    return YS_OK;
}
