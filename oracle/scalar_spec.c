/* ORACLE (test infrastructure, not product code).
 *
 * Plain-C scalar restatement of the index-producing parts of the hot path of
 * Pratye/yolo-scratch: the task-aligned assigner, greedy NMS and
 * the two pairwise IoU forms.  Compiled by oracle/Makefile with
 * -ffp-contract=off so every +,-,*,/ rounds to fp32 exactly once, which is what
 * the reference's chain of separate ATen elementwise kernels does.  It is
 * checked against the reference-generated fixtures in tests/golden by
 * tests/test_oracle_golden_c.py.  Only tests/, smoke() and bench.py's CPU legs
 * may load it.
 *
 * file:line citations are relative to /root/reference/yolo_scratch_cuda/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline float fmin2(float a, float b) { return a < b ? a : b; }
static inline float fmax2(float a, float b) { return a > b ? a : b; }

/* losses/yolo_v8_loss.py:31-44 (xywh=False, plain IoU, eps 1e-7), then clamp_(0) at :199 */
float orc_iou_assign(const float *p, const float *q) {
    const float eps = 1e-7f;
    float w1 = p[2] - p[0], h1 = p[3] - p[1] + eps;
    float w2 = q[2] - q[0], h2 = q[3] - q[1] + eps;
    float iw = fmin2(p[2], q[2]) - fmax2(p[0], q[0]);
    float ih = fmin2(p[3], q[3]) - fmax2(p[1], q[1]);
    if (iw < 0.f) iw = 0.f;
    if (ih < 0.f) ih = 0.f;
    float inter = iw * ih;
    float uni = w1 * h1 + w2 * h2 - inter + eps;
    float iou = inter / uni;
    return iou < 0.f ? 0.f : iou; /* NaN stays NaN, as clamp_ does */
}

/* train_yolo11_cuda.py:418-435 and utils/metrics.py:65-79: eps 1e-6 on the union only */
float orc_iou_nms(const float *a, const float *b) {
    float x1 = fmax2(a[0], b[0]), y1 = fmax2(a[1], b[1]);
    float x2 = fmin2(a[2], b[2]), y2 = fmin2(a[3], b[3]);
    float iw = x2 - x1, ih = y2 - y1;
    if (iw < 0.f) iw = 0.f;
    if (ih < 0.f) ih = 0.f;
    float inter = iw * ih;
    float aa = (a[2] - a[0]) * (a[3] - a[1]);
    float ab = (b[2] - b[0]) * (b[3] - b[1]);
    float uni = aa + ab - inter;
    return inter / (uni + 1e-6f);
}

void orc_iou_matrix(int n, int m, const float *b1, const float *b2, float *out) {
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) out[(size_t)i * m + j] = orc_iou_nms(b1 + 4 * i, b2 + 4 * j);
}

/* per-anchor resolve, losses/yolo_v8_loss.py:226-244 */
static void resolve_rows(int A, int G, uint8_t *mp, const float *ov, int32_t *tgi, int32_t *fg) {
    for (int a = 0; a < A; ++a) {
        uint8_t *row = mp + (size_t)a * G;
        const float *o = ov + (size_t)a * G;
        int cnt = 0;
        for (int g = 0; g < G; ++g) cnt += row[g];
        if (cnt > 1) {
            int best = 0;
            for (int g = 1; g < G; ++g)
                if (o[g] > o[best]) best = g; /* first maximum */
            memset(row, 0, (size_t)G);
            row[best] = 1;
            cnt = 1;
        }
        int first = 0;
        for (int g = 0; g < G; ++g)
            if (row[g]) { first = g; break; }
        tgi[a] = first;
        fg[a] = cnt;
    }
}

static int argmax_col(int A, int G, const float *ov, int g) {
    int k = 0;
    float best = ov[g];
    for (int a = 1; a < A; ++a) {
        float v = ov[(size_t)a * G + g];
        if (v > best) { best = v; k = a; }
    }
    return k;
}

/* TaskAlignedAssigner.forward, losses/yolo_v8_loss.py:78-180; steps as in SURVEY §8.1.
 * Outputs per (b,a): tgi (slot index), fg (0/1), norm (the single non-zero of the
 * target_scores row, at class clamp(label)), label (float), tbox (4 floats). */
int orc_assign(int B, int A, int G, int nc, const float *pd_scores, const float *pd_bboxes,
               const float *anc, const float *gt_labels, const float *gt_bboxes, const float *mask_gt,
               float alpha, float beta, float eps, int32_t *tgi_out, uint8_t *fg_out, float *norm_out,
               float *label_out, float *tbox_out) {
    if (G <= 0) return 1;
    float *ov = (float *)malloc(sizeof(float) * (size_t)A * G);
    uint8_t *mp = (uint8_t *)malloc((size_t)A * G);
    int32_t *tgi = (int32_t *)malloc(sizeof(int32_t) * A);
    int32_t *fg = (int32_t *)malloc(sizeof(int32_t) * A);
    if (!ov || !mp || !tgi || !fg) return 2;
    for (int b = 0; b < B; ++b) {
        const float *pb = pd_bboxes + (size_t)b * A * 4;
        const float *gb = gt_bboxes + (size_t)b * G * 4;
        const float *mg = mask_gt + (size_t)b * G;
        const float *gl = gt_labels + (size_t)b * G;
        /* steps 1-2 */
        for (int a = 0; a < A; ++a) {
            float ax = anc[2 * a], ay = anc[2 * a + 1];
            for (int g = 0; g < G; ++g) {
                const float *q = gb + 4 * g;
                ov[(size_t)a * G + g] = orc_iou_assign(pb + 4 * a, q);
                float d = fmin2(fmin2(ax - q[0], ay - q[1]), fmin2(q[2] - ax, q[3] - ay));
                float in = d > eps ? 1.f : 0.f;
                mp[(size_t)a * G + g] = (in * mg[g]) != 0.f; /* :192 product of two 0/1 floats */
            }
        }
        /* step 3 (:117-139) */
        for (int g = 0; g < G; ++g) {
            if (!(mg[g] != 0.f)) continue;
            int any = 0;
            for (int a = 0; a < A && !any; ++a) any = mp[(size_t)a * G + g];
            if (!any) mp[(size_t)argmax_col(A, G, ov, g) * G + g] = 1;
        }
        /* step 4 */
        resolve_rows(A, G, mp, ov, tgi, fg);
        /* step 5 (:146-162), sequential */
        for (int g = 0; g < G; ++g) {
            if (!(mg[g] != 0.f)) continue;
            int owned = 0;
            for (int a = 0; a < A && !owned; ++a) owned = (fg[a] > 0 && tgi[a] == g);
            if (!owned) {
                int k = argmax_col(A, G, ov, g);
                mp[(size_t)k * G + g] = 1;
                tgi[k] = g;
                fg[k] = 1;
            }
        }
        /* step 6 */
        resolve_rows(A, G, mp, ov, tgi, fg);
        /* steps 7-8 */
        for (int a = 0; a < A; ++a) {
            int g = tgi[a];
            float lab = gl[g];
            if (lab < 0.f) lab = 0.f;
            if (lab > (float)nc) lab = (float)nc;
            size_t o = (size_t)b * A + a;
            tgi_out[o] = g;
            fg_out[o] = fg[a] > 0;
            label_out[o] = lab;
            memcpy(tbox_out + 4 * o, gb + 4 * g, 4 * sizeof(float));
            float nrm = 0.f;
            if (fg[a] > 0) {
                float s = pd_scores[((size_t)b * A + a) * nc + (int)gl[g]];
                float v = ov[(size_t)a * G + g];
                float am = powf(s, alpha) * powf(v, beta);
                nrm = am * v / (am + eps);
            }
            norm_out[o] = nrm;
        }
    }
    free(ov); free(mp); free(tgi); free(fg);
    return 0;
}

/* nms_simple, train_yolo11_cuda.py:361-399.  Order: score descending, ties -> lower
 * index first (the reference's argsort is unstable).  Suppress when
 * IoU > thr (":396 keeps <= thr").  If labels != NULL only same-label pairs
 * suppress (class-aware mode).  Returns the number kept. */
typedef struct { float s; int32_t i; } orc_key;
static int key_cmp(const void *a, const void *b) {
    const orc_key *x = (const orc_key *)a, *y = (const orc_key *)b;
    if (x->s > y->s) return -1;
    if (x->s < y->s) return 1;
    return (x->i > y->i) - (x->i < y->i);
}

int orc_nms(int n, const float *boxes, const float *scores, const int32_t *labels, float thr, int32_t *keep) {
    if (n <= 0) return 0;
    orc_key *k = (orc_key *)malloc(sizeof(orc_key) * n);
    uint8_t *dead = (uint8_t *)calloc(n, 1);
    for (int i = 0; i < n; ++i) { k[i].s = scores[i]; k[i].i = i; }
    qsort(k, n, sizeof(orc_key), key_cmp);
    int nk = 0;
    for (int r = 0; r < n; ++r) {
        if (dead[r]) continue;
        int i = k[r].i;
        keep[nk++] = i;
        for (int c = r + 1; c < n; ++c) {
            if (dead[c]) continue;
            int j = k[c].i;
            if (labels && labels[i] != labels[j]) continue;
            if (!(orc_iou_nms(boxes + 4 * i, boxes + 4 * j) <= thr)) dead[c] = 1;
        }
    }
    free(k); free(dead);
    return nk;
}

/* decode_predictions_for_metrics, train_yolo11_cuda.py:296-356, one image:
 * pred (A, 4+nc) xywh px + scores -> filtered xyxy / score / class / source anchor.
 * Returns the number of candidates with best score > conf. */
int orc_filter_candidates(int A, int nc, const float *pred, float conf, float *xyxy, float *score,
                          int32_t *cls, int32_t *src) {
    int n = 0;
    for (int a = 0; a < A; ++a) {
        const float *r = pred + (size_t)a * (4 + nc);
        int best = 0;
        for (int c = 1; c < nc; ++c)
            if (r[4 + c] > r[4 + best]) best = c;
        float s = r[4 + best];
        if (!(s > conf)) continue;
        float hw = r[2] / 2, hh = r[3] / 2;
        xyxy[4 * n + 0] = r[0] - hw;
        xyxy[4 * n + 1] = r[1] - hh;
        xyxy[4 * n + 2] = r[0] + hw;
        xyxy[4 * n + 3] = r[1] + hh;
        score[n] = s;
        cls[n] = best;
        src[n] = a;
        ++n;
    }
    return n;
}
