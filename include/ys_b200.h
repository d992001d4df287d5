/* ys_b200.h -- C ABI of the B200-native detection-head hot path.
 *
 * One shared library (yolo_scratch_b200/libys_b200.so, built by
 * yolo_scratch_b200/build.py with nvcc -gencode arch=compute_100a,code=sm_100a).
 * The reference (Pratye/yolo-scratch, yolo_scratch_cuda) is pure Python and has no
 * FFI; every entry point below names the reference function(s) it replaces
 * (paths relative to /root/reference/yolo_scratch_cuda/).  INTEGRATION.md shows the
 * ctypes binding a maintainer would add.
 *
 * Conventions
 *  - All tensor pointers are DEVICE pointers unless the parameter name ends in
 *    _host.  Plain pointers + sizes only; no torch types.
 *  - No allocation, no host synchronisation, no state a result depends on: scratch
 *    lives in a caller-provided workspace (size from ys_*_workspace_bytes; zero it
 *    once after allocation), the last argument is the cudaStream_t to launch on (as
 *    void*).  Process-wide data are limited to per-device caches held in atomics
 *    (SM count, shared-memory opt-in done); any thread, any device.  Calls are
 *    CUDA-graph-capturable.
 *  - Return 0 on success; YS_ERR_* (<0) for bad arguments; a positive value is a
 *    cudaError_t from a launch.  ys_last_error() gives a thread-local message.
 *  - Never throws across the ABI.  Re-entrant across streams as long as each
 *    in-flight call has its own workspace.
 */
#ifndef YS_B200_H
#define YS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YS_OK 0
#define YS_ERR_BAD_ARG (-1)
#define YS_ERR_WORKSPACE (-2)
#define YS_ERR_UNSUPPORTED (-3)

#define YS_MAX_LEVELS 8

/* Library identity. */
int ys_version(void);
const char *ys_last_error(void);
/* SM count / arch the library was built for and sees at run time (host-side query;
 * returns 0 and fills the ints, or a cudaError_t). */
int ys_device_info(int *sm_count, int *cc_major, int *cc_minor);

/* ---------------------------------------------------------------------------
 * Fused detection loss: assign + BCE + CIoU + DFL, forward AND backward.
 * Replaces v8DetectionLoss.__call__ (losses/yolo_v8_loss.py:372-499) including
 * preprocess (:501-527), make_anchors (:541-552), bbox_decode (:529-538),
 * TaskAlignedAssigner.forward (:78-180) and BboxLoss.forward (:280-324), plus
 * the autograd backward of all of it (train_yolo11_cuda.py:57 loss.backward()).
 *
 *  level_x_host[n_levels]    device pointers to the NCHW head outputs (B, 4*reg_max+nc, h, w), fp32, contiguous;
 *                            any h, w (levels with h*w % 4 != 0 or a base pointer that is not 16-byte aligned take
 *                            plain-load tile loaders instead of TMA)
 *  level_grad_host[n_levels] device pointers to same-shaped gradient buffers, or NULL => forward only
 *  level_h/w/stride_host     per-level sizes and strides (host arrays)
 *  batch_idx (N) i64, cls (N) i64, bboxes (N,4) fp32 normalised xyxy: the flat GT list of the batch dict
 *                            (datasets/crater_dataset_cuda.py:336-341); any order; rows with batch_idx
 *                            outside [0,B) are ignored as the reference's mask does (:519)
 *  scale_x, scale_y          pixel scale applied to (x1,x2) and (y1,y2): the reference uses
 *                            imgsz = feats[0].shape[2:]*stride[0] and scales x by imgsz[0], y by imgsz[1] (:400,:515)
 *  hyp_box/cls/dfl           7.5 / 0.5 / 1.5 (:368-370); alpha/beta 0.5 / 4.0 (:363)
 *  loss_out[4]               device: [loss.sum()*B, box*hyp, cls*hyp, dfl*hyp]  (:495-499)
 *  Gradients are d(loss.sum()*B)/d(level_x) (i.e. upstream grad 1.0).
 *  Optional exports (may be NULL): pred_boxes_px (B,A,4) decoded boxes in pixels (:412),
 *  target_gt_idx (B,A) i32 (-1 = background), norm (B,A) the per-anchor target score (:173-178).
 *  stages: 0 or YS_STAGE_ALL runs the whole path; a subset re-runs only those kernels on the state a
 *  previous full call left in the SAME workspace (bench.py times the dominant kernel alone this way).
 */
#define YS_STAGE_PACK 1   /* k_pack_gt  flat GT list -> per-image CSR, pixel boxes */
#define YS_STAGE_ASSIGN 2 /* k_assign   decode + scan + forced-assignment fix-up + target_scores_sum */
#define YS_STAGE_LOSS 4   /* k_loss     BCE + CIoU + DFL forward and backward + ordered final reduction */
#define YS_STAGE_ALL 7
#define YS_STAGE_DIAG_NOFIXUP 256 /* profiling only: k_assign skips the forced-assignment fix-up (results then differ) */
size_t ys_detection_loss_workspace_bytes(int B, int A, int n_gt, int nc);
int ys_detection_loss(int n_levels, const float *const *level_x_host, float *const *level_grad_host,
                      const int *level_h_host, const int *level_w_host, const float *level_stride_host,
                      int B, int nc, int reg_max,
                      const int64_t *batch_idx, const int64_t *cls, const float *bboxes, int n_gt,
                      float scale_x, float scale_y,
                      float hyp_box, float hyp_cls, float hyp_dfl, float alpha, float beta,
                      float *loss_out,
                      float *pred_boxes_px, int32_t *target_gt_idx, float *norm, int stages,
                      void *workspace, size_t workspace_bytes, void *stream);

/* In-place `grad *= *scale` unless *scale == 1.0f (device scalar; no host sync).
 * Used by the autograd wrapper when the upstream gradient is not 1. */
int ys_scale_inplace(float *data, size_t n, const float *scale_dev, void *stream);

/* ---------------------------------------------------------------------------
 * TaskAlignedAssigner.forward on dense inputs (losses/yolo_v8_loss.py:78-180,
 * helpers :182-270).  Shapes as the reference: pd_scores (B,A,nc) probabilities,
 * pd_bboxes (B,A,4) px xyxy, anc_points (A,2) px, gt_labels (B,G) f32,
 * gt_bboxes (B,G,4), mask_gt (B,G) f32.  Outputs: target_labels (B,A) f32,
 * target_bboxes (B,A,4), target_scores (B,A,nc) (may be NULL), fg_mask (B,A) u8,
 * target_gt_idx (B,A) i64, norm (B,A) f32 (may be NULL).  G must be > 0 (the G==0 return of
 * :100-108 is produced by the Python shim without a launch).
 */
size_t ys_assign_workspace_bytes(int B, int A, int G);
int ys_assign(int B, int A, int G, int nc,
              const float *pd_scores, const float *pd_bboxes, const float *anc_points,
              const float *gt_labels, const float *gt_bboxes, const float *mask_gt,
              float alpha, float beta, float eps,
              float *target_labels, float *target_bboxes, float *target_scores,
              uint8_t *fg_mask, int64_t *target_gt_idx, float *norm,
              void *workspace, size_t workspace_bytes, void *stream);

/* The same entry point with the two non-default assignment modes north_star names (the reference runs
 * neither: it stores topk and never reads it, losses/yolo_v8_loss.py:70-76, and uses plain IoU, :199):
 *  topk > 0          per GT, only the topk anchors with the largest alignment metric among the anchors inside it (metric
 *                    > 0) stay in the positive mask before the forced-assignment loops (k_topk_select: radix select on
 *                    the metric bits + warp ballots).  Parity unpinned by the reference; oracle: oracle/ref_port.assign.
 *  ciou_assign != 0  the pair metric is clamp(bbox_iou(CIoU=True), 0) (:46-56) instead of clamp(IoU, 0).
 * topk = 0, ciou_assign = 0 is ys_assign. */
size_t ys_assign_ex_workspace_bytes(int B, int A, int G, int topk);
int ys_assign_ex(int B, int A, int G, int nc,
                 const float *pd_scores, const float *pd_bboxes, const float *anc_points,
                 const float *gt_labels, const float *gt_bboxes, const float *mask_gt,
                 float alpha, float beta, float eps, int topk, int ciou_assign,
                 float *target_labels, float *target_bboxes, float *target_scores,
                 uint8_t *fg_mask, int64_t *target_gt_idx, float *norm,
                 void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------
 * Inference decode: DFL.forward + Detect.inference + dist2bbox(xywh=True, dim=1)
 * (models/yolo11_modules.py:180-192, :248-266, :303-312).
 * out: (B, 4+nc, A) channels-first [cx,cy,w,h]*stride, sigmoid(cls)  (layout 0)
 *   or (B, A, 4+nc) anchor-major, the layout decode_predictions_for_metrics documents (layout 1).
 */
int ys_decode_infer(int n_levels, const float *const *level_x_host, const int *level_h_host,
                    const int *level_w_host, const float *level_stride_host, int B, int nc, int reg_max,
                    float *out, int layout, void *stream);

/* ---------------------------------------------------------------------------
 * Batched NMS: decode_predictions_for_metrics + nms_simple + calculate_iou_batch_simple
 * (train_yolo11_cuda.py:265-358, :361-399, :402-437).
 *  pred     (B, A, 4+nc) if layout==1, (B, 4+nc, A) if layout==0; xywh px + class probabilities
 *  per image: best class (first max), score > conf, xywh->xyxy, greedy NMS in score-descending
 *  order (ties: lower anchor index first), suppress when IoU > iou_thr; class_aware != 0 restricts
 *  suppression to same-class pairs (non-default).
 *  Any A (above 16384 candidates per image the sort runs as segment sorts + merge).
 *  Outputs, each with row stride A per image, first keep_count[b] rows valid, score-descending:
 *   keep_anchor (B,A) i32 source anchor index, keep_boxes (B,A,4) = clamp(xyxy/img_size,0,1),
 *   keep_scores (B,A), keep_labels (B,A) i32, keep_count (B) i32.
 */
size_t ys_nms_workspace_bytes(int B, int A);
int ys_nms_batched(const float *pred, int layout, int B, int A, int nc,
                   float conf_thr, float iou_thr, float img_size, int class_aware,
                   int32_t *keep_anchor, float *keep_boxes, float *keep_scores, int32_t *keep_labels,
                   int32_t *keep_count, void *workspace, size_t workspace_bytes, void *stream);

/* nms_simple on one already-filtered set (train_yolo11_cuda.py:361-399): boxes (N,4) xyxy, scores (N).
 * labels may be NULL (class-agnostic).  keep (N) i32 receives indices into the input, score-descending;
 * keep_count (1) i32. */
int ys_nms_single(const float *boxes, const float *scores, const int32_t *labels, int N, float iou_thr,
                  int32_t *keep, int32_t *keep_count, void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------
 * Pairwise IoU matrices.
 *  kind 0: utils.metrics.calculate_iou_batch / calculate_iou_batch_simple
 *          (utils/metrics.py:49-81, train_yolo11_cuda.py:402-437): inter/(a1+a2-inter+1e-6)
 *  kind 1: bbox_iou(xywh=False) broadcast form used by the assigner (losses/yolo_v8_loss.py:31-44)
 *  boxes1 (N,4), boxes2 (M,4) -> out (N,M)
 */
int ys_pairwise_iou(const float *boxes1, int N, const float *boxes2, int M, int kind, float *out, void *stream);

/* Elementwise bbox_iou on paired boxes (losses/yolo_v8_loss.py:12-61): box1,box2 (n,4) -> out (n).
 * mode 0 IoU, 1 GIoU, 2 DIoU, 3 CIoU; xywh != 0 => inputs are cx,cy,w,h.
 * If grad_box1 != NULL also writes d(sum_i out_i * gout_i)/d box1 (n,4) (alpha constant, :54-55);
 * gout may be NULL (= ones). */
int ys_bbox_iou(const float *box1, const float *box2, int n, int xywh, int mode, float eps, float *out,
                const float *gout, float *grad_box1, void *stream);

/* ---------------------------------------------------------------------------
 * BboxLoss.forward with _df_loss and bbox2dist (losses/yolo_v8_loss.py:280-310, :312-324, :327-330) on dense,
 * already-assigned tensors (stand-alone use; ys_detection_loss fuses the same arithmetic into its own kernels).
 *  pred_dist (n, 4*reg_max) logits, pred_bboxes (n,4) xyxy and target_bboxes (n,4) xyxy in grid units,
 *  anchor_points (A,2) grid units with n = B*A (row i uses anchor i % A), target_scores (n, nc),
 *  fg_mask (n) u8, target_scores_sum: DEVICE scalar.
 *  loss_out[2] = {loss_iou, loss_dfl} (loss_dfl = 0 when reg_max == 1, :307-308).
 *  grad_pred_bboxes (n,4) = d loss_iou / d pred_bboxes and grad_pred_dist (n, 4*reg_max) = d loss_dfl / d pred_dist
 *  (zero rows for background); either may be NULL.
 */
size_t ys_bbox_loss_workspace_bytes(int n);
int ys_bbox_loss(const float *pred_dist, const float *pred_bboxes, const float *anchor_points,
                 const float *target_bboxes, const float *target_scores, const uint8_t *fg_mask,
                 const float *target_scores_sum, int n, int A, int nc, int reg_max, float *loss_out,
                 float *grad_pred_bboxes, float *grad_pred_dist, void *workspace, size_t workspace_bytes, void *stream);

/* ys_bbox_loss without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments, results and error behaviour as ys_bbox_loss minus
 * workspace / workspace_bytes; the kernel's scratch (arrival counter + per-block partial sums) is a file-scope
 * __device__ array of 256 + 8 * YS_BBOX_LOSS_STATIC_MAX_ROWS / 64 bytes (128.25 KiB) that every GPU loading the library
 * holds until exit.  n > YS_BBOX_LOSS_STATIC_MAX_ROWS returns YS_ERR_UNSUPPORTED (use ys_bbox_loss).  Calls no
 * cudaMalloc / cudaFree or allocating library routine.  All calls on one device share the scratch: calls that overlap
 * on the GPU (two streams, or a replaying graph next to an eager call) race; serialise them. */
#define YS_BBOX_LOSS_STATIC_MAX_ROWS 1048576 /* n = B*A rows; B=64 at 1280 px (33600 anchors) is 2150400: use ys_bbox_loss */
int ys_bbox_loss_static_max_rows(void);
int ys_bbox_loss_static(const float *pred_dist, const float *pred_bboxes, const float *anchor_points,
                        const float *target_bboxes, const float *target_scores, const uint8_t *fg_mask,
                        const float *target_scores_sum, int n, int A, int nc, int reg_max, float *loss_out,
                        float *grad_pred_bboxes, float *grad_pred_dist, void *stream);

/* ---------------------------------------------------------------------------
 * Greedy detection-to-GT matching of evaluate_detections (utils/metrics.py:147-205, :211-256): per image,
 * predictions in score-descending order, each takes the best-IoU unmatched GT (first maximum, class agnostic); true
 * positive iff that IoU >= threshold (compared in double, as the reference does), and only then is the GT consumed.
 *  pred_boxes (B, pred_stride, 4) xyxy, pred_scores (B, pred_stride) SORTED DESCENDING per image, pred_count (B) i32
 *  (the layout ys_nms_batched writes); predictions with score < conf_thr are ignored (a suffix);
 *  gt_boxes (N,4) + gt_off (B+1) i32: per-image CSR list, at most 1024 GT per image (max_gt = the largest count);
 *  iou_thresholds_host: n_thresholds (<= 16) doubles on the HOST.
 *  tp (B, pred_stride, n_thresholds) u8: 1 = true positive, for the first n_valid[b] predictions of image b.
 */
int ys_match_detections(const float *pred_boxes, const float *pred_scores, const int32_t *pred_count, int B,
                        int pred_stride, const float *gt_boxes, const int32_t *gt_off, int max_gt, float conf_thr,
                        const double *iou_thresholds_host, int n_thresholds, uint8_t *tp, int32_t *n_valid,
                        void *stream);

/* ---------------------------------------------------------------------------
 * cda_nms (utils/metrics.py:484-620 with _calcYmat :440-453 and _calc_dGA :456-481): greedy suppression by the
 * Gaussian-angle distance between the ellipses inscribed in the boxes (fp64), at most max_det (<= 32) survivors.
 *  boxes (B, stride, 4) xyxy SORTED by score descending per image, angles (B, stride) radians or NULL (= 0),
 *  count (B) i32.  keep_idx (B, max_det) i32: positions in the sorted order; keep_count (B) i32.
 */
int ys_cda_nms(const float *boxes, const float *angles, const int32_t *count, int B, int stride, int max_det,
               double dga_thresh, int32_t *keep_idx, int32_t *keep_count, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* YS_B200_H */
