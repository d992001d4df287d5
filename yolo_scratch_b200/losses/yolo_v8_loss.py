"""Drop-in for the reference's `losses/yolo_v8_loss.py` (same names, same call signatures, same
return types), backed by the sm_100a kernels in libys_b200.so.

Reference interface mirrored here (paths relative to /root/reference/yolo_scratch_cuda/):
  v8DetectionLoss            losses/yolo_v8_loss.py:333-538
  TaskAlignedAssigner        losses/yolo_v8_loss.py:64-270
  BboxLoss                   losses/yolo_v8_loss.py:273-324
  bbox_iou / bbox2dist / make_anchors / dist2bbox   :12-61 / :327-330 / :541-552 / :555-564

`v8DetectionLoss.__call__` runs ONE fused C-ABI call (ys_detection_loss) that does target packing,
decode, assignment, BCE + CIoU + DFL forward and the full backward, reading the head tensors in their
native NCHW layout.  There is no CPU or eager-PyTorch fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import math

import torch
import torch.nn as nn

from .. import _lib
from .._lib import YsError, check, ptr, stream_ptr


# ----------------------------------------------------------------------------------------------
# small tensor helpers of the reference API (layout only; not on the hot path)
# ----------------------------------------------------------------------------------------------
def make_anchors(feats, strides, grid_cell_offset=0.5):
    """Anchor centres in grid units, level-major then row-major, and the stride column
    (reference losses/yolo_v8_loss.py:541-552)."""
    pts, sts = [], []
    dtype, device = feats[0].dtype, feats[0].device
    for i, stride in enumerate(strides):
        h, w = feats[i].shape[2:]
        ys_, xs_ = torch.meshgrid(torch.arange(h, device=device, dtype=dtype) + grid_cell_offset,
                                  torch.arange(w, device=device, dtype=dtype) + grid_cell_offset, indexing="ij")
        pts.append(torch.stack((xs_, ys_), -1).view(-1, 2))
        sts.append(torch.full((h * w, 1), float(stride), dtype=dtype, device=device))
    return torch.cat(pts), torch.cat(sts)


def dist2bbox(distance, anchor_points, xywh=True, dim=-1):
    """ltrb distances -> xywh or xyxy boxes (reference losses/yolo_v8_loss.py:555-564)."""
    lt, rb = distance.chunk(2, dim)
    x1y1, x2y2 = anchor_points - lt, anchor_points + rb
    if xywh:
        return torch.cat(((x1y1 + x2y2) / 2, x2y2 - x1y1), dim)
    return torch.cat((x1y1, x2y2), dim)


def bbox2dist(anchor_points, bbox, reg_max):
    """xyxy boxes -> clamped ltrb targets (reference losses/yolo_v8_loss.py:327-330)."""
    x1y1, x2y2 = bbox.chunk(2, -1)
    return torch.cat((anchor_points - x1y1, x2y2 - anchor_points), -1).clamp_(0, reg_max - 0.01)


class _BboxIoU(torch.autograd.Function):
    """Paired-box IoU family through ys_bbox_iou; gradient wrt box1 for the CIoU/xyxy form BboxLoss uses."""

    @staticmethod
    def forward(ctx, box1, box2, xywh, mode, eps, want_grad):
        lib = _lib.lib()
        n = box1.shape[0]
        out = torch.empty(n, device=box1.device, dtype=torch.float32)
        grad = torch.empty_like(box1) if want_grad else None
        with _lib.on_device(box1, box2):
            check(lib.ys_bbox_iou(ptr(box1), ptr(box2), n, int(bool(xywh)), mode, float(eps), ptr(out), 0, ptr(grad),
                                  stream_ptr(box1.device)), "ys_bbox_iou")
        ctx.save_for_backward(grad) if grad is not None else None
        ctx.has_grad = grad is not None
        return out

    @staticmethod
    def backward(ctx, gout):
        if not ctx.has_grad:
            return None, None, None, None, None, None
        (grad,) = ctx.saved_tensors
        return grad * gout.unsqueeze(-1), None, None, None, None, None


def bbox_iou(box1, box2, xywh=True, GIoU=False, DIoU=False, CIoU=False, eps=1e-7):
    """IoU / GIoU / DIoU / CIoU of broadcastable boxes (..., 4) -> (..., 1)
    (reference losses/yolo_v8_loss.py:12-61)."""
    _lib.require_cuda(box1, "box1")
    _lib.require_cuda(box2, "box2")
    if torch.is_grad_enabled():
        # The device kernel differentiates the BboxLoss form only: wrt box1, xyxy, CIoU, eps 1e-7 (reference :298).
        # Anything else would silently drop gradients the reference's autograd provides, so it raises here.
        if box2.requires_grad:
            raise YsError("bbox_iou: gradient wrt box2 is not implemented (pass box2.detach(); targets carry no grad "
                          "in the reference's loss, losses/yolo_v8_loss.py:416-417)")
        if box1.requires_grad and not ((not xywh) and CIoU and abs(eps - 1e-7) < 1e-12):
            raise YsError("bbox_iou: gradient wrt box1 is implemented for xywh=False, CIoU=True, eps=1e-7 only "
                          "(the BboxLoss use); detach box1 or use that form")
    shape = torch.broadcast_shapes(box1.shape[:-1], box2.shape[:-1])
    b1 = box1.expand(*shape, 4).reshape(-1, 4).contiguous().float()
    b2 = box2.expand(*shape, 4).reshape(-1, 4).contiguous().float()
    mode = 3 if CIoU else 2 if DIoU else 1 if GIoU else 0
    out = _BboxIoU.apply(b1, b2, bool(xywh), mode, float(eps), torch.is_grad_enabled() and box1.requires_grad)
    return out.view(*shape, 1)


# ----------------------------------------------------------------------------------------------
# TaskAlignedAssigner
# ----------------------------------------------------------------------------------------------
class TaskAlignedAssigner(nn.Module):
    """Same constructor and forward contract as the reference class (losses/yolo_v8_loss.py:64-180).
    `topk` is accepted and, as in the reference, never used -- unless one of the two keyword-only,
    default-off modes north_star names is switched on:
      use_topk=True     per GT, only the `topk` anchors with the largest alignment metric among the anchors inside it
                        stay positive before the forced-assignment loops (what `topk` would mean if the reference read
                        it; semantics defined by oracle/ref_port.assign(topk=...), parity unpinned by the reference);
      ciou_assign=True  the pair metric is the reference's own bbox_iou(CIoU=True) clamped at 0 instead of plain IoU."""

    def __init__(self, topk=13, num_classes=80, alpha=1.0, beta=6.0, eps=1e-9, *, use_topk=False, ciou_assign=False):
        super().__init__()
        self.topk = topk
        self.num_classes = num_classes
        self.alpha = alpha
        self.beta = beta
        self.eps = eps
        self.use_topk = bool(use_topk)
        self.ciou_assign = bool(ciou_assign)
        self._ws = _lib.Workspace()

    @torch.no_grad()
    def forward(self, pd_scores, pd_bboxes, anc_points, gt_labels, gt_bboxes, mask_gt):
        _lib.require_cuda(pd_scores, "pd_scores")
        self.bs = pd_scores.shape[0]
        self.n_max_boxes = gt_bboxes.shape[1]
        B, A, nc = pd_scores.shape
        G = self.n_max_boxes
        if G == 0:   # reference :100-108 (float fg_mask and a float tensor in place of target_gt_idx)
            z = torch.zeros_like(pd_scores[..., 0])
            return (torch.full_like(z, self.num_classes).long(), torch.zeros_like(pd_bboxes),
                    torch.zeros_like(pd_scores), z, z.clone())
        dev = pd_scores.device
        f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        ps, pb, an, gl, gb, mg = f(pd_scores), f(pd_bboxes), f(anc_points), f(gt_labels), f(gt_bboxes), f(mask_gt)
        lib = _lib.lib()
        tl = torch.empty(B, A, device=dev)
        tb = torch.empty(B, A, 4, device=dev)
        ts = torch.empty(B, A, self.num_classes, device=dev)
        fg = torch.empty(B, A, device=dev, dtype=torch.uint8)
        tgi = torch.empty(B, A, device=dev, dtype=torch.int64)
        topk = int(self.topk) if self.use_topk else 0
        ws = self._ws.get(lib.ys_assign_ex_workspace_bytes(B, A, G, topk), dev)
        with _lib.on_device(ps):
            check(lib.ys_assign_ex(B, A, G, self.num_classes, ptr(ps), ptr(pb), ptr(an), ptr(gl), ptr(gb), ptr(mg),
                                   float(self.alpha), float(self.beta), float(self.eps), topk, int(self.ciou_assign),
                                   ptr(tl), ptr(tb), ptr(ts), ptr(fg), ptr(tgi), 0, ptr(ws), ws.numel(), stream_ptr(dev)),
                  "ys_assign_ex")
        return tl, tb, ts, fg.bool(), tgi


# ----------------------------------------------------------------------------------------------
# BboxLoss (stand-alone use; the fused loss below does not go through it)
# ----------------------------------------------------------------------------------------------
class _BboxLossFn(torch.autograd.Function):
    """ys_bbox_loss: both loss terms and their gradients wrt pred_dist / pred_bboxes in one launch."""

    @staticmethod
    def forward(ctx, pred_dist, pred_bboxes, anchor_points, target_bboxes, target_scores, tss, fg_mask, reg_max, want_grad):
        lib = _lib.lib()
        dev = pred_dist.device
        A = anchor_points.shape[0]
        nc = target_scores.shape[-1]
        n = fg_mask.numel()
        f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        pd, pb, ap, tb, ts = f(pred_dist), f(pred_bboxes), f(anchor_points), f(target_bboxes), f(target_scores)
        fg = fg_mask.detach().to(device=dev, dtype=torch.uint8).contiguous()
        tss_d = f(tss).reshape(1)
        out = torch.empty(2, device=dev)
        g_pd = torch.empty_like(pd) if (want_grad and pred_dist.requires_grad) else None
        g_pb = torch.empty_like(pb) if (want_grad and pred_bboxes.requires_grad) else None
        ws = torch.empty(lib.ys_bbox_loss_workspace_bytes(n), dtype=torch.uint8, device=dev)
        with _lib.on_device(pd):
            check(lib.ys_bbox_loss(ptr(pd), ptr(pb), ptr(ap), ptr(tb), ptr(ts), ptr(fg), ptr(tss_d), n, A, nc, int(reg_max),
                                   ptr(out), ptr(g_pb), ptr(g_pd), ptr(ws), ws.numel(), stream_ptr(dev)), "ys_bbox_loss")
        ctx.g_pb = g_pb.view_as(pred_bboxes) if g_pb is not None else None
        ctx.g_pd = g_pd.view_as(pred_dist) if g_pd is not None else None
        return out[0], out[1]

    @staticmethod
    def backward(ctx, g_iou, g_dfl):
        g_pb = ctx.g_pb * g_iou if ctx.g_pb is not None else None
        g_pd = ctx.g_pd * g_dfl if ctx.g_pd is not None else None
        return g_pd, g_pb, None, None, None, None, None, None, None


class BboxLoss(nn.Module):
    """CIoU + DFL on already-assigned dense tensors: same constructor and forward contract as the reference class
    (losses/yolo_v8_loss.py:273-324), computed by one kernel (ys_bbox_loss) instead of boolean-mask gathers.
    Gradients flow to pred_dist and pred_bboxes; the targets carry none (they come out of a no_grad assigner,
    reference :78, :416-417)."""

    def __init__(self, reg_max=16):
        super().__init__()
        self.reg_max = reg_max

    def forward(self, pred_dist, pred_bboxes, anchor_points, target_bboxes, target_scores, target_scores_sum, fg_mask):
        _lib.require_cuda(pred_dist, "pred_dist")
        if torch.is_grad_enabled() and (target_bboxes.requires_grad or target_scores.requires_grad):
            raise YsError("BboxLoss: targets must not require grad (the assigner's outputs are detached in the reference)")
        tss = target_scores_sum if torch.is_tensor(target_scores_sum) else torch.tensor(float(target_scores_sum))
        return _BboxLossFn.apply(pred_dist, pred_bboxes, anchor_points, target_bboxes, target_scores, tss, fg_mask,
                                 self.reg_max, torch.is_grad_enabled())


# ----------------------------------------------------------------------------------------------
# fused detection loss
# ----------------------------------------------------------------------------------------------
class _FusedDetectionLoss(torch.autograd.Function):
    """forward = ys_detection_loss (forward AND backward kernels); backward = grad * upstream scalar."""

    @staticmethod
    def forward(ctx, owner, batch_idx, cls, bboxes, export, *feats):
        lib = _lib.lib()
        dev = feats[0].device
        n = len(feats)
        B = feats[0].shape[0]
        want_grad = any(f.requires_grad for f in feats)
        xs = [f.detach() for f in feats]
        grads = [torch.empty_like(x) for x in xs] if want_grad else None
        hs = (C.c_int * n)(*[x.shape[2] for x in xs])
        ws_ = (C.c_int * n)(*[x.shape[3] for x in xs])
        st = (C.c_float * n)(*[float(s) for s in owner.stride[:n]])
        xp = (C.c_void_p * n)(*[x.data_ptr() for x in xs])
        gp = (C.c_void_p * n)(*[g.data_ptr() for g in grads]) if want_grad else None
        A = sum(x.shape[2] * x.shape[3] for x in xs)
        n_gt = int(batch_idx.shape[0])
        out = torch.empty(4, device=dev, dtype=torch.float32)
        ws = owner._ws.get(lib.ys_detection_loss_workspace_bytes(B, A, n_gt, owner.nc), dev)
        pb = tgi = nrm = None
        if export:
            pb = torch.empty(B, A, 4, device=dev)
            tgi = torch.empty(B, A, device=dev, dtype=torch.int32)
            nrm = torch.empty(B, A, device=dev)
        # imgsz = feats[0].shape[2:] * stride[0]; x is scaled by imgsz[0], y by imgsz[1] (reference :400, :515)
        sx = float(xs[0].shape[2]) * float(owner.stride[0])
        sy = float(xs[0].shape[3]) * float(owner.stride[0])
        with _lib.on_device(*xs, batch_idx, cls, bboxes):
            check(lib.ys_detection_loss(n, xp, gp, hs, ws_, st, B, owner.nc, owner.reg_max, ptr(batch_idx), ptr(cls),
                                        ptr(bboxes), n_gt, sx, sy, owner.hyp_box, owner.hyp_cls, owner.hyp_dfl,
                                        float(owner.assigner.alpha), float(owner.assigner.beta), ptr(out), ptr(pb),
                                        ptr(tgi), ptr(nrm), 0, ptr(ws), ws.numel(), stream_ptr(dev)), "ys_detection_loss")
        if export:
            owner.last_export = {"pred_bboxes_px": pb, "target_gt_idx": tgi, "norm": nrm}
        ctx.grads = grads
        ctx.n_feats = n
        ctx.consumed = False
        total, items = out[0], out[1:4]
        ctx.mark_non_differentiable(items)
        return total, items

    @staticmethod
    def backward(ctx, g_total, g_items):
        if ctx.consumed:
            # the gradients were computed by the forward call and handed over (scaled in place) by the first backward;
            # like torch without retain_graph, a second pass needs a fresh forward
            raise YsError("v8DetectionLoss: backward through the same loss twice is not supported (the fused kernel's "
                          "gradient buffers are handed to autograd by the first backward); call the criterion again")
        grads = ctx.grads
        if grads is None:        # no head tensor required grad
            return (None,) * (5 + ctx.n_feats)
        lib = _lib.lib()
        g = g_total.detach().to(torch.float32).contiguous()
        with _lib.on_device(*grads, g):
            for t in grads:   # in place, skipped on the device when the upstream scalar is exactly 1
                check(lib.ys_scale_inplace(ptr(t), t.numel(), ptr(g), stream_ptr(t.device)), "ys_scale_inplace")
        ctx.grads = None
        ctx.consumed = True
        return (None, None, None, None, None) + tuple(grads)


class v8DetectionLoss:
    """Same constructor and call contract as the reference (losses/yolo_v8_loss.py:333-499):
    `criterion = v8DetectionLoss(model, tal_topk=10)`; `loss, loss_items = criterion(preds, batch)`."""

    def __init__(self, model, tal_topk=10):
        self.model = model
        self.tal_topk = tal_topk
        detect = None
        if hasattr(model, "model") and hasattr(model.model, "detect"):
            detect = model.model.detect
        elif hasattr(model, "detect"):
            detect = model.detect
        else:
            for m in model.modules():
                if type(m).__name__ == "Detect":
                    detect = m
                    break
        if detect is None:
            raise YsError("model has no Detect head")
        self.nc = detect.nc
        self.reg_max = detect.reg_max
        self.stride = detect.stride
        self.device = next(model.parameters()).device
        self.assigner = TaskAlignedAssigner(topk=50, num_classes=self.nc, alpha=0.5, beta=4.0)
        self.bbox_loss = BboxLoss(self.reg_max)
        self.epoch = 0
        self.hyp_box = 7.5
        self.hyp_cls = 0.5
        self.hyp_dfl = 1.5
        self.export_assignment = False      # tests: keep decoded boxes / assignment of the last call
        self.last_export = None
        self._ws = _lib.Workspace()

    def __call__(self, preds, batch):
        feats = preds
        if isinstance(feats, tuple):
            feats = feats[1] if len(feats) == 2 else feats
        feats = list(feats)
        for i, f in enumerate(feats):
            _lib.require_cuda(f, f"preds[{i}]")
            if f.dtype != torch.float32 or not f.is_contiguous():
                feats[i] = f.float().contiguous()
        dev = feats[0].device
        bi = batch["batch_idx"].to(device=dev, dtype=torch.int64, non_blocking=True).contiguous().view(-1)
        cl = batch["cls"].to(device=dev, dtype=torch.int64, non_blocking=True).contiguous().view(-1)
        bb = batch["bboxes"].to(device=dev, dtype=torch.float32, non_blocking=True).contiguous().view(-1, 4)
        total, items = _FusedDetectionLoss.apply(self, bi, cl, bb, self.export_assignment, *feats)
        return total, items
