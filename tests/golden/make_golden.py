"""Generate golden fixtures by running the UNMODIFIED reference (imported from
/root/reference/yolo_scratch_cuda) on seeded synthetic inputs.

Run in the build container only (the reference does not travel):
    python tests/golden/make_golden.py
Writes tests/golden/*.npz.  Small cases store inputs and outputs; C1-sized cases
(B=2, 8400 anchors) store the generator arguments, an input checksum and the
outputs (integer outputs in full, float tensors as checksums + strided samples).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/yolo_scratch_cuda")

from losses.yolo_v8_loss import (TaskAlignedAssigner, v8DetectionLoss, bbox_iou,  # noqa: E402  (reference)
                                 bbox2dist, make_anchors, dist2bbox)
from train_yolo11_cuda import (nms_simple, decode_predictions_for_metrics,       # noqa: E402  (reference)
                               calculate_iou_batch_simple)
from utils.metrics import calculate_iou_batch                                    # noqa: E402  (reference)
from models.yolo11_modules import Detect as RefDetect                            # noqa: E402  (reference)

from yolo_scratch_b200 import synth                                              # noqa: E402

torch.set_num_threads(8)


class Detect(torch.nn.Module):
    """Stand-in carrying the three attributes v8DetectionLoss.__init__ reads (:348-356)."""

    def __init__(self, nc, reg_max, strides):
        super().__init__()
        self.nc, self.reg_max = nc, reg_max
        self.stride = torch.tensor([float(s) for s in strides])
        self.p = torch.nn.Parameter(torch.zeros(1))


def tensor_digest(t: torch.Tensor, n: int = 4096):
    f = t.detach().double().flatten()
    step = max(1, f.numel() // n)
    return {"sum": f.sum().item(), "abssum": f.abs().sum().item(), "sqsum": (f * f).sum().item(),
            "absmax": f.abs().max().item() if f.numel() else 0.0,
            "step": step, "samples": t.detach().flatten()[::step].float().numpy().copy()}


def save(name, **kw):
    flat = {}
    for k, v in kw.items():
        if isinstance(v, dict):
            for kk, vv in v.items():
                flat[f"{k}.{kk}"] = np.asarray(vv)
        elif isinstance(v, torch.Tensor):
            flat[k] = v.detach().numpy()
        else:
            flat[k] = np.asarray(v)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **flat)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


# ------------------------------------------------------------------ assigner
ASSIGN_SMALL = [
    # name, batch, shapes, nc, max_gt, seed, extra gt kwargs
    ("assign_small_s10", 3, ((16, 16), (8, 8), (4, 4)), 8, 6, 10, {}),
    ("assign_small_s11_forced", 4, ((16, 16), (8, 8), (4, 4)), 8, 8, 11, {"p_tiny": 0.4, "p_nested": 0.3}),
    ("assign_small_s12_g1", 2, ((16, 16), (8, 8), (4, 4)), 5, 1, 12, {"p_empty": 0.0}),
    ("assign_small_s13_empty_img", 4, ((16, 16), (8, 8), (4, 4)), 8, 5, 13, {"p_empty": 0.5}),
    ("assign_small_s14_dense", 2, ((20, 20), (10, 10), (5, 5)), 4, 40, 14, {"p_nested": 0.3, "p_tiny": 0.2}),
]


def run_ref_assigner(inp, nc, alpha=0.5, beta=4.0):
    a = TaskAlignedAssigner(topk=50, num_classes=nc, alpha=alpha, beta=beta)
    ps, pb, anc, gl, gb, mg = (t.clone() for t in inp)
    return a.forward(ps, pb, anc, gl, gb, mg)


def golden_assigner():
    for name, B, shapes, nc, max_gt, seed, kw in ASSIGN_SMALL:
        inp = synth.make_assigner_inputs(B, shapes, synth.LEVEL_STRIDES, nc, max_gt, seed, **kw)
        tl, tb, ts, fg, tgi = run_ref_assigner(inp, nc)
        save(name, pd_scores=inp[0], pd_bboxes=inp[1], anc_points=inp[2], gt_labels=inp[3], gt_bboxes=inp[4],
             mask_gt=inp[5], nc=nc, target_labels=tl, target_bboxes=tb, target_scores=ts, fg_mask=fg,
             target_gt_idx=tgi.to(torch.int32))
    # duplicate GT boxes -> argmax ties resolve to the lowest index
    inp = list(synth.make_assigner_inputs(2, ((16, 16), (8, 8), (4, 4)), synth.LEVEL_STRIDES, 8, 4, 15, p_empty=0.0))
    inp[4][:, 1] = inp[4][:, 0]
    inp[5][:, :2] = 1.0
    tl, tb, ts, fg, tgi = run_ref_assigner(inp, 8)
    save("assign_small_s15_dupgt", pd_scores=inp[0], pd_bboxes=inp[1], anc_points=inp[2], gt_labels=inp[3],
         gt_bboxes=inp[4], mask_gt=inp[5], nc=8, target_labels=tl, target_bboxes=tb, target_scores=ts, fg_mask=fg,
         target_gt_idx=tgi.to(torch.int32))
    # default alpha/beta of the class (1.0 / 6.0, :70)
    inp = synth.make_assigner_inputs(2, ((16, 16), (8, 8), (4, 4)), synth.LEVEL_STRIDES, 8, 6, 16)
    tl, tb, ts, fg, tgi = run_ref_assigner(inp, 8, alpha=1.0, beta=6.0)
    save("assign_small_s16_a1b6", pd_scores=inp[0], pd_bboxes=inp[1], anc_points=inp[2], gt_labels=inp[3],
         gt_bboxes=inp[4], mask_gt=inp[5], nc=8, alpha=1.0, beta=6.0, target_labels=tl, target_bboxes=tb,
         target_scores=ts, fg_mask=fg, target_gt_idx=tgi.to(torch.int32))
    # C1-sized: B=2, A=8400, nc=80 -- outputs only
    for seed in (20, 21):
        kw = {"p_tiny": 0.2, "p_nested": 0.2} if seed == 21 else {}
        inp = synth.make_assigner_inputs(2, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, 20, seed, **kw)
        tl, tb, ts, fg, tgi = run_ref_assigner(inp, 80)
        nz = ts.nonzero()
        save(f"assign_c1_s{seed}", seed=seed, gt_kw_keys=list(kw.keys()), gt_kw_vals=list(kw.values()),
             in_digest=tensor_digest(inp[1], 64), target_labels=tl.to(torch.int16), fg_mask=fg,
             target_gt_idx=tgi.to(torch.int16), ts_nz_idx=nz.to(torch.int32), ts_nz_val=ts[ts != 0],
             tb_digest=tensor_digest(tb))


# ------------------------------------------------------------------ loss
def run_ref_loss(feats, batch, nc, reg_max, strides):
    crit = v8DetectionLoss(Detect(nc, reg_max, strides))
    leaves = [f.clone().requires_grad_(True) for f in feats]
    total, items = crit(leaves, {k: v.clone() for k, v in batch.items()})
    total.backward()
    return total.detach(), items, [l.grad for l in leaves]


def golden_loss():
    small = [
        ("loss_small_s30", 3, 128, 8, 6, 30, {}),
        ("loss_small_s31_forced", 4, 128, 8, 8, 31, {"p_tiny": 0.4, "p_nested": 0.3}),
        ("loss_small_s32_empty_img", 4, 128, 6, 5, 32, {"p_empty": 0.5}),
    ]
    for name, B, imgsz, nc, max_gt, seed, kw in small:
        feats = synth.make_head_levels(B, imgsz, nc, 16, seed=seed)
        batch = synth.make_gt_batch(B, nc, max_gt, seed=seed + 1, imgsz=imgsz, **kw)
        total, items, grads = run_ref_loss(feats, batch, nc, 16, synth.LEVEL_STRIDES)
        save(name, nc=nc, imgsz=imgsz, f0=feats[0], f1=feats[1], f2=feats[2], batch_idx=batch["batch_idx"],
             cls=batch["cls"], bboxes=batch["bboxes"], total=total, items=items, g0=grads[0], g1=grads[1], g2=grads[2])
    # fully empty batch: cls loss only (:503-504, :100-108, :489-492)
    feats = synth.make_head_levels(2, 128, 8, 16, seed=33)
    batch = {"batch_idx": torch.zeros((0,), dtype=torch.long), "cls": torch.zeros((0, 1), dtype=torch.long),
             "bboxes": torch.zeros((0, 4))}
    total, items, grads = run_ref_loss(feats, batch, 8, 16, synth.LEVEL_STRIDES)
    save("loss_small_s33_empty_batch", nc=8, imgsz=128, f0=feats[0], f1=feats[1], f2=feats[2],
         batch_idx=batch["batch_idx"], cls=batch["cls"], bboxes=batch["bboxes"], total=total, items=items,
         g0=grads[0], g1=grads[1], g2=grads[2])
    # C1: B=2, 640 px, nc=80, <=20 GT/img  (BASELINE.json configs[0]) -- outputs only
    for seed in (40, 41):
        kw = {"p_empty": 0.0}
        feats = synth.make_head_levels(2, 640, 80, 16, seed=seed)
        batch = synth.make_gt_batch(2, 80, 20, seed=seed + 1, **kw)
        total, items, grads = run_ref_loss(feats, batch, 80, 16, synth.LEVEL_STRIDES)
        save(f"loss_c1_s{seed}", seed=seed, in_digest=tensor_digest(feats[0], 64), n_gt=batch["cls"].shape[0],
             total=total, items=items, g0=tensor_digest(grads[0]), g1=tensor_digest(grads[1]),
             g2=tensor_digest(grads[2]))


# ------------------------------------------------------------------ IoU helpers, anchors, decode
def golden_misc():
    g = torch.Generator().manual_seed(50)
    xy = torch.rand(64, 2, generator=g) * 100
    wh = torch.rand(64, 2, generator=g) * 40 + 1
    b1 = torch.cat((xy, xy + wh), 1)
    xy2 = xy + torch.randn(64, 2, generator=g) * 10
    wh2 = wh * (0.5 + torch.rand(64, 2, generator=g))
    b2 = torch.cat((xy2, xy2 + wh2), 1)
    out = {"b1": b1, "b2": b2}
    out["iou_xyxy"] = bbox_iou(b1, b2, xywh=False)
    out["ciou_xyxy"] = bbox_iou(b1, b2, xywh=False, CIoU=True)
    out["diou_xyxy"] = bbox_iou(b1, b2, xywh=False, DIoU=True)
    out["giou_xyxy"] = bbox_iou(b1, b2, xywh=False, GIoU=True)
    c1 = torch.cat(((b1[:, :2] + b1[:, 2:]) / 2, b1[:, 2:] - b1[:, :2]), 1)
    c2 = torch.cat(((b2[:, :2] + b2[:, 2:]) / 2, b2[:, 2:] - b2[:, :2]), 1)
    out["c1"], out["c2"] = c1, c2
    out["iou_xywh"] = bbox_iou(c1, c2, xywh=True)
    out["ciou_xywh"] = bbox_iou(c1, c2, xywh=True, CIoU=True)
    out["pair_simple"] = calculate_iou_batch_simple(b1[:1], b2)
    out["pair_matrix"] = calculate_iou_batch(b1[:16], b2)
    anc = torch.rand(64, 2, generator=g) * 20
    out["anc"] = anc
    out["bbox2dist"] = bbox2dist(anc, b1 / 8.0, 15)
    d = torch.rand(2, 64, 4, generator=g) * 5
    out["dist"] = d
    out["dist2bbox_xyxy"] = dist2bbox(d, anc.unsqueeze(0), xywh=False, dim=-1)
    out["dist2bbox_xywh"] = dist2bbox(d, anc.unsqueeze(0), xywh=True, dim=-1)
    feats = [torch.zeros(1, 1, 5, 7), torch.zeros(1, 1, 3, 4)]
    ap, st = make_anchors(feats, [8.0, 16.0], 0.5)
    out["anchors_5x7_3x4"], out["strides_5x7_3x4"] = ap, st
    save("misc_iou_anchor", **out)

    # Detect.inference (a16) through the reference module with its real DFL conv
    det = RefDetect(nc=8, ch=(16, 16, 16))
    det.stride = torch.tensor([8.0, 16.0, 32.0])
    det.eval()
    feats = synth.make_head_levels(2, 128, 8, 16, seed=51)
    y, _ = det.inference([f.clone() for f in feats])
    save("detect_inference_s51", nc=8, imgsz=128, f0=feats[0], f1=feats[1], f2=feats[2], y=y)


# ------------------------------------------------------------------ NMS
def golden_nms():
    cases = [("nms_s60_uniform", 2, 600, 8, 60, False), ("nms_s61_clustered", 2, 1500, 8, 61, True),
             ("nms_s62_c4shape", 1, 8400, 80, 62, True)]
    for name, B, A, nc, seed, clustered in cases:
        pred = synth.make_nms_predictions(B, A, nc, 640, seed=seed, clustered=clustered)
        for conf, thr in ((0.001, 0.7), (0.25, 0.45)):
            res = decode_predictions_for_metrics(pred.clone(), 640, conf, thr, torch.device("cpu"))
            keep_rows = []
            for b in range(B):
                best, cid = pred[b, :, 4:].max(1)
                m = best > conf
                x, y, w, h = pred[b, :, :4][m].unbind(1)
                xyxy = torch.stack((x - w / 2, y - h / 2, x + w / 2, y + h / 2), 1)
                k = nms_simple(xyxy, best[m], thr)
                keep_rows.append(m.nonzero().squeeze(1)[k] if len(k) else torch.zeros(0, dtype=torch.long))
            tag = f"{name}_conf{conf}_iou{thr}"
            kw = {"seed": seed, "B": B, "A": A, "nc": nc, "clustered": clustered, "conf": conf, "iou": thr,
                  "in_digest": tensor_digest(pred, 64)}
            if A <= 1500:
                kw["pred"] = pred
            for b in range(B):
                kw[f"keep{b}"] = keep_rows[b].to(torch.int32)
                kw[f"boxes{b}"] = res[b]["boxes"]
                kw[f"scores{b}"] = res[b]["scores"]
                kw[f"labels{b}"] = res[b]["labels"].to(torch.int32)
            save(tag, **kw)


if __name__ == "__main__":
    which = sys.argv[1:] or ["assign", "loss", "misc", "nms"]
    if "assign" in which:
        golden_assigner()
    if "loss" in which:
        golden_loss()
    if "misc" in which:
        golden_misc()
    if "nms" in which:
        golden_nms()
