"""Golden fixtures for the mAP matching row: the UNMODIFIED reference evaluate_detections / calculate_ap
(imported from /root/reference/yolo_scratch_cuda/utils/metrics.py) run on seeded synthetic predictions and targets.
Build container only:  python tests/golden/make_golden_metrics.py   -> tests/golden/evalmetrics_s*.npz"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/yolo_scratch_cuda")
from utils.metrics import evaluate_detections, calculate_ap      # noqa: E402  (reference)


def make_case(seed, B, max_gt, n_rand, jitter):
    g = torch.Generator().manual_seed(seed)
    preds, tgts = [], []
    for b in range(B):
        n = int(torch.randint(0, max_gt + 1, (), generator=g))
        c = torch.rand(n, 2, generator=g)
        wh = 0.03 + 0.3 * torch.rand(n, 2, generator=g)
        gt = torch.cat((c - wh / 2, c + wh / 2), 1).clamp(0, 1)
        # predictions: jittered copies of some GT (several per GT: duplicates must become FPs) + random boxes
        k = torch.randint(0, 3, (n,), generator=g)
        rep = gt.repeat_interleave(k, 0)
        rep = (rep + jitter * (torch.rand(rep.shape, generator=g) - 0.5)).clamp(0, 1)
        m = int(torch.randint(0, n_rand + 1, (), generator=g)) if b != 1 else 0
        rc = torch.rand(m, 2, generator=g)
        rwh = 0.03 + 0.3 * torch.rand(m, 2, generator=g)
        rnd = torch.cat((rc - rwh / 2, rc + rwh / 2), 1).clamp(0, 1)
        boxes = torch.cat((rep, rnd), 0) if b != 2 else torch.zeros(0, 4)
        P = boxes.shape[0]
        scores = (torch.randperm(4096, generator=g)[:P].float() + 1) / 4097.0          # unique
        preds.append({"boxes": boxes, "scores": scores, "labels": torch.zeros(P, dtype=torch.long)})
        tgts.append({"boxes": gt, "labels": torch.zeros(n, dtype=torch.long)})
    return preds, tgts


def pack(lst, key, width, fill=0.0):
    cols = 4 if key == "boxes" else 1
    out = torch.full((len(lst), width, cols), fill)
    for i, d in enumerate(lst):
        n = len(d[key])
        if n:
            out[i, :n] = d[key].float().view(n, cols)
    return out.numpy()


for seed, B, max_gt, n_rand, jitter, conf, iou in [(70, 6, 12, 20, 0.04, 0.25, 0.5), (71, 8, 30, 60, 0.08, 0.001, 0.5),
                                                   (72, 5, 6, 10, 0.02, 0.4, 0.6), (73, 3, 0, 8, 0.0, 0.1, 0.5)]:
    preds, tgts = make_case(seed, B, max_gt, n_rand, jitter)
    ref = evaluate_detections(preds, tgts, conf_threshold=conf, iou_threshold=iou)
    P = max(1, max(len(p["boxes"]) for p in preds))
    G = max(1, max(len(t["boxes"]) for t in tgts))
    np.savez_compressed(os.path.join(HERE, f"evalmetrics_s{seed}.npz"), seed=seed, conf=conf, iou=iou,
                        pred_boxes=pack(preds, "boxes", P), pred_scores=pack(preds, "scores", P)[..., 0],
                        pred_count=np.array([len(p["boxes"]) for p in preds]),
                        gt_boxes=pack(tgts, "boxes", G), gt_count=np.array([len(t["boxes"]) for t in tgts]),
                        precision=ref["precision"], recall=ref["recall"], mAP50=float(ref["mAP50"]),
                        mAP50_95=float(ref["mAP50-95"]))
    print(seed, ref)

# calculate_ap known answers
g = torch.Generator().manual_seed(7)
cases = []
for n_tp, n_fp, n_gt in [(5, 3, 9), (0, 4, 3), (6, 0, 6), (0, 0, 5), (3, 3, 0), (40, 55, 61)]:
    tp = ((torch.randperm(1000, generator=g)[:n_tp].float() + 1) / 1001).tolist()
    fp = ((torch.randperm(1000, generator=g)[:n_fp].float() + 1) / 1001).tolist()
    cases.append((tp, fp, n_gt, float(calculate_ap(tp, fp, n_gt))))
np.savez_compressed(os.path.join(HERE, "calculate_ap_known.npz"),
                    tp=np.array([np.array(c[0] + [-1.0] * (64 - len(c[0]))) for c in cases]),
                    fp=np.array([np.array(c[1] + [-1.0] * (64 - len(c[1]))) for c in cases]),
                    n_gt=np.array([c[2] for c in cases]), ap=np.array([c[3] for c in cases]))
print([c[3] for c in cases])

# ---- cda_nms: reference keep lists on clustered detections, with and without angles
from utils.metrics import cda_nms      # noqa: E402  (reference)

for seed, n, n_clusters, with_angles, max_det, thr in [(80, 200, 6, False, 10, 0.15), (81, 400, 25, True, 10, 0.15),
                                                       (82, 60, 3, False, 5, 0.3), (83, 500, 40, True, 20, 0.05)]:
    g = torch.Generator().manual_seed(seed)
    cc = 0.15 + 0.7 * torch.rand(n_clusters, 2, generator=g)
    cw = 0.04 + 0.12 * torch.rand(n_clusters, 2, generator=g)
    k = torch.randint(0, n_clusters, (n,), generator=g)
    c = cc[k] + 0.01 * torch.randn(n, 2, generator=g)
    wh = cw[k] * (1 + 0.1 * torch.randn(n, 2, generator=g)).clamp(0.5, 1.5)
    boxes = torch.cat((c - wh / 2, c + wh / 2), 1).clamp(0, 1)
    scores = (torch.randperm(8192, generator=g)[:n].float() + 1) / 8193.0
    pred = {"boxes": boxes, "scores": scores, "labels": k.clone()}
    if with_angles:
        pred["angles"] = (torch.rand(n, generator=g) - 0.5) * 3.0
    out = cda_nms(pred, max_detections=max_det, dga_thresh=thr)
    # identify the kept rows by their (unique) scores
    kept = [int((scores == s).nonzero()[0, 0]) for s in out["scores"]]
    np.savez_compressed(os.path.join(HERE, f"cda_nms_s{seed}.npz"), boxes=boxes.numpy(), scores=scores.numpy(),
                        labels=k.numpy(), angles=pred.get("angles", torch.zeros(0)).numpy(), with_angles=with_angles,
                        max_det=max_det, thr=thr, kept=np.array(kept), kept_boxes=out["boxes"].numpy())
    print("cda_nms", seed, kept)
