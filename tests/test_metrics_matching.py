"""mAP matching.  Fixtures come from the UNMODIFIED reference (tests/golden/make_golden_metrics.py):
calculate_ap known answers (CPU) and whole evaluate_detections result dicts (GPU, through ys_match_detections)."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_calculate_ap_known_answers():
    from yolo_scratch_b200.utils.metrics import calculate_ap
    g = np.load(os.path.join(GOLD, "calculate_ap_known.npz"))
    for tp, fp, n_gt, ap in zip(g["tp"], g["fp"], g["n_gt"], g["ap"]):
        got = calculate_ap([x for x in tp.tolist() if x >= 0], [x for x in fp.tolist() if x >= 0], int(n_gt))
        assert abs(float(got) - float(ap)) <= 1e-12, (got, ap)


def test_evaluate_detections_refuses_cpu_tensors():
    from yolo_scratch_b200 import _lib
    from yolo_scratch_b200.utils.metrics import evaluate_detections
    p = [{"boxes": torch.zeros(1, 4), "scores": torch.ones(1), "labels": torch.zeros(1, dtype=torch.long)}]
    t = [{"boxes": torch.zeros(1, 4), "labels": torch.zeros(1, dtype=torch.long)}]
    with pytest.raises(_lib.YsError):
        evaluate_detections(p, t)


def _load(seed):
    g = np.load(os.path.join(GOLD, f"evalmetrics_s{seed}.npz"))
    preds, tgts = [], []
    for b in range(len(g["pred_count"])):
        n, m = int(g["pred_count"][b]), int(g["gt_count"][b])
        preds.append({"boxes": torch.tensor(g["pred_boxes"][b, :n]).view(n, 4).cuda(),
                      "scores": torch.tensor(g["pred_scores"][b, :n]).cuda(),
                      "labels": torch.zeros(n, dtype=torch.long).cuda()})
        tgts.append({"boxes": torch.tensor(g["gt_boxes"][b, :m]).view(m, 4).cuda(), "labels": torch.zeros(m, dtype=torch.long).cuda()})
    return g, preds, tgts


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [70, 71, 72, 73])
def test_evaluate_detections_golden(cuda_lib, seed):
    from yolo_scratch_b200.utils.metrics import evaluate_detections
    g, preds, tgts = _load(seed)
    got = evaluate_detections(preds, tgts, conf_threshold=float(g["conf"]), iou_threshold=float(g["iou"]))
    assert got["precision"] == pytest.approx(float(g["precision"]), abs=1e-12)
    assert got["recall"] == pytest.approx(float(g["recall"]), abs=1e-12)
    assert float(got["mAP50"]) == pytest.approx(float(g["mAP50"]), abs=1e-12)
    assert float(got["mAP50-95"]) == pytest.approx(float(g["mAP50_95"]), abs=1e-12)


@pytest.mark.gpu
def test_matching_properties_after_nms(cuda_lib):
    """End of the validation path at full size: NMS output (score-descending) -> matching.  Every GT is consumed at most
    once per threshold, scores come back sorted, unsorted input gives the same flags as sorted input."""
    from yolo_scratch_b200 import synth
    from yolo_scratch_b200.train_yolo11_cuda import decode_predictions_for_metrics
    from yolo_scratch_b200.utils.metrics import match_detections
    pred = synth.make_nms_predictions(4, 8400, 80, 640, seed=91, clustered=True).cuda()
    dets = decode_predictions_for_metrics(pred, 640, 0.25, 0.45, pred.device)
    g = torch.Generator().manual_seed(5)
    tgts = []
    for d in dets:
        k = min(15, d["boxes"].shape[0])
        idx = torch.randperm(d["boxes"].shape[0], generator=g)[:k].to(d["boxes"].device)
        tgts.append({"boxes": (d["boxes"][idx] + 0.01 * torch.rand(k, 4, generator=g).to(d["boxes"].device)).clamp(0, 1),
                     "labels": torch.zeros(k, dtype=torch.long)})
    thr = [0.5, 0.6, 0.7, 0.8, 0.9]
    tp, scores, nv, n_gt = match_detections(dets, tgts, 0.25, thr)
    for b in range(len(dets)):
        per_thr = tp[b, :nv[b]].sum(0)
        assert (per_thr <= len(tgts[b]["boxes"])).all()
        assert (np.diff(scores[b, :nv[b]]) <= 0).all()
    shuffled = []
    for d in dets:
        perm = torch.randperm(d["boxes"].shape[0], generator=g).to(d["boxes"].device)
        shuffled.append({k: v[perm] for k, v in d.items()})
    tp2, scores2, nv2, _ = match_detections(shuffled, tgts, 0.25, thr)
    assert np.array_equal(nv, nv2) and np.array_equal(tp, tp2) and np.array_equal(scores, scores2)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [80, 81, 82, 83])
def test_cda_nms_golden(cuda_lib, seed):
    """keep list of the reference's cda_nms (dGA-distance greedy NMS, top-k) on clustered detections."""
    from yolo_scratch_b200.utils.metrics import cda_nms
    g = np.load(os.path.join(GOLD, f"cda_nms_s{seed}.npz"))
    pred = {"boxes": torch.tensor(g["boxes"]).cuda(), "scores": torch.tensor(g["scores"]).cuda(),
            "labels": torch.tensor(g["labels"]).cuda()}
    if bool(g["with_angles"]):
        pred["angles"] = torch.tensor(g["angles"]).cuda()
    out = cda_nms(pred, max_detections=int(g["max_det"]), dga_thresh=float(g["thr"]))
    kept = [int((pred["scores"] == s).nonzero()[0, 0]) for s in out["scores"]]
    assert kept == g["kept"].tolist()
    assert torch.equal(out["boxes"].cpu(), torch.tensor(g["kept_boxes"]))
    assert torch.equal(out["labels"].cpu(), torch.tensor(g["labels"])[g["kept"]])
    assert set(out.keys()) == {"boxes", "scores", "angles", "labels"}
