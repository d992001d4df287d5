"""CPU: the oracle port (oracle/ref_port.py) replayed against fixtures produced by the
unmodified reference (tests/golden/make_golden.py).  Integer outputs bit-exact,
float outputs to 1e-6 (same ATen ops, possibly different op order)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, golden_names, t, digest_close
from oracle import ref_port as orc
from yolo_scratch_b200 import synth


def _assign_inputs_from_seed(g, seed):
    kw = dict(zip([str(k) for k in g["gt_kw_keys"]], [float(v) for v in g["gt_kw_vals"]]))
    inp = synth.make_assigner_inputs(2, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, 20, seed, **kw)
    digest_close(inp[1], g, "in_digest", 0.0)
    return inp


@pytest.mark.parametrize("name", golden_names("assign_small"))
def test_assigner_small(name):
    g = load_golden(name)
    alpha = float(g["alpha"]) if "alpha" in g else 0.5
    beta = float(g["beta"]) if "beta" in g else 4.0
    tl, tb, ts, fg, tgi = orc.assign(t(g["pd_scores"]), t(g["pd_bboxes"]), t(g["anc_points"]), t(g["gt_labels"]),
                                     t(g["gt_bboxes"]), t(g["mask_gt"]), int(g["nc"]), alpha, beta)
    assert torch.equal(fg, t(g["fg_mask"]))
    assert torch.equal(tgi.int(), t(g["target_gt_idx"]))
    assert torch.equal(tl, t(g["target_labels"]))
    assert torch.equal(tb, t(g["target_bboxes"]))
    assert torch.allclose(ts, t(g["target_scores"]), rtol=1e-6, atol=1e-9)
    assert fg.any()


@pytest.mark.parametrize("name", golden_names("assign_c1"))
def test_assigner_c1(name):
    g = load_golden(name)
    inp = _assign_inputs_from_seed(g, int(g["seed"]))
    tl, tb, ts, fg, tgi = orc.assign(*inp, 80)
    assert torch.equal(fg, t(g["fg_mask"]))
    assert torch.equal(tgi.to(torch.int16), t(g["target_gt_idx"]))
    assert torch.equal(tl.to(torch.int16), t(g["target_labels"]))
    nz = ts.nonzero().int()
    assert torch.equal(nz, t(g["ts_nz_idx"]))
    assert torch.allclose(ts[ts != 0], t(g["ts_nz_val"]), rtol=1e-6, atol=0)
    digest_close(tb, g, "tb_digest", 0.0)


def _loss_inputs(g):
    feats = [t(g["f0"]), t(g["f1"]), t(g["f2"])]
    batch = {"batch_idx": t(g["batch_idx"]), "cls": t(g["cls"]), "bboxes": t(g["bboxes"])}
    return feats, batch


@pytest.mark.parametrize("name", golden_names("loss_small"))
def test_loss_small(name):
    g = load_golden(name)
    feats, batch = _loss_inputs(g)
    leaves = [f.clone().requires_grad_(True) for f in feats]
    total, items = orc.detection_loss(leaves, batch, int(g["nc"]), 16, synth.LEVEL_STRIDES)
    total.backward()
    assert torch.allclose(items, t(g["items"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(total.detach(), t(g["total"]), rtol=1e-6)
    for leaf, key in zip(leaves, ("g0", "g1", "g2")):
        ref = t(g[key])
        scale = ref.abs().max().item() or 1.0
        assert (leaf.grad - ref).abs().max().item() <= 1e-6 * scale


@pytest.mark.parametrize("name", golden_names("loss_c1"))
def test_loss_c1(name):
    g = load_golden(name)
    seed = int(g["seed"])
    feats = synth.make_head_levels(2, 640, 80, 16, seed=seed)
    digest_close(feats[0], g, "in_digest", 0.0)
    batch = synth.make_gt_batch(2, 80, 20, seed=seed + 1, p_empty=0.0)
    assert batch["cls"].shape[0] == int(g["n_gt"])
    leaves = [f.clone().requires_grad_(True) for f in feats]
    total, items = orc.detection_loss(leaves, batch, 80, 16, synth.LEVEL_STRIDES)
    total.backward()
    assert torch.allclose(items, t(g["items"]), rtol=1e-6)
    assert torch.allclose(total.detach(), t(g["total"]), rtol=1e-6)
    for leaf, key in zip(leaves, ("g0", "g1", "g2")):
        digest_close(leaf.grad, g, key, 1e-6)


def test_iou_helpers_and_anchors():
    g = load_golden("misc_iou_anchor")
    b1, b2, c1, c2 = t(g["b1"]), t(g["b2"]), t(g["c1"]), t(g["c2"])
    assert torch.equal(orc.pair_iou_xyxy(b1, b2), t(g["iou_xyxy"]))
    assert torch.allclose(orc.pair_iou_xyxy(b1, b2, ciou=True), t(g["ciou_xyxy"]), rtol=1e-6, atol=1e-7)
    assert torch.equal(orc.iou_general(b1, b2, xywh=False), t(g["iou_xyxy"]))
    assert torch.allclose(orc.iou_general(b1, b2, xywh=False, DIoU=True), t(g["diou_xyxy"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(orc.iou_general(b1, b2, xywh=False, GIoU=True), t(g["giou_xyxy"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(orc.iou_general(c1, c2, xywh=True), t(g["iou_xywh"]), rtol=1e-6, atol=1e-7)
    assert torch.allclose(orc.iou_general(c1, c2, xywh=True, CIoU=True), t(g["ciou_xywh"]), rtol=1e-6, atol=1e-7)
    assert torch.equal(orc.iou_matrix(b1[:1], b2).squeeze(0), t(g["pair_simple"]))
    assert torch.equal(orc.iou_matrix(b1[:16], b2), t(g["pair_matrix"]))
    ap, st = orc.anchors_for([(5, 7), (3, 4)], [8.0, 16.0])
    assert torch.equal(ap, t(g["anchors_5x7_3x4"])) and torch.equal(st, t(g["strides_5x7_3x4"]))
    d, anc = t(g["dist"]), t(g["anc"])
    assert torch.equal(orc.ltrb_to_xyxy(d, anc.unsqueeze(0)), t(g["dist2bbox_xyxy"]))


def test_detect_inference():
    g = load_golden("detect_inference_s51")
    feats = [t(g["f0"]), t(g["f1"]), t(g["f2"])]
    y = orc.detect_inference(feats, int(g["nc"]), 16, synth.LEVEL_STRIDES)
    assert y.shape == t(g["y"]).shape
    assert torch.allclose(y, t(g["y"]), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("name", golden_names("nms_"))
def test_nms(name):
    g = load_golden(name)
    B, A, nc = int(g["B"]), int(g["A"]), int(g["nc"])
    if "pred" in g:
        pred = t(g["pred"])
    else:
        pred = synth.make_nms_predictions(B, A, nc, 640, seed=int(g["seed"]), clustered=bool(g["clustered"]))
        digest_close(pred, g, "in_digest", 0.0)
    if A > 2000 and float(g["conf"]) < 0.01:
        pytest.skip("full-size greedy NMS through the torch port is covered by the C oracle test")
    res = orc.decode_for_metrics(pred, 640, float(g["conf"]), float(g["iou"]), return_keep=True)
    for b in range(B):
        assert torch.equal(res[b]["keep"].int(), t(g[f"keep{b}"]))
        assert torch.equal(res[b]["boxes"], t(g[f"boxes{b}"]))
        assert torch.equal(res[b]["scores"], t(g[f"scores{b}"]))
        assert torch.equal(res[b]["labels"].int(), t(g[f"labels{b}"]))
