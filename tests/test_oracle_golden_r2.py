"""CPU: the oracles replayed against the round-2 fixtures (tests/golden/make_golden_r2.py, outputs of the unmodified
reference): stand-alone BboxLoss, the CIoU gradient, class-aware NMS as a composition of the reference's nms_simple,
the bench configuration itself (B=64, seed 1000) and image sizes whose P5 level has h*w % 4 != 0."""
import numpy as np
import pytest
import torch

from conftest import load_golden, golden_names, t, digest_close
from oracle import ref_port as orc, c_oracle
from yolo_scratch_b200 import synth


@pytest.mark.parametrize("name", golden_names("bboxloss"))
def test_bbox_loss_port(name):
    g = load_golden(name)
    pd = t(g["pred_dist"]).requires_grad_(True)
    pb = t(g["pred_bboxes"]).requires_grad_(True)
    li, ld = orc.bbox_loss(pd, pb, t(g["anchor_points"]), t(g["target_bboxes"]), t(g["target_scores"]), t(g["tss"]),
                           t(g["fg_mask"]), int(g["reg_max"]))
    assert torch.allclose(li.detach(), t(g["loss_iou"]), rtol=1e-6)
    assert torch.allclose(ld.detach(), t(g["loss_dfl"]), rtol=1e-6)
    (li + ld).backward()
    assert torch.allclose(pd.grad, t(g["g_pred_dist"]), rtol=1e-5, atol=1e-9)
    assert torch.allclose(pb.grad, t(g["g_pred_bboxes"]), rtol=1e-5, atol=1e-9)


def test_ciou_grad_port():
    g = load_golden("ciou_grad_s92")
    x = t(g["b1"]).requires_grad_(True)
    out = orc.pair_iou_xyxy(x, t(g["b2"]), ciou=True)
    assert torch.allclose(out.detach(), t(g["ciou"]), rtol=1e-6, atol=1e-7)
    (out * t(g["gout"])).sum().backward()
    assert torch.allclose(x.grad, t(g["g_b1"]), rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize("name", golden_names("nmsclass"))
def test_class_aware_nms_oracles_pinned(name):
    """Both oracles of the class-aware mode -- the C spec's labelled branch and the torch port -- against keep lists
    produced by the reference's own nms_simple run per class subset (and the class-agnostic lists on the same inputs)."""
    g = load_golden(name)
    pred = t(g["pred"])
    conf, thr = float(g["conf"]), float(g["iou"])
    for b in range(pred.shape[0]):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], conf)
        keep_c = src[c_oracle.nms(xyxy, sc, thr, labels=cl)]
        assert np.array_equal(keep_c, g[f"keep{b}"]), "C oracle (labels) vs reference composition"
        keep_p = orc.nms_class_aware(torch.from_numpy(xyxy), torch.from_numpy(sc), torch.from_numpy(cl).long(), thr)
        assert np.array_equal(src[keep_p], g[f"keep{b}"]), "torch port vs reference composition"
        assert np.array_equal(src[c_oracle.nms(xyxy, sc, thr)], g[f"keep_agnostic{b}"])
        d = orc.decode_for_metrics(pred[b:b + 1], 640, conf, thr, class_aware=True, return_keep=True)[0]
        assert np.array_equal(d["keep"].numpy(), g[f"keep{b}"])


def test_loss_c2_port():
    """BASELINE configs[1] exactly as bench.py builds it (B=64, 640 px, seed 1000)."""
    g = load_golden("loss_c2_s1000")
    feats = synth.make_head_levels(64, 640, 80, 16, seed=1000)
    batch = synth.make_gt_batch(64, 80, 20, seed=1001)
    digest_close(feats[0], g, "in_digest", 0.0)
    assert batch["cls"].shape[0] == int(g["n_gt"])
    leaves = [f.requires_grad_(True) for f in feats]
    total, items = orc.detection_loss(leaves, batch, 80, 16, synth.LEVEL_STRIDES)
    total.backward()
    assert torch.allclose(items, t(g["items"]), rtol=1e-6)
    assert torch.allclose(total.detach(), t(g["total"]), rtol=1e-6)
    for leaf, key in zip(leaves, ("g0", "g1", "g2")):
        digest_close(leaf.grad, g, key, 1e-6)


def imgsz_case(g):
    shapes = [tuple(int(v) for v in s) for s in g["shapes"]]
    nc, seed = int(g["nc"]), int(g["seed"])
    if "f0" in g:
        feats = [t(g["f0"]), t(g["f1"]), t(g["f2"])]
    else:
        feats = synth.make_head_levels(int(g["bboxes"].shape[0] and t(g["batch_idx"]).max().item() + 1), 0, nc, 16,
                                       seed=seed, shapes=shapes)
        digest_close(feats[0], g, "in_digest", 0.0)
    batch = {"batch_idx": t(g["batch_idx"]), "cls": t(g["cls"]), "bboxes": t(g["bboxes"])}
    return feats, batch, nc


def check_imgsz_grads(grads, g, rtol):
    for gr, key in zip(grads, ("g0", "g1", "g2")):
        if key in g:
            ref = t(g[key])
            scale = ref.abs().max().item() or 1.0
            assert (gr - ref).abs().max().item() <= rtol * scale, key
        else:
            digest_close(gr, g, key, rtol)


@pytest.mark.parametrize("name", golden_names("loss_imgsz") + golden_names("loss_rect"))
def test_loss_unaligned_levels_port(name):
    g = load_golden(name)
    feats, batch, nc = imgsz_case(g)
    leaves = [f.clone().requires_grad_(True) for f in feats]
    total, items = orc.detection_loss(leaves, batch, nc, 16, synth.LEVEL_STRIDES)
    total.backward()
    assert torch.allclose(items, t(g["items"]), rtol=1e-6)
    assert torch.allclose(total.detach(), t(g["total"]), rtol=1e-6)
    check_imgsz_grads([l.grad for l in leaves], g, 1e-6)


MODE_FIXTURES = golden_names("assign_ciou") + golden_names("assign_topk")


@pytest.mark.parametrize("name", MODE_FIXTURES)
def test_assigner_modes_port(name):
    """Non-default assigner modes (north_star wording) of the port against compositions of reference code
    (make_golden_r2.py: the reference assigner with its own bbox_iou(CIoU=True) as the pair metric; the reference forward
    with mask_pos restricted to the per-GT top-k of align_metric).  top-k: parity unpinned by the reference itself."""
    g = load_golden(name)
    inp = [t(g[k]) for k in ("pd_scores", "pd_bboxes", "anc_points", "gt_labels", "gt_bboxes", "mask_gt")]
    tl, tb, ts, fg, tgi = orc.assign(*inp, int(g["nc"]), 0.5, 4.0, ciou_assign=bool(int(g["ciou"])), topk=int(g["topk"]))
    assert torch.equal(fg, t(g["fg_mask"]))
    assert torch.equal(tgi.int(), t(g["target_gt_idx"]))
    assert torch.equal(tl, t(g["target_labels"]))
    assert torch.equal(tb, t(g["target_bboxes"]))
    assert torch.allclose(ts, t(g["target_scores"]), rtol=1e-6, atol=1e-9)
    base = orc.assign(*inp, int(g["nc"]), 0.5, 4.0)
    assert not torch.equal(base[4], tgi), "the fixture must exercise the mode"
