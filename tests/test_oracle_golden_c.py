"""CPU: the plain-C scalar oracle (oracle/scalar_spec.c) against the reference fixtures:
bit-exact indices / keep lists / IoU values."""
import numpy as np
import pytest
import torch

from conftest import load_golden, golden_names, t, digest_close
from oracle import c_oracle
from yolo_scratch_b200 import synth


@pytest.mark.parametrize("name", golden_names("assign_small"))
def test_c_assigner_small(name):
    g = load_golden(name)
    alpha = float(g["alpha"]) if "alpha" in g else 0.5
    beta = float(g["beta"]) if "beta" in g else 4.0
    r = c_oracle.assign(g["pd_scores"], g["pd_bboxes"], g["anc_points"], g["gt_labels"], g["gt_bboxes"],
                        g["mask_gt"], int(g["nc"]), alpha, beta)
    assert torch.equal(r["fg_mask"], t(g["fg_mask"]))
    assert torch.equal(r["target_gt_idx"], t(g["target_gt_idx"]))
    assert torch.equal(r["target_labels"], t(g["target_labels"]))
    assert torch.equal(r["target_bboxes"], t(g["target_bboxes"]))
    ts = t(g["target_scores"])
    assert torch.allclose(r["norm"], ts.sum(-1), rtol=2e-6, atol=1e-9)


@pytest.mark.parametrize("name", golden_names("assign_c1"))
def test_c_assigner_c1(name):
    g = load_golden(name)
    kw = dict(zip([str(k) for k in g["gt_kw_keys"]], [float(v) for v in g["gt_kw_vals"]]))
    inp = synth.make_assigner_inputs(2, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, 20,
                                     int(g["seed"]), **kw)
    r = c_oracle.assign(*inp, 80)
    assert torch.equal(r["fg_mask"], t(g["fg_mask"]))
    assert torch.equal(r["target_gt_idx"].to(torch.int16), t(g["target_gt_idx"]))
    assert torch.equal(r["target_labels"].to(torch.int16), t(g["target_labels"]))


@pytest.mark.parametrize("name", golden_names("nms_"))
def test_c_nms(name):
    g = load_golden(name)
    B, A, nc = int(g["B"]), int(g["A"]), int(g["nc"])
    if "pred" in g:
        pred = t(g["pred"])
    else:
        pred = synth.make_nms_predictions(B, A, nc, 640, seed=int(g["seed"]), clustered=bool(g["clustered"]))
        digest_close(pred, g, "in_digest", 0.0)
    for b in range(B):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], float(g["conf"]))
        keep = c_oracle.nms(xyxy, sc, float(g["iou"]))
        assert np.array_equal(src[keep], g[f"keep{b}"])
        assert np.array_equal(cl[keep], g[f"labels{b}"])
        assert np.array_equal(sc[keep], g[f"scores{b}"])
        boxes = np.clip(xyxy[keep] / np.float32(640), 0.0, 1.0)
        assert np.array_equal(boxes, g[f"boxes{b}"])


def test_c_iou_matrix():
    g = load_golden("misc_iou_anchor")
    assert torch.equal(c_oracle.iou_matrix(g["b1"][:16], g["b2"]), t(g["pair_matrix"]))
