"""GPU parity: dense TaskAlignedAssigner (ys_assign through the drop-in class) vs the reference fixtures
and the oracles.  Indices / masks / labels / boxes bit-exact; target_scores to 1e-5 relative."""
import pytest
import torch

from conftest import load_golden, golden_names, t
from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu


def _run(inp, nc, alpha=0.5, beta=4.0):
    from yolo_scratch_b200.losses import TaskAlignedAssigner
    a = TaskAlignedAssigner(topk=50, num_classes=nc, alpha=alpha, beta=beta)
    out = a.forward(*[x.cuda() for x in inp])
    torch.cuda.synchronize()
    return [o.cpu() for o in out]


def _check(out, ref, rtol=1e-5):
    tl, tb, ts, fg, tgi = out
    rtl, rtb, rts, rfg, rtgi = ref
    assert fg.dtype == torch.bool and tgi.dtype == torch.int64 and tl.dtype == torch.float32
    assert torch.equal(fg, rfg), f"fg_mask differs at {(fg != rfg).nonzero()[:5].tolist()}"
    assert torch.equal(tgi, rtgi.long()), f"target_gt_idx differs at {(tgi != rtgi.long()).nonzero()[:5].tolist()}"
    assert torch.equal(tl, rtl)
    assert torch.equal(tb, rtb)
    assert torch.equal(ts != 0, rts != 0)
    assert torch.allclose(ts, rts, rtol=rtol, atol=1e-9)


@pytest.mark.parametrize("name", golden_names("assign_small"))
def test_assigner_golden_small(cuda_lib, name):
    g = load_golden(name)
    alpha = float(g["alpha"]) if "alpha" in g else 0.5
    beta = float(g["beta"]) if "beta" in g else 4.0
    inp = [t(g[k]) for k in ("pd_scores", "pd_bboxes", "anc_points", "gt_labels", "gt_bboxes", "mask_gt")]
    out = _run(inp, int(g["nc"]), alpha, beta)
    ref = [t(g[k]) for k in ("target_labels", "target_bboxes", "target_scores", "fg_mask", "target_gt_idx")]
    _check(out, ref)


@pytest.mark.parametrize("name", golden_names("assign_c1"))
def test_assigner_golden_c1(cuda_lib, name):
    g = load_golden(name)
    kw = dict(zip([str(k) for k in g["gt_kw_keys"]], [float(v) for v in g["gt_kw_vals"]]))
    inp = synth.make_assigner_inputs(2, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, 20, int(g["seed"]), **kw)
    tl, tb, ts, fg, tgi = _run(inp, 80)
    assert torch.equal(fg, t(g["fg_mask"]))
    assert torch.equal(tgi.to(torch.int16), t(g["target_gt_idx"]))
    assert torch.equal(tl.to(torch.int16), t(g["target_labels"]))
    assert torch.equal(ts.nonzero().int(), t(g["ts_nz_idx"]))
    assert torch.allclose(ts[ts != 0], t(g["ts_nz_val"]), rtol=1e-5, atol=0)


@pytest.mark.parametrize("seed,B,max_gt,kw", [
    (100, 8, 20, {}),
    (101, 8, 20, {"p_tiny": 0.3, "p_nested": 0.3}),
    (102, 4, 60, {"p_tiny": 0.2, "p_nested": 0.4, "p_empty": 0.25}),
    (103, 16, 3, {"p_tiny": 0.5}),
])
def test_assigner_vs_oracles_640(cuda_lib, seed, B, max_gt, kw):
    """A=8400: CUDA vs the torch port (all five outputs) and vs the plain-C scalar oracle (indices)."""
    from oracle import ref_port as orc, c_oracle
    inp = synth.make_assigner_inputs(B, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, max_gt, seed, **kw)
    out = _run(inp, 80)
    ref = orc.assign(*inp, 80)
    _check(out, ref)
    c = c_oracle.assign(*inp, 80)
    assert torch.equal(out[3], c["fg_mask"]) and torch.equal(out[4].int(), c["target_gt_idx"])


def test_assigner_dense_scene_300gt(cuda_lib):
    """Dense-scene shape (multi-tile GT loop: G > 256) on a reduced grid, vs the C oracle."""
    from oracle import c_oracle
    inp = synth.make_assigner_inputs(2, ((40, 40), (20, 20), (10, 10)), synth.LEVEL_STRIDES, 80, 300, 110,
                                     fixed_count=300, dense=True, p_nested=0.2, p_tiny=0.1)
    out = _run(inp, 80)
    c = c_oracle.assign(*inp, 80)
    assert torch.equal(out[3], c["fg_mask"]) and torch.equal(out[4].int(), c["target_gt_idx"])
    assert torch.equal(out[0], c["target_labels"]) and torch.equal(out[1], c["target_bboxes"])
    assert torch.allclose(out[2].sum(-1), c["norm"], rtol=1e-5, atol=1e-9)


def test_assigner_empty_gt_dim(cuda_lib):
    from yolo_scratch_b200.losses import TaskAlignedAssigner
    a = TaskAlignedAssigner(num_classes=8)
    ps = torch.rand(2, 336, 8).cuda()
    out = a.forward(ps, torch.rand(2, 336, 4).cuda(), torch.rand(336, 2).cuda(), torch.zeros(2, 0).cuda(),
                    torch.zeros(2, 0, 4).cuda(), torch.zeros(2, 0).cuda())
    assert out[0].dtype == torch.int64 and (out[0] == 8).all() and out[3].dtype == torch.float32 and out[3].sum() == 0


def test_every_valid_gt_owns_an_anchor_or_was_stolen(cuda_lib):
    """Property: after assignment each valid GT owns >= 1 fg anchor unless its forced anchor went to
    another GT in the final resolve; fg anchors point at valid GT unless resolved to an all-zero-overlap slot."""
    inp = synth.make_assigner_inputs(8, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, 20, 120, p_tiny=0.3)
    tl, tb, ts, fg, tgi = _run(inp, 80)
    mask_gt = inp[5]
    owned = torch.zeros_like(mask_gt)
    for b in range(8):
        idx = tgi[b][fg[b]]
        owned[b].scatter_(0, idx, 1.0)
    frac = (owned * mask_gt).sum() / mask_gt.sum()
    assert frac > 0.9
    assert (ts.sum(-1)[~fg] == 0).all()


@pytest.mark.parametrize("name", golden_names("assign_ciou") + golden_names("assign_topk"))
def test_assigner_modes_golden(cuda_lib, name):
    """use_topk / ciou_assign (keyword-only, default off) against fixtures built from compositions of reference code."""
    from yolo_scratch_b200.losses import TaskAlignedAssigner
    g = load_golden(name)
    inp = [t(g[k]) for k in ("pd_scores", "pd_bboxes", "anc_points", "gt_labels", "gt_bboxes", "mask_gt")]
    topk = int(g["topk"])
    a = TaskAlignedAssigner(topk=topk or 50, num_classes=int(g["nc"]), alpha=0.5, beta=4.0, use_topk=topk > 0,
                            ciou_assign=bool(int(g["ciou"])))
    out = [o.cpu() for o in a.forward(*[x.cuda() for x in inp])]
    ref = [t(g[k]) for k in ("target_labels", "target_bboxes", "target_scores", "fg_mask", "target_gt_idx")]
    _check(out, ref)


@pytest.mark.parametrize("seed,topk,ciou,B,max_gt,kw", [
    (120, 10, False, 4, 20, {}),
    (121, 13, False, 4, 20, {"p_tiny": 0.2, "p_nested": 0.3}),
    (122, 1, False, 3, 40, {"p_nested": 0.3}),
    (123, 0, True, 4, 20, {"p_tiny": 0.2, "p_nested": 0.3}),
    (124, 5, True, 2, 30, {"p_nested": 0.4}),
    (125, 9000, False, 2, 10, {}),            # k larger than the anchor count: every positive candidate stays
])
def test_assigner_modes_640_vs_port(cuda_lib, seed, topk, ciou, B, max_gt, kw):
    """The modes at the 640 px shape (A = 8400) against the torch oracle (oracle/ref_port.assign): indices bit-exact
    (continuous random inputs: alignment metrics are unique)."""
    from oracle import ref_port as orc
    from yolo_scratch_b200.losses import TaskAlignedAssigner
    inp = synth.make_assigner_inputs(B, ((80, 80), (40, 40), (20, 20)), synth.LEVEL_STRIDES, 80, max_gt, seed, **kw)
    a = TaskAlignedAssigner(topk=topk or 50, num_classes=80, alpha=0.5, beta=4.0, use_topk=topk > 0, ciou_assign=ciou)
    out = [o.cpu() for o in a.forward(*[x.cuda() for x in inp])]
    ref = orc.assign(*inp, 80, 0.5, 4.0, ciou_assign=ciou, topk=topk)
    _check(out, ref)
