"""GPU parity of the stand-alone entry points of losses/: bbox_iou in every mode (ys_bbox_iou), its CIoU gradient,
BboxLoss.forward + backward (ys_bbox_loss), ys_pairwise_iou kind 1, and the tensor helpers bbox2dist / dist2bbox /
make_anchors -- all against outputs of the unmodified reference (tests/golden/misc_iou_anchor.npz from
make_golden.py; bboxloss_s*.npz and ciou_grad_s92.npz from make_golden_r2.py;
reference losses/yolo_v8_loss.py:12-61, :280-330, :541-564)."""
import pytest
import torch

from conftest import load_golden, golden_names, t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def misc():
    return load_golden("misc_iou_anchor")


@pytest.mark.parametrize("key,kw", [
    ("iou_xyxy", dict(xywh=False)),
    ("giou_xyxy", dict(xywh=False, GIoU=True)),
    ("diou_xyxy", dict(xywh=False, DIoU=True)),
    ("ciou_xyxy", dict(xywh=False, CIoU=True)),
    ("iou_xywh", dict(xywh=True)),
    ("ciou_xywh", dict(xywh=True, CIoU=True)),
])
def test_bbox_iou_modes_golden(cuda_lib, misc, key, kw):
    from yolo_scratch_b200.losses import bbox_iou
    a, b = ("c1", "c2") if kw["xywh"] else ("b1", "b2")
    out = bbox_iou(t(misc[a]).cuda(), t(misc[b]).cuda(), **kw)
    ref = t(misc[key])
    assert out.shape == ref.shape == (64, 1)
    # +, -, *, / in the reference's order: bit-exact except where atan enters (CIoU), 1e-6 there
    if key in ("iou_xyxy", "iou_xywh"):
        assert torch.equal(out.cpu(), ref)
    else:
        assert torch.allclose(out.cpu(), ref, rtol=1e-6, atol=1e-7)


def test_bbox_iou_broadcast_shapes(cuda_lib, misc):
    """The assigner's call shape (reference :199): (B,A,1,4) x (B,1,G,4) -> (B,A,G,1)."""
    from yolo_scratch_b200.losses import bbox_iou
    from oracle import ref_port as orc
    b1, b2 = t(misc["b1"]).view(2, 32, 1, 4), t(misc["b2"])[:10].view(1, 1, 10, 4).expand(2, 1, 10, 4)
    out = bbox_iou(b1.cuda(), b2.cuda(), xywh=False)
    ref = orc.pair_iou_xyxy(b1, b2)
    assert out.shape == (2, 32, 10, 1)
    assert torch.equal(out.cpu(), ref)


def test_bbox_iou_ciou_gradient_golden(cuda_lib):
    from yolo_scratch_b200.losses import bbox_iou
    g = load_golden("ciou_grad_s92")
    x = t(g["b1"]).cuda().requires_grad_(True)
    out = bbox_iou(x, t(g["b2"]).cuda(), xywh=False, CIoU=True)
    assert torch.allclose(out.detach().cpu(), t(g["ciou"]), rtol=1e-6, atol=1e-7)
    (out * t(g["gout"]).cuda()).sum().backward()
    ref = t(g["g_b1"])
    assert (x.grad.cpu() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()


def test_bbox_iou_unsupported_gradients_raise(cuda_lib, misc):
    """Anything the device kernel cannot differentiate raises at call time instead of dropping a gradient."""
    from yolo_scratch_b200.losses import bbox_iou
    from yolo_scratch_b200._lib import YsError
    b1, b2 = t(misc["b1"]).cuda(), t(misc["b2"]).cuda()
    with pytest.raises(YsError):
        bbox_iou(b1, b2.clone().requires_grad_(True), xywh=False, CIoU=True)
    with pytest.raises(YsError):
        bbox_iou(b1.clone().requires_grad_(True), b2, xywh=False, GIoU=True)
    with pytest.raises(YsError):
        bbox_iou(b1.clone().requires_grad_(True), b2, xywh=True, CIoU=True)
    with torch.no_grad():      # no graph, no restriction
        bbox_iou(b1.clone().requires_grad_(True), b2, xywh=False, GIoU=True)
    with pytest.raises(YsError):
        bbox_iou(b1.cpu(), b2.cpu(), xywh=False)


@pytest.mark.parametrize("name", golden_names("bboxloss"))
def test_bbox_loss_golden(cuda_lib, name):
    from yolo_scratch_b200.losses import BboxLoss
    g = load_golden(name)
    reg_max = int(g["reg_max"])
    args = [t(g[k]).cuda() for k in ("anchor_points", "target_bboxes", "target_scores")]
    fg = t(g["fg_mask"]).cuda()
    for wi, wd, kd, kb in ((1.0, 1.0, "g_pred_dist", "g_pred_bboxes"), (2.0, 0.5, "g_pred_dist_w", "g_pred_bboxes_w")):
        pd = t(g["pred_dist"]).cuda().requires_grad_(True)
        pb = t(g["pred_bboxes"]).cuda().requires_grad_(True)
        # target_scores_sum as the reference passes it: max(target_scores.sum(), 1) -- a tensor or the int 1
        li, ld = BboxLoss(reg_max)(pd, pb, args[0], args[1].clone(), args[2], t(g["tss"]).cuda(), fg)
        assert torch.allclose(li.detach().cpu(), t(g["loss_iou"]), rtol=1e-5)
        assert torch.allclose(ld.detach().cpu(), t(g["loss_dfl"]), rtol=1e-5)
        (li * wi + ld * wd).backward()
        for got, key in ((pd.grad, kd), (pb.grad, kb)):
            ref = t(g[key])
            assert (got.cpu() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item(), key
    # python-number target_scores_sum (the `max(..., 1)` branch that returns the int)
    with torch.no_grad():
        li2, _ = BboxLoss(reg_max)(t(g["pred_dist"]).cuda(), t(g["pred_bboxes"]).cuda(), args[0], args[1], args[2],
                                   float(g["tss"]), fg)
    assert torch.allclose(li2.cpu(), t(g["loss_iou"]), rtol=1e-5)


def test_bbox_loss_reg_max_one(cuda_lib):
    """reg_max == 1: the DFL term is the constant 0 (reference :307-308)."""
    from yolo_scratch_b200.losses import BboxLoss
    g = load_golden("bboxloss_s91")
    n = g["pred_bboxes"].shape
    pd = torch.zeros(n[0], n[1], 4).cuda()
    li, ld = BboxLoss(1)(pd, t(g["pred_bboxes"]).cuda(), t(g["anchor_points"]).cuda(), t(g["target_bboxes"]).cuda(),
                         t(g["target_scores"]).cuda(), t(g["tss"]).cuda(), t(g["fg_mask"]).cuda())
    assert float(ld) == 0.0
    assert torch.allclose(li.cpu(), t(g["loss_iou"]), rtol=1e-5)


def test_pairwise_iou_kinds(cuda_lib, misc):
    """kind 0 = calculate_iou_batch(_simple) (utils/metrics.py:49-81, train_yolo11_cuda.py:402-437); kind 1 = the
    broadcast bbox_iou(xywh=False) of the assigner (losses/yolo_v8_loss.py:31-44)."""
    from oracle import ref_port as orc
    from yolo_scratch_b200 import _lib
    from yolo_scratch_b200.train_yolo11_cuda import calculate_iou_batch_simple
    from yolo_scratch_b200.utils.metrics import calculate_iou_batch
    b1, b2 = t(misc["b1"]), t(misc["b2"])
    assert torch.equal(calculate_iou_batch(b1[:16].cuda(), b2.cuda()).cpu(), t(misc["pair_matrix"]))
    assert torch.equal(calculate_iou_batch_simple(b1[:1].cuda(), b2.cuda()).cpu(), t(misc["pair_simple"]))
    one = calculate_iou_batch_simple(b1[0].cuda(), b2[:1].cuda())          # (4,) x (1,4): the reference squeezes to 0-dim
    assert one.dim() == 0 and float(one) == float(misc["pair_simple"][0])
    out = torch.empty(64, 64).cuda()
    d1, d2 = b1.cuda(), b2.cuda()
    rc = _lib.lib().ys_pairwise_iou(d1.data_ptr(), 64, d2.data_ptr(), 64, 1, out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    ref = orc.pair_iou_xyxy(b1.view(64, 1, 4), b2.view(1, 64, 4)).squeeze(-1)
    assert torch.equal(out.cpu(), ref)
    assert torch.equal(out.cpu().diagonal(), t(misc["iou_xyxy"]).squeeze(-1))


def test_tensor_helpers_golden(cuda_lib, misc):
    from yolo_scratch_b200.losses import bbox2dist, dist2bbox, make_anchors
    anc = t(misc["anc"]).cuda()
    assert torch.equal(bbox2dist(anc, t(misc["b1"]).cuda() / 8.0, 15).cpu(), t(misc["bbox2dist"]))
    d = t(misc["dist"]).cuda()
    assert torch.equal(dist2bbox(d, anc.unsqueeze(0), xywh=False, dim=-1).cpu(), t(misc["dist2bbox_xyxy"]))
    assert torch.equal(dist2bbox(d, anc.unsqueeze(0), xywh=True, dim=-1).cpu(), t(misc["dist2bbox_xywh"]))
    feats = [torch.zeros(1, 1, 5, 7).cuda(), torch.zeros(1, 1, 3, 4).cuda()]
    ap, st = make_anchors(feats, [8.0, 16.0], 0.5)
    assert ap.is_cuda and torch.equal(ap.cpu(), t(misc["anchors_5x7_3x4"])) and torch.equal(st.cpu(), t(misc["strides_5x7_3x4"]))
