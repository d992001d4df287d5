"""GPU parity: fused detection loss (ys_detection_loss through the drop-in v8DetectionLoss) vs the reference
fixtures and the oracle port.  Loss items / total / gradients within 1e-5 relative (gradients relative to the
gradient's max-norm); assignment indices bit-exact through the spliced-oracle test (§4.2)."""
import pytest
import torch

from conftest import load_golden, golden_names, t, digest_close
from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5


class Detect(torch.nn.Module):
    def __init__(self, nc, reg_max=16, strides=synth.LEVEL_STRIDES):
        super().__init__()
        self.nc, self.reg_max = nc, reg_max
        self.stride = torch.tensor([float(s) for s in strides])
        self.p = torch.nn.Parameter(torch.zeros(1))


def make_criterion(nc):
    from yolo_scratch_b200.losses import v8DetectionLoss
    return v8DetectionLoss(Detect(nc).cuda())


def run_cuda(feats, batch, nc, export=False):
    crit = make_criterion(nc)
    crit.export_assignment = export
    leaves = [f.cuda().requires_grad_(True) for f in feats]
    total, items = crit(leaves, {k: v.cuda() for k, v in batch.items()})
    total.backward()
    torch.cuda.synchronize()
    return total.detach().cpu(), items.cpu(), [l.grad.cpu() for l in leaves], crit


def assert_grads_close(got, ref, rtol=RTOL):
    for i, (a, b) in enumerate(zip(got, ref)):
        scale = b.abs().max().item() or 1.0
        err = (a - b).abs().max().item() / scale
        assert err <= rtol, f"level {i}: grad err {err:.3e} (rel. to max-norm {scale:.3e})"


@pytest.mark.parametrize("name", golden_names("loss_small"))
def test_loss_golden_small(cuda_lib, name):
    g = load_golden(name)
    feats = [t(g["f0"]), t(g["f1"]), t(g["f2"])]
    batch = {"batch_idx": t(g["batch_idx"]), "cls": t(g["cls"]), "bboxes": t(g["bboxes"])}
    total, items, grads, _ = run_cuda(feats, batch, int(g["nc"]))
    assert items.dtype == torch.float32 and items.shape == (3,)
    assert torch.allclose(items, t(g["items"]), rtol=RTOL, atol=1e-7), (items, g["items"])
    assert torch.allclose(total, t(g["total"]), rtol=RTOL)
    assert_grads_close(grads, [t(g["g0"]), t(g["g1"]), t(g["g2"])])


@pytest.mark.parametrize("name", golden_names("loss_c1"))
def test_loss_golden_c1(cuda_lib, name):
    """BASELINE.json configs[0]: B=2, 640 px, nc=80, <=20 GT/img -- against the reference's own outputs."""
    g = load_golden(name)
    seed = int(g["seed"])
    feats = synth.make_head_levels(2, 640, 80, 16, seed=seed)
    batch = synth.make_gt_batch(2, 80, 20, seed=seed + 1, p_empty=0.0)
    total, items, grads, _ = run_cuda(feats, batch, 80)
    assert torch.allclose(items, t(g["items"]), rtol=RTOL), (items, g["items"])
    assert torch.allclose(total, t(g["total"]), rtol=RTOL)
    for gr, key in zip(grads, ("g0", "g1", "g2")):
        digest_close(gr, g, key, RTOL)


@pytest.mark.parametrize("seed,B,max_gt,kw", [
    (200, 4, 20, {}),
    (201, 8, 20, {"p_tiny": 0.3, "p_nested": 0.3}),
    (202, 3, 40, {"p_empty": 0.34, "p_nested": 0.3}),
])
def test_loss_vs_oracle_640_with_spliced_assignment(cuda_lib, seed, B, max_gt, kw):
    spliced_case(seed, B, 640, max_gt, kw)


@pytest.mark.parametrize("seed,B,imgsz,kw", [
    # BASELINE.json configs[4] shape: 1280 px (A = 33600), exactly 300 GT per image -- more GT than one shared-memory
    # tile (256) holds, so the multi-tile scan and the global-scratch fix-up run; many GT own nothing (loop 2)
    (210, 1, 1280, {"fixed_count": 300, "dense": True}),
    # same GT density on a small image: almost every GT is at risk or forced
    (211, 2, 320, {"fixed_count": 300, "dense": True, "p_tiny": 0.1, "p_nested": 0.2}),
    (212, 2, 320, {"fixed_count": 70, "p_tiny": 0.2, "p_nested": 0.3}),     # just above the containment-hint limit (64)
    # more than 1024 GT in the batch: the general k_pack_gt (not the one-GT-per-thread fast path) packs them
    (213, 8, 320, {"fixed_count": 300, "dense": True, "p_tiny": 0.05, "p_nested": 0.1}),
    # more GT in ONE image than the fix-up's shared-memory scratch holds (1536): global-scratch fix-up, 7 GT tiles
    (214, 1, 320, {"fixed_count": 1700, "dense": True, "p_tiny": 0.02, "p_nested": 0.05}),
])
def test_loss_dense_scene_vs_oracle(cuda_lib, seed, B, imgsz, kw):
    spliced_case(seed, B, imgsz, kw.get("fixed_count", 20), kw)


def test_loss_ownerless_gt_outside_the_risk_hint(cuda_lib):
    """A GT with hundreds of anchor centres inside that is not contained in any other GT, yet owns nothing: two
    smaller GT cover it and win every arg-max contest (the predicted boxes are made tiny, so IoU ~ 1 / GT area).
    k_pack_gt's launch-order hint does not see such a GT; loop 2 (:146-162) fires for it all the same.  Indices bit-exact
    against the spliced oracle, in a sparse image and next to an image with more than 64 GT."""
    B, imgsz, nc = 3, 640, 80
    feats = synth.make_head_levels(B, imgsz, nc, 16, seed=260)
    for f in feats:
        f[:, 0:64:16] += 9.0            # bin 0 of every side dominates: predicted boxes of a fraction of a cell
    covered = torch.tensor([[100.0, 100.0, 300.0, 300.0],       # A: 625 stride-8 centres inside, not inside B1 or B2
                            [99.0, 99.0, 301.0, 200.5],         # B1 and B2 cover A and are smaller
                            [99.0, 199.5, 301.0, 301.0],
                            [400.0, 50.0, 600.0, 250.0],        # a second such triple, listed cover-first
                            [399.0, 49.0, 500.5, 251.0],
                            [499.5, 49.0, 601.0, 251.0],
                            [20.0, 400.0, 90.0, 470.0]]) / imgsz
    order = torch.tensor([1, 0, 2, 4, 5, 3, 6])
    dense = synth.make_gt_batch(1, nc, 70, seed=262, imgsz=imgsz, fixed_count=70, p_nested=0.3)
    batch = {"batch_idx": torch.cat((torch.zeros(7, dtype=torch.long), torch.ones(7, dtype=torch.long),
                                     torch.full((70,), 2, dtype=torch.long))),
             "cls": torch.cat((torch.arange(7).view(7, 1), torch.arange(7).view(7, 1) + 10, dense["cls"])),
             "bboxes": torch.cat((covered, covered[order], dense["bboxes"]))}
    ex, parts = spliced_check(feats, batch, B, imgsz, nc)
    # loop 2 fires for the covered GT, and the final resolve (:165) hands the forced anchor back to the smaller cover
    tgi = ex["target_gt_idx"]
    assert int((tgi[0] == 0).sum()) == 0 and int((tgi[0] == 3).sum()) == 0
    assert int((tgi[1] == 1).sum()) == 0 and int((tgi[1] == 5).sum()) == 0
    assert int((tgi[0] == 1).sum()) > 300 and int((tgi[1] == 0).sum()) > 300


def test_loss_many_images_mixed_counts_multiblock_pack(cuda_lib):
    """More than 1024 GT spread over 64 images with 0..60 GT each (empty images in the middle and at the end): the
    multi-block k_pack_gt's sorted path -- image boundaries found per row, gaps in gt_off filled, containment hint
    from the halo'd rows -- against the oracle."""
    batch = synth.make_gt_batch(64, 80, 60, seed=217, imgsz=320, p_empty=0.2, p_nested=0.3, p_tiny=0.1)
    assert batch["cls"].shape[0] > 1024 and int(batch["batch_idx"].max()) < 63
    spliced_case(216, 64, 320, 60, {"p_empty": 0.2, "p_nested": 0.3, "p_tiny": 0.1})


@pytest.mark.parametrize("seed,nc", [(220, 1), (221, 13), (222, 23)])
def test_loss_vs_oracle_other_class_counts(cuda_lib, seed, nc):
    """Class counts that are not a multiple of the class rows per work item (10): one partial group (nc = 1), a full
    and a partial one (13), two full and a partial one (23)."""
    spliced_case(seed, 2, 320, 12, {"p_nested": 0.2}, nc=nc)


def spliced_case(seed, B, imgsz, max_gt, kw, nc=80, shapes=None):
    feats = synth.make_head_levels(B, imgsz, nc, 16, seed=seed, shapes=shapes)
    batch = synth.make_gt_batch(B, nc, max_gt, seed=seed + 1, imgsz=imgsz, **kw)
    return spliced_check(feats, batch, B, imgsz, nc)


def spliced_check(feats, batch, B, imgsz, nc=80):
    from oracle import ref_port as orc
    total, items, grads, crit = run_cuda(feats, batch, nc, export=True)
    leaves = [f.clone().requires_grad_(True) for f in feats]
    rtotal, ritems, parts = orc.detection_loss(leaves, batch, nc, 16, synth.LEVEL_STRIDES, return_parts=True)
    rtotal.backward()
    ex = {k: v.cpu() for k, v in crit.last_export.items()}
    # (1) decoded boxes to 1e-6 of the image size
    assert (ex["pred_bboxes_px"] - parts["pred_bboxes_px"]).abs().max().item() <= 1e-6 * imgsz * 4
    # (2) spliced oracle: the oracle assigner fed with OUR decoded boxes must give OUR indices bit-exactly
    A = ex["pred_bboxes_px"].shape[1]
    tg = orc.pack_targets(batch["cls"], batch["bboxes"], batch["batch_idx"], B,
                          torch.tensor([float(feats[0].shape[2]), float(feats[0].shape[3])]) * synth.LEVEL_STRIDES[0])
    anc, st = orc.anchors_for([f.shape[2:] for f in feats], synth.LEVEL_STRIDES)
    flat = torch.cat([f.view(B, 64 + nc, -1) for f in feats], 2)
    scores = flat[:, 64:].permute(0, 2, 1).sigmoid()
    if tg.shape[1] > 0:
        _, _, _, fg, tgi = orc.assign(scores, ex["pred_bboxes_px"], anc * st, tg[..., 0], tg[..., 1:5],
                                      (tg[..., 5] > 0).float(), nc)
        ours_fg = ex["target_gt_idx"] >= 0
        assert torch.equal(ours_fg, fg.bool())
        assert torch.equal(ex["target_gt_idx"].clamp(min=0).long(), tgi)
    # (3) loss and gradients against the oracle with the spliced assignment (a decode difference of one ulp may flip a
    #     near-tie between two GT of one anchor; with 10^7 pairs per dense image that happens), and against the oracle's
    #     own decode as well whenever the two oracle runs agree on every index
    sleaves = [f.clone().requires_grad_(True) for f in feats]
    stotal, sitems, sparts = orc.detection_loss(sleaves, batch, nc, 16, synth.LEVEL_STRIDES, return_parts=True,
                                                assign_boxes_px=ex["pred_bboxes_px"])
    stotal.backward()
    assert torch.allclose(items, sitems, rtol=RTOL, atol=1e-7), (items, sitems)
    assert torch.allclose(total, stotal.detach(), rtol=RTOL)
    assert_grads_close(grads, [l.grad for l in sleaves])
    same = torch.equal(sparts["target_gt_idx"], parts["target_gt_idx"]) and torch.equal(sparts["fg_mask"], parts["fg_mask"])
    flips = (sparts["target_gt_idx"] != parts["target_gt_idx"]).sum().item()
    assert flips <= 1e-5 * parts["target_gt_idx"].numel(), f"{flips} assignments differ between the two decodes"
    if same:
        assert torch.allclose(items, ritems, rtol=RTOL, atol=1e-7), (items, ritems)
        assert torch.allclose(total, rtotal.detach(), rtol=RTOL)
        assert_grads_close(grads, [l.grad for l in leaves])
    return ex, parts


def test_loss_config2_exact_bench_inputs(cuda_lib):
    """BASELINE configs[1] with the very inputs bench.py times (B=64, 640 px, seeds 1000/1001): loss items, total and
    gradient digests against the unmodified reference's outputs (tests/golden/loss_c2_s1000.npz), then the spliced-oracle
    comparison (decoded boxes, bit-exact indices, items, gradients) against the port on the same tensors."""
    g = load_golden("loss_c2_s1000")
    feats = synth.make_head_levels(64, 640, 80, 16, seed=1000)
    batch = synth.make_gt_batch(64, 80, 20, seed=1001)
    digest_close(feats[0], g, "in_digest", 0.0)
    total, items, grads, _ = run_cuda(feats, batch, 80)
    assert torch.allclose(items, t(g["items"]), rtol=RTOL), (items, g["items"])
    assert torch.allclose(total, t(g["total"]), rtol=RTOL)
    for gr, key in zip(grads, ("g0", "g1", "g2")):
        digest_close(gr, g, key, RTOL)
    spliced_case(1000, 64, 640, 20, {})


def test_loss_config5_shape_batch8(cuda_lib):
    """BASELINE configs[4] shape at B=8: 1280 px (A = 33600), exactly 300 GT per image -- 2400 GT take the general
    k_pack_gt, the dense-scene ring depth and the multi-tile GT scan together."""
    spliced_case(215, 8, 1280, 300, {"fixed_count": 300, "dense": True})


@pytest.mark.parametrize("name", golden_names("loss_imgsz") + golden_names("loss_rect"))
def test_loss_unaligned_levels_golden(cuda_lib, name):
    """Image sizes the reference accepts and whose levels have h*w % 4 != 0 (imgsz 672: P5 = 21x21; 96: 6x6 and 3x3;
    a non-square head): reference outputs from make_golden_r2.py.  These levels take the non-TMA tile loaders."""
    from test_oracle_golden_r2 import imgsz_case, check_imgsz_grads
    g = load_golden(name)
    feats, batch, nc = imgsz_case(g)
    total, items, grads, _ = run_cuda(feats, batch, nc)
    assert torch.allclose(items, t(g["items"]), rtol=RTOL), (items, g["items"])
    assert torch.allclose(total, t(g["total"]), rtol=RTOL)
    check_imgsz_grads(grads, g, RTOL)


@pytest.mark.parametrize("seed,B,shapes,nc", [
    (250, 2, [(84, 84), (42, 42), (21, 21)], 80),        # imgsz 672
    (251, 3, [(100, 100), (50, 50), (25, 25)], 13),      # imgsz 800: 625 = 1 mod 4
    (252, 2, [(9, 7), (5, 3), (3, 1)], 5),               # every level unaligned, tiny
    (253, 2, [(18, 18), (9, 9), (5, 5)], 80),            # 324 ok, 81 and 25 not
])
def test_loss_unaligned_levels_vs_oracle(cuda_lib, seed, B, shapes, nc):
    spliced_case(seed, B, shapes[0][0] * 8, 10, {"p_tiny": 0.15, "p_nested": 0.2, "p_empty": 0.0}, nc=nc, shapes=shapes)


def test_loss_second_backward_raises(cuda_lib):
    """The fused call hands its gradient buffers over on the first backward; a second pass through the same graph must
    raise (not return None gradients)."""
    from yolo_scratch_b200._lib import YsError
    feats = synth.make_head_levels(2, 128, 8, 16, seed=260)
    batch = synth.make_gt_batch(2, 8, 5, seed=261, imgsz=128, p_empty=0.0)
    crit = make_criterion(8)
    leaves = [f.cuda().requires_grad_(True) for f in feats]
    total, _ = crit(leaves, {k: v.cuda() for k, v in batch.items()})
    total.backward(retain_graph=True)
    g0 = [l.grad.clone() for l in leaves]
    with pytest.raises(YsError):
        total.backward()
    for l, g in zip(leaves, g0):
        assert torch.equal(l.grad, g)


def test_loss_forward_only_and_tuple_input(cuda_lib):
    """validate(): criterion((y, x), batch) under no_grad takes preds[1] (reference :387-389) and skips the backward."""
    from oracle import ref_port as orc
    feats = synth.make_head_levels(2, 128, 8, 16, seed=210)
    batch = synth.make_gt_batch(2, 8, 5, seed=211, imgsz=128, p_empty=0.0)
    crit = make_criterion(8)
    with torch.no_grad():
        total, items = crit((torch.zeros(1).cuda(), [f.cuda() for f in feats]), {k: v.cuda() for k, v in batch.items()})
    rtotal, ritems = orc.detection_loss(feats, batch, 8, 16, synth.LEVEL_STRIDES)
    assert not total.requires_grad
    assert torch.allclose(items.cpu(), ritems, rtol=RTOL, atol=1e-7)


def test_loss_upstream_gradient_scaling_and_unsorted_targets(cuda_lib):
    """(loss * 0.25).backward() scales the stored gradients on the device; shuffled GT rows give the same result
    when per-image order is preserved (the reference's mask-based gather is order-preserving per image)."""
    feats = synth.make_head_levels(3, 128, 8, 16, seed=220)
    batch = synth.make_gt_batch(3, 8, 6, seed=221, imgsz=128, p_empty=0.0)
    _, items, grads, _ = run_cuda(feats, batch, 8)
    crit = make_criterion(8)
    leaves = [f.cuda().requires_grad_(True) for f in feats]
    # interleave the images' rows round-robin: batch_idx becomes unsorted, within-image order is kept
    per_img = [(batch["batch_idx"] == b).nonzero().squeeze(1).tolist() for b in range(3)]
    perm = []
    while any(per_img):
        for lst in per_img:
            if lst:
                perm.append(lst.pop(0))
    perm = torch.tensor(perm)
    sh = {k: v[perm].cuda() for k, v in batch.items()}
    assert not torch.equal(sh["batch_idx"].cpu(), batch["batch_idx"])
    total, items2 = crit(leaves, sh)
    (total * 0.25).backward()
    torch.cuda.synchronize()
    assert torch.allclose(items2.cpu(), items, rtol=1e-6)
    for a, b in zip(leaves, grads):
        assert torch.allclose(a.grad.cpu(), b * 0.25, rtol=1e-6, atol=1e-12)
    # rows with batch_idx outside [0, B) are ignored, as the reference's `batch_idx == i` masks do
    junk = {"batch_idx": torch.cat((sh["batch_idx"], torch.tensor([7, -1]).cuda())),
            "cls": torch.cat((sh["cls"], torch.zeros(2, 1, dtype=torch.long).cuda())),
            "bboxes": torch.cat((sh["bboxes"], torch.rand(2, 4).cuda()))}
    with torch.no_grad():
        _, items3 = crit([f.cuda() for f in feats], junk)
    assert torch.allclose(items3.cpu(), items, rtol=1e-6)


def test_loss_full_config2_properties(cuda_lib):
    """BASELINE.json configs[1] shape (B=64, 640 px): size-independent properties -- shard additivity (the path is
    per-image independent): unnormalised per-shard sums recombine to the full-batch loss; zero box
    gradients on background anchors; finite everything."""
    B = 64
    feats = synth.make_head_levels(B, 640, 80, 16, seed=230)
    batch = synth.make_gt_batch(B, 80, 20, seed=231)
    crit = make_criterion(80)
    crit.export_assignment = True
    leaves = [f.cuda().requires_grad_(True) for f in feats]
    total, items = crit(leaves, {k: v.cuda() for k, v in batch.items()})
    total.backward()
    ex = crit.last_export
    fg = ex["target_gt_idx"] >= 0
    tss = max(ex["norm"].double().sum().item(), 1.0)
    assert torch.isfinite(items).all() and all(torch.isfinite(l.grad).all() for l in leaves)
    # background anchors get exactly zero gradient on the 64 box channels
    off = 0
    for l in leaves:
        hw = l.shape[2] * l.shape[3]
        bg = ~fg[:, off:off + hw]
        gbox = l.grad[:, :64].reshape(B, 64, hw)
        assert (gbox.abs().amax(1)[bg] == 0).all()
        off += hw
    # shards of 16 images: sum_s items_s * tss_s == items * tss
    acc = torch.zeros(3, dtype=torch.float64)
    for s in range(4):
        sel = (batch["batch_idx"] >= 16 * s) & (batch["batch_idx"] < 16 * (s + 1))
        sb = {"batch_idx": (batch["batch_idx"][sel] - 16 * s).cuda(), "cls": batch["cls"][sel].cuda(),
              "bboxes": batch["bboxes"][sel].cuda()}
        c2 = make_criterion(80)
        c2.export_assignment = True
        with torch.no_grad():
            _, it = c2([f[16 * s:16 * (s + 1)].cuda() for f in feats], sb)
        tss_s = max(c2.last_export["norm"].double().sum().item(), 1.0)
        assert torch.equal(c2.last_export["target_gt_idx"], ex["target_gt_idx"][16 * s:16 * (s + 1)])
        acc += it.double().cpu() * tss_s
    assert torch.allclose(acc / tss, items.double().cpu(), rtol=1e-5)


def test_general_pack_unsorted_many_gt(cuda_lib):
    """More than 1024 GT with interleaved batch_idx: the general k_pack_gt's stable-rank walk.  Within-image order is
    kept, so the result must equal the sorted call's bit for bit (same kernels, same inputs after packing); rows with
    batch_idx outside [0, B) are dropped."""
    B = 6
    feats = synth.make_head_levels(B, 320, 80, 16, seed=240)
    batch = synth.make_gt_batch(B, 80, 260, seed=241, imgsz=320, fixed_count=260, dense=True)
    assert batch["cls"].shape[0] > 1024
    total, items, grads, _ = run_cuda(feats, batch, 80)
    per_img = [list(torch.nonzero(batch["batch_idx"] == b).flatten().tolist()) for b in range(B)]
    perm = []
    while any(per_img):
        for lst in per_img:
            if lst:
                perm.append(lst.pop(0))
    perm = torch.tensor(perm)
    sh = {k: v[perm] for k, v in batch.items()}
    sh = {"batch_idx": torch.cat((sh["batch_idx"], torch.tensor([B, -3]))),
          "cls": torch.cat((sh["cls"], torch.zeros(2, 1, dtype=torch.long))),
          "bboxes": torch.cat((sh["bboxes"], torch.rand(2, 4)))}
    total2, items2, grads2, _ = run_cuda(feats, sh, 80)
    assert torch.equal(items, items2) and torch.equal(total, total2)
    for a, b in zip(grads, grads2):
        assert torch.equal(a, b)
