"""GPU parity: batched NMS / decode-for-metrics / nms_simple vs the reference fixtures and the C oracle.
Keep lists bit-exact; kept boxes / scores / labels bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden, golden_names, t, digest_close
from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu


def _pred_of(g):
    B, A, nc = int(g["B"]), int(g["A"]), int(g["nc"])
    if "pred" in g:
        return t(g["pred"])
    pred = synth.make_nms_predictions(B, A, nc, 640, seed=int(g["seed"]), clustered=bool(g["clustered"]))
    digest_close(pred, g, "in_digest", 0.0)
    return pred


@pytest.mark.parametrize("name", golden_names("nms_"))
def test_decode_for_metrics_golden(cuda_lib, name):
    from yolo_scratch_b200.train_yolo11_cuda import decode_predictions_for_metrics, nms_batched
    g = load_golden(name)
    pred = _pred_of(g).cuda()
    conf, thr = float(g["conf"]), float(g["iou"])
    res = decode_predictions_for_metrics(pred, 640, conf, thr, pred.device)
    ka, _, _, _, kc = nms_batched(pred, conf, thr, 640)
    for b in range(int(g["B"])):
        k = int(kc[b])
        assert np.array_equal(ka[b, :k].cpu().numpy(), g[f"keep{b}"]), f"keep list differs (image {b})"
        assert torch.equal(res[b]["boxes"].cpu(), t(g[f"boxes{b}"]))
        assert torch.equal(res[b]["scores"].cpu(), t(g[f"scores{b}"]))
        assert torch.equal(res[b]["labels"].cpu().int(), t(g[f"labels{b}"]))
        assert res[b]["labels"].dtype == torch.int64


@pytest.mark.parametrize("clustered", [False, True])
@pytest.mark.parametrize("layout", [1, 0])
def test_nms_config4_shape_vs_c_oracle(cuda_lib, clustered, layout):
    """BASELINE.json configs[3] shape per image (8400 candidates x 80 classes, conf 0.001 / IoU 0.7), 6 images."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    B = 6
    pred = synth.make_nms_predictions(B, 8400, 80, 640, seed=300 + int(clustered), clustered=clustered)
    dev = pred.cuda() if layout == 1 else pred.permute(0, 2, 1).contiguous().cuda()
    ka, kb, ks, kl, kc = nms_batched(dev, 0.001, 0.7, 640, layout=layout)
    for b in range(B):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], 0.001)
        keep = c_oracle.nms(xyxy, sc, 0.7)
        k = int(kc[b])
        assert k == len(keep)
        assert np.array_equal(ka[b, :k].cpu().numpy(), src[keep])
        assert np.array_equal(kl[b, :k].cpu().numpy(), cl[keep])
        assert np.array_equal(ks[b, :k].cpu().numpy(), sc[keep])
        assert np.array_equal(kb[b, :k].cpu().numpy(), np.clip(xyxy[keep] / np.float32(640), 0, 1))
        s = ks[b, :k]
        assert (s[1:] <= s[:-1]).all()      # score-sorted


def test_nms_class_aware_vs_oracle(cuda_lib):
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    pred = synth.make_nms_predictions(3, 3000, 6, 640, seed=310, clustered=True)
    ka, _, _, _, kc = nms_batched(pred.cuda(), 0.01, 0.5, 640, class_aware=True)
    for b in range(3):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], 0.01)
        keep = c_oracle.nms(xyxy, sc, 0.5, labels=cl)
        assert np.array_equal(ka[b, :int(kc[b])].cpu().numpy(), src[keep])


@pytest.mark.parametrize("name", golden_names("nmsclass"))
def test_nms_class_aware_golden(cuda_lib, name):
    """Class-aware mode against keep lists produced by the reference's own nms_simple run once per class
    subset (tests/golden/make_golden_r2.py), through the batched entry point and through nms_simple(labels=)."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched, nms_simple
    g = load_golden(name)
    pred = t(g["pred"])
    conf, thr = float(g["conf"]), float(g["iou"])
    ka, _, _, kl, kc = nms_batched(pred.cuda(), conf, thr, 640, class_aware=True)
    ka0, _, _, _, kc0 = nms_batched(pred.cuda(), conf, thr, 640)
    for b in range(pred.shape[0]):
        assert np.array_equal(ka[b, :int(kc[b])].cpu().numpy(), g[f"keep{b}"])
        assert np.array_equal(ka0[b, :int(kc0[b])].cpu().numpy(), g[f"keep_agnostic{b}"])
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], conf)
        keep = nms_simple(torch.from_numpy(xyxy).cuda(), torch.from_numpy(sc).cuda(), thr, labels=torch.from_numpy(cl).cuda())
        assert np.array_equal(src[keep], g[f"keep{b}"])
        assert np.array_equal(kl[b, :int(kc[b])].cpu().numpy(), cl[np.searchsorted(src, g[f"keep{b}"])])


def test_nms_simple_and_iou_helpers(cuda_lib):
    from oracle import ref_port as orc, c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_simple, calculate_iou_batch_simple
    from yolo_scratch_b200.utils import calculate_iou_batch
    g = torch.Generator().manual_seed(320)
    for n in (1, 2, 33, 700, 5000):
        xy = torch.rand(n, 2, generator=g) * 300
        wh = torch.rand(n, 2, generator=g) * 80 + 2
        boxes = torch.cat((xy, xy + wh), 1)
        scores = torch.rand(n, generator=g)
        scores[::5] = scores[0]        # ties -> lower index first
        keep = nms_simple(boxes.cuda(), scores.cuda(), 0.45)
        assert keep == c_oracle.nms(boxes, scores, 0.45)
        if n <= 700:
            assert keep == orc.nms_greedy(boxes, scores, 0.45)
    assert nms_simple(torch.zeros(0, 4).cuda(), torch.zeros(0).cuda(), 0.5) == []
    m = calculate_iou_batch(boxes[:50].cuda(), boxes[:70].cuda()).cpu()
    assert torch.equal(m, orc.iou_matrix(boxes[:50], boxes[:70]))
    v = calculate_iou_batch_simple(boxes[:1].cuda(), boxes[:70].cuda()).cpu()
    assert v.shape == (70,) and torch.equal(v, orc.iou_matrix(boxes[:1], boxes[:70]).squeeze(0))


def test_nms_nothing_passes_and_everything_suppressed(cuda_lib):
    from yolo_scratch_b200.train_yolo11_cuda import decode_predictions_for_metrics
    pred = synth.make_nms_predictions(2, 500, 4, 640, seed=330)
    res = decode_predictions_for_metrics(pred.cuda(), 640, 2.0, 0.5, torch.device("cuda"))
    assert all(r["boxes"].shape == (0, 4) and r["labels"].dtype == torch.int64 for r in res)
    same = pred.clone()
    same[:, :, :4] = torch.tensor([100.0, 100.0, 50.0, 50.0])
    res = decode_predictions_for_metrics(same.cuda(), 640, 0.0, 0.5, torch.device("cuda"))
    assert all(r["boxes"].shape == (1, 4) for r in res)


def test_nms_full_batch_256_properties(cuda_lib):
    """BASELINE.json configs[3] at full size (B=256): keep lists are score-sorted, pairwise IoU of kept boxes <= thr
    (checked on a sample), image 0 and image 255 match the C oracle."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    from yolo_scratch_b200.utils import calculate_iou_batch
    B = 256
    base = synth.make_nms_predictions(8, 8400, 80, 640, seed=340)
    pred = base.repeat(32, 1, 1)
    pred = pred[torch.arange(B).roll(3)]
    dev = pred.cuda()
    ka, kb, ks, kl, kc = nms_batched(dev, 0.001, 0.7, 640)
    torch.cuda.synchronize()
    for b in (0, 255):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], 0.001)
        keep = c_oracle.nms(xyxy, sc, 0.7)
        assert np.array_equal(ka[b, :int(kc[b])].cpu().numpy(), src[keep])
    for b in (17, 130):
        k = int(kc[b])
        assert (ks[b, 1:k] <= ks[b, :k - 1]).all()
        a = ka[b, :min(k, 1500)].long()
        x, y, w, h = dev[b, a, 0], dev[b, a, 1], dev[b, a, 2], dev[b, a, 3]
        bx = torch.stack((x - w / 2, y - h / 2, x + w / 2, y + h / 2), 1)
        iou = calculate_iou_batch(bx, bx)
        iou.fill_diagonal_(0)
        assert (iou <= 0.7).all()


def _adversarial_boxes(n, seed, scale=1.0, offset=0.0):
    """Geometry that stresses the candidate filters of the suppress kernels: lattice boxes sharing edges (differences
    of exactly +0), duplicates, zero-width / zero-height / zero-area boxes, contained boxes whose IoU sits at the
    threshold (i inside j, same height, width = thr * W * (1 + d) for tiny d), shifted copies, tiny boxes."""
    g = torch.Generator().manual_seed(seed)
    cell = 16.0
    ij = torch.randint(0, 24, (n, 2), generator=g).float()
    wh = torch.randint(1, 4, (n, 2), generator=g).float()
    boxes = torch.cat((ij * cell, (ij + wh) * cell), 1)                      # lattice: shared edges everywhere
    k = n // 8
    boxes[k:2 * k] = boxes[:k]                                                # exact duplicates
    boxes[2 * k:3 * k, 2] = boxes[2 * k:3 * k, 0]                             # zero width
    boxes[3 * k:4 * k, 3] = boxes[3 * k:4 * k, 1]                             # zero height
    for thr, lo in ((0.7, 4 * k), (0.3, 5 * k)):                              # IoU at the threshold, both sides of it
        base = boxes[:k].clone()
        W = base[:, 2] - base[:, 0]
        d = (torch.rand(k, generator=g) - 0.5) * 2e-4
        a = torch.rand(k, generator=g) * (1 - thr) * 0.5 * W
        inner = base.clone()
        inner[:, 0] = base[:, 0] + a
        inner[:, 2] = inner[:, 0] + thr * W * (1 + d)
        boxes[lo:lo + k] = inner
    sh = boxes[:k].clone()                                                    # copies shifted along y: IoU = ih / (2h - ih)
    H = sh[:, 3] - sh[:, 1]
    sh[:, 1] += H * torch.rand(k, generator=g) * 0.4
    sh[:, 3] = sh[:, 1] + H
    boxes[6 * k:7 * k] = sh
    tiny = boxes[7 * k:].clone()                                              # tiny boxes, some nested in each other
    tiny[:, 2:] = tiny[:, :2] + torch.rand(tiny.shape[0], 2, generator=g) * 1e-3
    boxes[7 * k:] = tiny
    boxes = boxes * scale + offset
    scores = torch.rand(n, generator=g)
    scores[::7] = scores[3]                                                   # ties -> lower index first
    perm = torch.randperm(n, generator=g)
    return boxes[perm].contiguous(), scores[perm].contiguous()


@pytest.mark.parametrize("n,seed,scale,offset", [(3000, 340, 1.0, 0.0), (9000, 341, 1.0, 0.0), (4000, 342, 1.0, 30000.0),
                                                 (2500, 343, 1e-3, 0.0), (6000, 344, 2.5e4, -3e6)])
def test_nms_adversarial_geometry_vs_c_oracle(cuda_lib, n, seed, scale, offset):
    """Keep lists bit-exact against the C oracle on geometry built to break a conservative candidate filter (touching
    edges, degenerate boxes, IoU within 1e-4 of the threshold, tiny boxes far from the origin where one ulp of a
    coordinate is larger than the box), at thresholds on, inside and outside [0, 1]; 3000..9000 boxes cover the
    5- and 9-column instantiations."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_simple
    boxes, scores = _adversarial_boxes(n, seed, scale, offset)
    for thr in (0.7, 0.3, 0.0, 0.999, 1.0, 1.5):
        keep = nms_simple(boxes.cuda(), scores.cuda(), thr)
        assert keep == c_oracle.nms(boxes, scores, thr), (n, thr)
    cls = (torch.arange(n) % 3).int()
    keep = nms_simple(boxes.cuda(), scores.cuda(), 0.5, labels=cls.cuda())
    assert keep == c_oracle.nms(boxes, scores, 0.5, labels=cls.numpy())


def test_nms_non_finite_and_inverted_boxes_vs_c_oracle(cuda_lib):
    """Boxes with x2 < x1, infinities and NaN take the kernels' exact (unfiltered) path; the keep list still equals the
    oracle's, which applies the reference's `ious <= thr` literally (a NaN IoU drops the box, train_yolo11_cuda.py:396)."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_simple
    boxes, scores = _adversarial_boxes(3000, 350)
    boxes[5] = torch.tensor([50.0, 50.0, 10.0, 90.0])
    boxes[6] = torch.tensor([50.0, 90.0, 90.0, 10.0])
    keep = nms_simple(boxes.cuda(), scores.cuda(), 0.5)
    assert keep == c_oracle.nms(boxes, scores, 0.5)
    boxes[77] = torch.tensor([0.0, 0.0, float("inf"), 10.0])
    boxes[78] = torch.tensor([float("nan"), 0.0, 5.0, 10.0])
    keep = nms_simple(boxes.cuda(), scores.cuda(), 0.5)
    assert keep == c_oracle.nms(boxes, scores, 0.5)


@pytest.mark.parametrize("clustered", [False, True])
def test_nms_1280px_candidate_count_vs_c_oracle(cuda_lib, clustered):
    """33600 candidates per image (validation at 1280 px, the resolution of BASELINE configs[4]): more keys than one
    shared-memory sort holds -- segment sorts + merge, then the shared-memory-free suppress path.  The reference's
    nms_simple has no size limit (train_yolo11_cuda.py:361-399)."""
    from oracle import c_oracle
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched, nms_simple
    A = 33600
    pred = synth.make_nms_predictions(2, A, 4, 1280, seed=330 + int(clustered), clustered=clustered, n_clusters=400)
    ka, kb, ks, kl, kc = nms_batched(pred.cuda(), 0.001, 0.7, 1280)
    for b in range(2):
        xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], 0.001)
        assert len(src) > 16384
        keep = c_oracle.nms(xyxy, sc, 0.7)
        k = int(kc[b])
        assert k == len(keep)
        assert np.array_equal(ka[b, :k].cpu().numpy(), src[keep])
        assert np.array_equal(ks[b, :k].cpu().numpy(), sc[keep])
        assert np.array_equal(kl[b, :k].cpu().numpy(), cl[keep])
        if b == 0:      # the single-image entry point on the same candidates (N > 16384)
            k1 = nms_simple(torch.from_numpy(xyxy).cuda(), torch.from_numpy(sc).cuda(), 0.7)
            assert k1 == keep
