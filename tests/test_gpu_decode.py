"""GPU parity: inference decode (DFL + dist2bbox + sigmoid) vs the reference fixture and the oracle port,
1e-6 relative (softmax / conv summation order differs)."""
import pytest
import torch

from conftest import load_golden, t
from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu


def test_detect_inference_golden(cuda_lib):
    from yolo_scratch_b200.head_decode import detect_inference
    g = load_golden("detect_inference_s51")
    feats = [t(g["f0"]).cuda(), t(g["f1"]).cuda(), t(g["f2"]).cuda()]
    y = detect_inference(feats, int(g["nc"]))
    ref = t(g["y"])
    assert y.shape == ref.shape
    # boxes: 1e-6 of the image size (w,h are differences of two ~15-bin expectations times the stride)
    assert torch.allclose(y.cpu()[:, :4], ref[:, :4], rtol=2e-6, atol=1e-6 * 128)
    assert torch.allclose(y.cpu()[:, 4:], ref[:, 4:], rtol=2e-6, atol=1e-7)
    y2 = detect_inference(feats, int(g["nc"]), anchor_major=True)
    assert torch.equal(y2.permute(0, 2, 1).contiguous(), y)


def test_detect_inference_640_then_nms_end_to_end(cuda_lib):
    """levels -> decode (anchor-major) -> batched NMS equals oracle decode -> C-oracle NMS on the same scores."""
    import numpy as np
    from oracle import ref_port as orc, c_oracle
    from yolo_scratch_b200.head_decode import detect_inference
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    feats = synth.make_head_levels(2, 640, 80, 16, seed=400)
    for f in feats:
        f[:, 64:] += 3.0      # lift some class scores over the threshold
    y = detect_inference([f.cuda() for f in feats], 80, anchor_major=True)
    ref = orc.detect_inference(feats, 80, 16, synth.LEVEL_STRIDES)
    assert torch.allclose(y.cpu().permute(0, 2, 1), ref, rtol=2e-6, atol=1e-6 * 640)
    ka, kb, ks, kl, kc = nms_batched(y, 0.25, 0.45, 640)
    yc = y.cpu()
    for b in range(2):
        xyxy, sc, cl, src = c_oracle.filter_candidates(yc[b], 0.25)
        keep = c_oracle.nms(xyxy, sc, 0.45)
        assert np.array_equal(ka[b, :int(kc[b])].cpu().numpy(), src[keep])


@pytest.mark.parametrize("shapes,nc", [
    ([(84, 84), (42, 42), (21, 21)], 80),      # imgsz 672: P5 = 441 = 1 mod 4
    ([(9, 7), (5, 3), (3, 1)], 5),             # every level unaligned
    ([(20, 28), (10, 14), (5, 7)], 6),         # non-square head, 35 = 3 mod 4
    ([(16, 16), (8, 8), (4, 4)], 12),          # aligned levels, 4 + nc = 16: anchor-major rows staged with an odd pitch
    ([(24, 24), (12, 12), (6, 6)], 20),        # aligned levels, 4 + nc = 24 = 0 mod 8
    ([(13, 13), (7, 7), (4, 4)], 28),          # unaligned levels, 4 + nc = 32 = 0 mod 8: odd pitch
    ([(13, 11), (7, 5), (4, 3)], 8),           # unaligned levels, 4 + nc = 12 = 4 mod 8: 16-byte staging words
    ([(40, 40), (20, 20), (10, 10)], 3),       # nc not a multiple of 4: partial class group
])
def test_detect_inference_unaligned_levels(cuda_lib, shapes, nc):
    """Head shapes the float4 kernel cannot take (the reference's Detect.inference accepts any image size,
    models/yolo11_modules.py:248-266): the any-shape kernel, both output layouts, against the oracle port."""
    from oracle import ref_port as orc
    from yolo_scratch_b200.head_decode import detect_inference
    feats = synth.make_head_levels(2, 0, nc, 16, seed=410, shapes=shapes)
    ref = orc.detect_inference(feats, nc, 16, synth.LEVEL_STRIDES)
    dev = [f.cuda() for f in feats]
    y = detect_inference(dev, nc)
    scale = float(shapes[0][0] * 8)
    assert y.shape == ref.shape
    assert torch.allclose(y.cpu()[:, :4], ref[:, :4], rtol=2e-6, atol=1e-6 * scale)
    assert torch.allclose(y.cpu()[:, 4:], ref[:, 4:], rtol=2e-6, atol=1e-7)
    y2 = detect_inference(dev, nc, anchor_major=True)
    assert torch.equal(y2.permute(0, 2, 1).contiguous(), y)
    # a batch slice whose data pointer is not 16-byte aligned (contiguous view into a larger tensor)
    if (feats[2].shape[1] * feats[2].shape[2] * feats[2].shape[3]) % 4:
        y3 = detect_inference([f[1:] for f in dev], nc)
        assert torch.equal(y3, y[1:])
