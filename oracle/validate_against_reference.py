"""ORACLE (test infrastructure).  Build-container-only check: the torch port (oracle/ref_port.py) and the C scalar
spec (oracle/scalar_spec.c) against the LIVE, unmodified reference imported from /root/reference, on randomized
seeded inputs beyond the committed fixtures.  Exits 0 and prints a summary; skips if the reference is absent.

    python oracle/validate_against_reference.py [n_seeds]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/yolo_scratch_cuda"
if not os.path.isdir(REF):
    print("reference not present: nothing to validate here")
    sys.exit(0)
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from losses.yolo_v8_loss import TaskAlignedAssigner, v8DetectionLoss          # noqa: E402 (reference)
from train_yolo11_cuda import decode_predictions_for_metrics                 # noqa: E402 (reference)

from oracle import ref_port as orc, c_oracle                                 # noqa: E402
from yolo_scratch_b200 import synth                                          # noqa: E402


class Detect(torch.nn.Module):
    def __init__(self, nc):
        super().__init__()
        self.nc, self.reg_max = nc, 16
        self.stride = torch.tensor([8.0, 16.0, 32.0])
        self.p = torch.nn.Parameter(torch.zeros(1))


def main(n_seeds):
    torch.set_num_threads(8)
    worst = {"assign_ts": 0.0, "loss_items": 0.0, "grad": 0.0}
    n_forced = 0
    for s in range(n_seeds):
        shapes = ((16, 16), (8, 8), (4, 4)) if s % 3 else ((40, 40), (20, 20), (10, 10))
        nc = 4 + s % 7
        kw = {"p_tiny": 0.1 * (s % 5), "p_nested": 0.1 * (s % 4), "p_empty": 0.1 * (s % 3)}
        inp = synth.make_assigner_inputs(2 + s % 3, shapes, synth.LEVEL_STRIDES, nc, 1 + (s * 7) % 30, 7000 + s, **kw)
        if inp[4].shape[1] == 0:
            continue
        ref = TaskAlignedAssigner(topk=50, num_classes=nc, alpha=0.5, beta=4.0).forward(*[t.clone() for t in inp])
        got = orc.assign(*inp, nc)
        assert torch.equal(got[3], ref[3]) and torch.equal(got[4], ref[4]), f"assign indices seed {s}"
        assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1])
        worst["assign_ts"] = max(worst["assign_ts"], (got[2] - ref[2]).abs().max().item())
        c = c_oracle.assign(*inp, nc)
        assert torch.equal(c["fg_mask"], ref[3]) and torch.equal(c["target_gt_idx"].long(), ref[4]), f"C assign seed {s}"
        inside_any = (inp[5].sum() > 0)
        n_forced += int(inside_any)
        # loss
        imgsz = shapes[0][0] * 8
        feats = synth.make_head_levels(2 + s % 2, imgsz, nc, 16, seed=8000 + s)
        batch = synth.make_gt_batch(2 + s % 2, nc, 1 + (s * 5) % 12, seed=8100 + s, imgsz=imgsz, **kw)
        la = [f.clone().requires_grad_(True) for f in feats]
        lb = [f.clone().requires_grad_(True) for f in feats]
        t_ref, i_ref = v8DetectionLoss(Detect(nc))(la, {k: v.clone() for k, v in batch.items()})
        t_ref.backward()
        t_got, i_got = orc.detection_loss(lb, batch, nc, 16, synth.LEVEL_STRIDES)
        t_got.backward()
        worst["loss_items"] = max(worst["loss_items"], ((i_got - i_ref).abs() / i_ref.abs().clamp(min=1e-6)).max().item())
        for a, b in zip(la, lb):
            sc = a.grad.abs().max().item() or 1.0
            worst["grad"] = max(worst["grad"], (a.grad - b.grad).abs().max().item() / sc)
        # NMS
        pred = synth.make_nms_predictions(1, 300 + 50 * s, nc, 640, seed=9000 + s, clustered=bool(s % 2))
        thr, conf = (0.7, 0.001) if s % 2 else (0.45, 0.25)
        r = decode_predictions_for_metrics(pred.clone(), 640, conf, thr, torch.device("cpu"))[0]
        g = orc.decode_for_metrics(pred, 640, conf, thr, return_keep=True)[0]
        assert torch.equal(g["boxes"], r["boxes"]) and torch.equal(g["scores"], r["scores"]) and torch.equal(g["labels"], r["labels"])
        xyxy, sc_, cl, src = c_oracle.filter_candidates(pred[0], conf)
        keep = c_oracle.nms(xyxy, sc_, thr)
        assert np.array_equal(src[keep], g["keep"].numpy())
    print(f"validated {n_seeds} seeds against the live reference: indices/keep lists bit-exact; worst diffs {worst}")
    assert worst["loss_items"] < 2e-6 and worst["grad"] < 2e-6 and worst["assign_ts"] < 1e-6


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 24)
