#!/usr/bin/env python
"""Randomised parity sweep on the GPU box (not part of the test suite): random head shapes, class counts, batch sizes and
GT densities through the fused loss (spliced-oracle check of tests/test_gpu_loss.py: decoded boxes, bit-exact indices,
items, gradients), random candidate sets through the batched NMS (keep lists vs the C oracle), random heads through the
inference decode (both layouts vs the oracle port) and random dense inputs through TaskAlignedAssigner (default, use_topk and
ciou_assign modes vs the oracle port: indices bit-exact).
    python scripts/fuzz_parity.py [--seed 0] [--loss 40] [--nms 40] [--decode 30]"""
import argparse
import json
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from yolo_scratch_b200 import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--loss", type=int, default=40)
    ap.add_argument("--nms", type=int, default=40)
    ap.add_argument("--decode", type=int, default=30)
    ap.add_argument("--assign", type=int, default=30)
    a = ap.parse_args()
    rng = random.Random(a.seed)
    import test_gpu_loss as tl
    from oracle import c_oracle, ref_port as orc
    from yolo_scratch_b200.head_decode import detect_inference
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    out = {"seed": a.seed, "loss": [], "nms": [], "decode": [], "assign": []}
    t0 = time.time()
    for k in range(a.loss):
        h0 = rng.choice([4, 8, 9, 12, 13, 16, 20, 21, 28, 40])           # P3 side; P4 = ceil/2, P5 = ceil/4 (odd sizes included)
        w0 = h0 if rng.random() < 0.7 else rng.choice([4, 8, 12, 20, 28])
        shapes = [(h0, w0), ((h0 + 1) // 2, (w0 + 1) // 2), ((h0 + 3) // 4, (w0 + 3) // 4)]
        nc = rng.choice([1, 2, 3, 5, 8, 13, 20, 37, 80, 91])
        B = rng.randint(1, 5)
        max_gt = rng.choice([1, 3, 8, 20, 45, 70, 150])
        kw = {"p_empty": rng.choice([0.0, 0.2, 0.5]), "p_tiny": rng.choice([0.0, 0.1, 0.4]), "p_nested": rng.choice([0.0, 0.1, 0.4])}
        if rng.random() < 0.3:
            kw["fixed_count"] = max_gt
            kw["dense"] = rng.random() < 0.5
        seed = 10_000 + 17 * k + a.seed * 1000
        imgsz = max(h0, w0) * 8
        feats = synth.make_head_levels(B, 0, nc, 16, seed=seed, shapes=shapes)
        batch = synth.make_gt_batch(B, nc, max_gt, seed=seed + 1, imgsz=imgsz, **kw)
        case = {"shapes": shapes, "nc": nc, "B": B, "max_gt": max_gt, "kw": kw, "seed": seed, "n_gt": int(batch["cls"].shape[0])}
        try:
            tl.spliced_check(feats, batch, B, imgsz, nc)
            case["ok"] = True
        except Exception as ex:          # keep sweeping: the report lists every failing case
            case["ok"] = False
            case["error"] = repr(ex)[:300]
        out["loss"].append(case)
    for k in range(a.nms):
        A = rng.choice([1, 7, 33, 100, 600, 1500, 4000, 8400, 9000])
        nc = rng.choice([1, 3, 20, 80])
        Bn = rng.randint(1, 4)
        conf = rng.choice([0.001, 0.05, 0.25])
        thr = rng.choice([0.1, 0.45, 0.7, 0.9])
        clustered = rng.random() < 0.4
        class_aware = rng.random() < 0.3
        seed = 20_000 + 13 * k + a.seed * 1000
        pred = synth.make_nms_predictions(Bn, A, nc, 640, seed=seed, clustered=clustered)
        case = {"A": A, "nc": nc, "B": Bn, "conf": conf, "thr": thr, "clustered": clustered, "class_aware": class_aware, "seed": seed}
        try:
            ka, kb, ks, kl, kc = nms_batched(pred.cuda(), conf, thr, 640, class_aware=class_aware)
            for b in range(Bn):
                xyxy, sc, cl, src = c_oracle.filter_candidates(pred[b], conf)
                keep = c_oracle.nms(xyxy, sc, thr, labels=cl if class_aware else None)
                got = ka[b, :int(kc[b])].cpu().numpy()
                assert np.array_equal(got, src[keep]), f"image {b}: {len(got)} kept vs {len(keep)}"
            case["ok"] = True
        except Exception as ex:
            case["ok"] = False
            case["error"] = repr(ex)[:300]
        out["nms"].append(case)
    for k in range(a.decode):
        h0 = rng.choice([3, 4, 7, 8, 12, 13, 20, 21, 40, 80])
        w0 = h0 if rng.random() < 0.7 else rng.choice([3, 5, 8, 20])
        shapes = [(h0, w0), ((h0 + 1) // 2, (w0 + 1) // 2), ((h0 + 3) // 4, (w0 + 3) // 4)]
        nc = rng.choice([1, 2, 3, 4, 8, 12, 20, 28, 37, 80, 91])
        B = rng.randint(1, 4)
        seed = 30_000 + 11 * k + a.seed * 1000
        feats = synth.make_head_levels(B, 0, nc, 16, seed=seed, shapes=shapes)
        case = {"shapes": shapes, "nc": nc, "B": B, "seed": seed}
        try:
            ref = orc.detect_inference(feats, nc, 16, synth.LEVEL_STRIDES)
            dev = [f.cuda() for f in feats]
            y = detect_inference(dev, nc)
            y2 = detect_inference(dev, nc, anchor_major=True)
            # decoded extents reach 15 cells x stride 32 = 480 px whatever the image size, and centres are differences of
            # two such numbers: the absolute tolerance follows that magnitude, not the (possibly tiny) image
            scale = max(float(max(h0, w0) * 8), 16.0 * 32.0)
            assert torch.allclose(y.cpu()[:, :4], ref[:, :4], rtol=2e-6, atol=1e-6 * scale)
            assert torch.allclose(y.cpu()[:, 4:], ref[:, 4:], rtol=2e-6, atol=1e-7)
            assert torch.equal(y2.permute(0, 2, 1).contiguous(), y)
            case["ok"] = True
        except Exception as ex:
            case["ok"] = False
            case["error"] = repr(ex)[:300]
        out["decode"].append(case)
    import test_gpu_assign as ta
    from yolo_scratch_b200.losses import TaskAlignedAssigner
    for k in range(a.assign):
        h0 = rng.choice([4, 8, 12, 20, 40])
        shapes = ((h0, h0), ((h0 + 1) // 2, (h0 + 1) // 2), ((h0 + 3) // 4, (h0 + 3) // 4))
        nc = rng.choice([1, 3, 8, 20, 80])
        B = rng.randint(1, 4)
        max_gt = rng.choice([1, 4, 20, 60, 130])
        topk = rng.choice([0, 0, 5, 13])
        ciou = rng.random() < 0.3
        kw = {"p_empty": rng.choice([0.0, 0.3]), "p_tiny": rng.choice([0.0, 0.3]), "p_nested": rng.choice([0.0, 0.3])}
        seed = 40_000 + 7 * k + a.seed * 1000
        case = {"shapes": shapes, "nc": nc, "B": B, "max_gt": max_gt, "topk": topk, "ciou": ciou, "kw": kw, "seed": seed}
        try:
            inp = synth.make_assigner_inputs(B, shapes, synth.LEVEL_STRIDES, nc, max_gt, seed, **kw)
            asg = TaskAlignedAssigner(topk=topk or 50, num_classes=nc, alpha=0.5, beta=4.0, use_topk=topk > 0, ciou_assign=ciou)
            got = [o.cpu() for o in asg.forward(*[x.cuda() for x in inp])]
            ref = orc.assign(*inp, nc, 0.5, 4.0, ciou_assign=ciou, topk=topk)
            if inp[3].shape[1] > 0:
                ta._check(got, ref)
            case["ok"] = True
        except Exception as ex:
            case["ok"] = False
            case["error"] = repr(ex)[:300]
        out["assign"].append(case)
    torch.cuda.synchronize()
    summary = {k: {"cases": len(v), "failed": sum(1 for c in v if not c["ok"])} for k, v in out.items() if isinstance(v, list)}
    out["summary"] = summary
    out["seconds"] = time.time() - t0
    print(json.dumps(out))
    print(json.dumps(summary), file=sys.stderr)
    sys.exit(1 if any(s["failed"] for s in summary.values()) else 0)


if __name__ == "__main__":
    main()
