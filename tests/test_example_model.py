"""The caller-side YOLO11 of examples/ (used by the configs[2] DDP step) has the reference model's size and outputs:
parameter counts probed from /root/reference/yolo_scratch_cuda/models/yolo11_model.py build_yolo11(ch=3, nc=80) in the
build container -- n: 3 217 808, s: 11 397 344, each including the reference DFL's 16 frozen conv weights
(models/yolo11_modules.py:183-188), which the example head does not carry."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples"))


@pytest.mark.parametrize("scale,ref_params", [("n", 3217808), ("s", 11397344)])
def test_param_count_and_head_shapes(scale, ref_params):
    from yolo11_torch import YOLO11
    m = YOLO11(scale, 3, 80).train()
    assert sum(p.numel() for p in m.parameters()) == ref_params - 16
    out = m(torch.zeros(2, 3, 64, 64))
    assert [tuple(o.shape) for o in out] == [(2, 144, 8, 8), (2, 144, 4, 4), (2, 144, 2, 2)]
    head = [x for x in m.modules() if type(x).__name__ == "Detect"]
    assert len(head) == 1 and head[0].nc == 80 and head[0].reg_max == 16 and head[0].stride.tolist() == [8.0, 16.0, 32.0]


@pytest.mark.gpu
def test_train_step_runs_and_updates_weights():
    """One full step on the GPU: backbone forward, CUDA loss path, backward into the backbone, optimiser step."""
    from yolo11_torch import YOLO11
    from yolo_scratch_b200 import synth
    from yolo_scratch_b200.losses import v8DetectionLoss
    torch.manual_seed(0)
    m = YOLO11("n", 3, 80).cuda().train()
    crit = v8DetectionLoss(m)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
    img = torch.rand(4, 3, 256, 256, device="cuda")
    batch = {k: v.cuda() for k, v in synth.make_gt_batch(4, 80, 10, seed=5, p_empty=0.0).items()}
    w0 = m.head.cls[0][2].weight.detach().clone()
    losses = []
    for _ in range(3):
        opt.zero_grad(set_to_none=True)
        loss, items = crit(m(img), batch)
        loss.backward()
        gn = torch.nn.utils.clip_grad_norm_(m.parameters(), 10.0)
        opt.step()
        assert torch.isfinite(items).all() and torch.isfinite(gn)
        losses.append(float(loss))
    assert not torch.equal(w0, m.head.cls[0][2].weight.detach())
    assert all(p.grad is not None for p in m.parameters() if p.requires_grad)
