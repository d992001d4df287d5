"""Debug helper: where does the CUDA loss gradient differ most from the oracle port?  (GPU box)
    python scripts/debug_grad_diff.py SEED B IMGSZ FIXED_COUNT"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import ref_port as orc
from yolo_scratch_b200 import synth
from yolo_scratch_b200.losses import v8DetectionLoss

seed, B, imgsz, cnt = (int(x) for x in sys.argv[1:5])
nc = 80


class Detect(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.nc, self.reg_max = nc, 16
        self.stride = torch.tensor([8.0, 16.0, 32.0])
        self.p = torch.nn.Parameter(torch.zeros(1))


feats = synth.make_head_levels(B, imgsz, nc, 16, seed=seed)
batch = synth.make_gt_batch(B, nc, cnt, seed=seed + 1, imgsz=imgsz, fixed_count=cnt, dense=True)
crit = v8DetectionLoss(Detect().cuda())
crit.export_assignment = True
leaves = [f.cuda().requires_grad_(True) for f in feats]
total, items = crit(leaves, {k: v.cuda() for k, v in batch.items()})
total.backward()
torch.cuda.synchronize()
ref = [f.clone().requires_grad_(True) for f in feats]
rt, ri, parts = orc.detection_loss(ref, batch, nc, 16, synth.LEVEL_STRIDES, return_parts=True)
rt.backward()
print("items", items.tolist(), ri.tolist())
ex = {k: v.cpu() for k, v in crit.last_export.items()}
off = 0
for l, (a, b) in enumerate(zip(leaves, ref)):
    d = (a.grad.cpu() - b.grad).abs()
    scale = b.grad.abs().max().item()
    hw = a.shape[2] * a.shape[3]
    print(f"level {l}: max err {d.max().item():.3e} rel {d.max().item() / scale:.3e}")
    flat = d.view(B, -1, hw)
    top = torch.topk(flat.flatten(), 12)
    for v, idx in zip(top.values.tolist(), top.indices.tolist()):
        bi, rem = divmod(idx, flat.shape[1] * hw)
        c, i = divmod(rem, hw)
        an = off + i
        print(f"   b={bi} ch={c} i={i} err={v:.3e} ours={a.grad[bi].view(-1, hw)[c, i].item():.6e} ref={b.grad[bi].view(-1, hw)[c, i].item():.6e} "
              f"tgi={ex['target_gt_idx'][bi, an].item()} ref_tgi={parts['target_gt_idx'][bi, an].item()} fg={parts['fg_mask'][bi, an].item()} "
              f"norm={ex['norm'][bi, an].item():.5e} refnorm={parts['norm'][bi, an].item():.5e}")
    off += hw
