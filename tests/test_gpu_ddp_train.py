"""GPU (row f1): ddp_train.train_one_epoch with the REAL criterion (yolo_scratch_b200.losses.v8DetectionLoss, the CUDA
path) on a tiny convolutional model and a synthetic loader: the epoch means it returns equal the oracle's loss on the
very head tensors and targets the criterion saw (1e-5), training moves the weights, and the rank-0 checkpoint in the
reference's last.pt / best.pt format (train_yolo11_cuda.py:628-653) round-trips into a fresh model.  World size 1 (the
multi-rank plumbing runs on CPU over gloo in test_ddp_train_gloo.py and on 2 GPUs in examples/train_ddp.py
--check-grads)."""
import os
import tempfile

import pytest
import torch

from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu
NC = 4


class Detect(torch.nn.Module):
    """Named `Detect` on purpose: the criterion finds the head by class name, as the reference does (:348-351)."""

    def __init__(self, ch, nc):
        super().__init__()
        self.nc, self.reg_max = nc, 16
        self.stride = torch.tensor([8.0, 16.0, 32.0])
        self.heads = torch.nn.ModuleList(torch.nn.Conv2d(c, 64 + nc, 1) for c in ch)

    def forward(self, xs):
        return [h(x) for h, x in zip(self.heads, xs)]


class TinyNet(torch.nn.Module):
    def __init__(self, nc=NC):
        super().__init__()
        torch.manual_seed(5)
        self.s8 = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, 2, 1), torch.nn.SiLU(), torch.nn.Conv2d(8, 8, 3, 2, 1),
                                      torch.nn.SiLU(), torch.nn.Conv2d(8, 16, 3, 2, 1), torch.nn.SiLU())
        self.s16 = torch.nn.Sequential(torch.nn.Conv2d(16, 16, 3, 2, 1), torch.nn.SiLU())
        self.s32 = torch.nn.Sequential(torch.nn.Conv2d(16, 16, 3, 2, 1), torch.nn.SiLU())
        self.detect = Detect((16, 16, 16), nc)

    def forward(self, x):
        a = self.s8(x)
        b = self.s16(a)
        c = self.s32(b)
        return self.detect([a, b, c])


class SynthSet(torch.utils.data.Dataset):
    """12 synthetic 96x96 images with 1-4 boxes each, in the reference's per-item format (collate below builds the
    batch dict of datasets/crater_dataset_cuda.py:327-341)."""

    def __init__(self, n=12, imgsz=96):
        g = torch.Generator().manual_seed(9)
        self.img = torch.rand(n, 3, imgsz, imgsz, generator=g)
        self.gt = [synth.make_gt_batch(1, NC, 4, seed=40 + i, imgsz=imgsz, p_empty=0.0) for i in range(n)]

    def __len__(self):
        return self.img.shape[0]

    def __getitem__(self, i):
        return {"img": self.img[i], "cls": self.gt[i]["cls"], "bboxes": self.gt[i]["bboxes"]}


def collate(items):
    return {"img": torch.stack([it["img"] for it in items]),
            "batch_idx": torch.cat([torch.full((it["cls"].shape[0],), k, dtype=torch.long) for k, it in enumerate(items)]),
            "cls": torch.cat([it["cls"] for it in items]), "bboxes": torch.cat([it["bboxes"] for it in items])}


def test_train_one_epoch_with_cuda_criterion_vs_oracle(cuda_lib):
    from oracle import ref_port as orc
    from yolo_scratch_b200 import ddp_train as T
    from yolo_scratch_b200.losses import v8DetectionLoss
    rank, world, _, device = T.init_distributed()
    assert world == 1 and device.type == "cuda"
    model = T.wrap_model(TinyNet(), device)
    crit = v8DetectionLoss(model)
    seen = []

    def recording_criterion(preds, batch):       # the product criterion, with its inputs kept for the oracle
        seen.append(([p.detach().cpu() for p in preds], {k: batch[k].detach().cpu() for k in ("batch_idx", "cls", "bboxes")}))
        return crit(preds, batch)

    loader, sampler = T.make_loader(SynthSet(), batch_size=4, collate_fn=collate, shuffle=True, seed=1)
    assert sampler is None
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    before = [p.detach().clone() for p in model.parameters()]
    T.cosine_lr_schedule(opt, 0, 2)
    m0 = T.train_one_epoch(model, loader, opt, recording_criterion, device, 0, 2, sampler=sampler)
    assert len(seen) == 3
    # the oracle on the head tensors / targets of every step
    tot = torch.zeros(4, dtype=torch.float64)
    for preds, batch in seen:
        total, items = orc.detection_loss(preds, batch, NC, 16, synth.LEVEL_STRIDES)
        tot[0] += float(total)
        tot[1:] += items.double()
    tot /= len(seen)
    got = torch.tensor([m0["loss"], m0["box_loss"], m0["cls_loss"], m0["dfl_loss"]], dtype=torch.float64)
    assert torch.allclose(got, tot, rtol=1e-5), (got, tot)
    assert any(not torch.equal(a, b.detach()) for a, b in zip(before, model.parameters())), "the epoch must train"
    T.cosine_lr_schedule(opt, 1, 2)
    m1 = T.train_one_epoch(model, loader, opt, recording_criterion, device, 1, 2, sampler=sampler)
    assert all(torch.isfinite(torch.tensor(list(m1.values()))))

    # checkpoint: the reference's dict, loadable into a fresh, unwrapped model
    with tempfile.TemporaryDirectory() as tmp:
        best_loss, best_map = T.save_checkpoint(tmp, 1, model, opt, m1, {"loss": m1["loss"]}, float("inf"), 0.0)
        assert best_loss == m1["loss"] and os.path.exists(os.path.join(tmp, "last.pt")) and os.path.exists(os.path.join(tmp, "best.pt"))
        ck = torch.load(os.path.join(tmp, "last.pt"), map_location="cpu", weights_only=False)
        assert tuple(ck.keys()) == T.CHECKPOINT_KEYS and ck["epoch"] == 1
        fresh = TinyNet()
        T.load_checkpoint(os.path.join(tmp, "best.pt"), fresh)
        for (k, a), (_, b) in zip(fresh.state_dict().items(), T.unwrap(model).state_dict().items()):
            assert torch.equal(a, b.cpu()), k
        # a second save with a worse loss keeps best.pt
        t_best = os.path.getmtime(os.path.join(tmp, "best.pt"))
        bl2, _ = T.save_checkpoint(tmp, 2, model, opt, m1, {"loss": m1["loss"] + 1.0}, best_loss, best_map)
        assert bl2 == best_loss and os.path.getmtime(os.path.join(tmp, "best.pt")) == t_best
