"""SURVEY §8f row f1 host logic on CPU, world_size 2 over gloo: DistributedSampler shards, the DDP epoch loop,
all-rank loss means, and the rank-0 checkpoint in the reference's last.pt / best.pt format
(train_yolo11_cuda.py:31-98, :628-653).  The criterion here is a small differentiable stand-in with the reference's
return convention (scalar loss with grad, 3 detached items): the product criterion is CUDA-only and is exercised by
the gpu tests and examples/train_ddp.py; this test is about the plumbing around it."""
import math
import os
import sys
import tempfile

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class _Set(torch.utils.data.Dataset):
    def __init__(self, n):
        g = torch.Generator().manual_seed(3)
        self.img = torch.randn(n, 3, 8, 8, generator=g)

    def __len__(self):
        return self.img.shape[0]

    def __getitem__(self, i):
        return {"img": self.img[i], "idx": torch.tensor(i)}


def _collate(items):
    return {"img": torch.stack([it["img"] for it in items]), "idx": torch.stack([it["idx"] for it in items]), "note": "x"}


class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(11)
        self.conv = torch.nn.Conv2d(3, 4, 3, padding=1)
        self.bn = torch.nn.BatchNorm2d(4)

    def forward(self, x):
        return self.bn(self.conv(x))


def _criterion(seen):
    def crit(preds, batch):
        seen.extend(batch["idx"].tolist())
        B = preds.shape[0]
        parts = torch.stack([preds.pow(2).mean(), preds.abs().mean(), (preds - 1).pow(2).mean()])
        return parts.sum() * B, parts.detach()             # loss.sum() * batch_size, loss.detach()  (yolo_v8_loss.py:499)
    return crit


def _worker(rank, world, port, tmp, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(1)
    from yolo_scratch_b200 import ddp_train as T
    r, w, _, device = T.init_distributed("gloo")
    ok = (r, w, device.type) == (rank, world, "cpu")
    ds = _Set(10)
    loader, sampler = T.make_loader(ds, batch_size=2, collate_fn=_collate, shuffle=True, seed=5)
    model = T.wrap_model(_Net(), device)
    opt = torch.optim.SGD(model.parameters(), lr=0.01)
    seen = []
    m1 = T.train_one_epoch(model, loader, opt, _criterion(seen), device, 1, 2, sampler=sampler)
    ok &= set(m1) == {"loss", "box_loss", "cls_loss", "dfl_loss"} and all(math.isfinite(v) for v in m1.values())
    gathered = [None] * world
    flat = torch.cat([p.detach().flatten() for p in T.unwrap(model).parameters()])
    dist.all_gather_object(gathered, {"seen": seen, "m": m1, "p": flat})
    ok &= sorted(gathered[0]["seen"] + gathered[1]["seen"]) == list(range(10))        # disjoint shards that cover the set
    ok &= gathered[0]["m"] == gathered[1]["m"]                                         # every rank reports the job's mean
    ok &= torch.equal(gathered[0]["p"], gathered[1]["p"])                              # replicas stay in lock-step
    # second epoch reshuffles the shards
    seen2 = []
    T.train_one_epoch(model, loader, opt, _criterion(seen2), device, 2, 2, sampler=sampler)
    ok &= seen2 != seen
    # checkpoint: mAP50 rule, then the loss rule; rank 0 alone writes
    best = T.save_checkpoint(tmp, 0, model, opt, m1, {"loss": 3.0, "mAP50": 0.25}, float("inf"), 0.0)
    ok &= best == (float("inf"), 0.25)
    best = T.save_checkpoint(tmp, 1, model, opt, m1, {"loss": 2.0, "mAP50": 0.20}, *best)
    ok &= best == (float("inf"), 0.25)
    if rank == 0:
        last = torch.load(os.path.join(tmp, "last.pt"), weights_only=False)
        bestf = torch.load(os.path.join(tmp, "best.pt"), weights_only=False)
        ok &= tuple(last.keys()) == T.CHECKPOINT_KEYS == tuple(bestf.keys())
        ok &= last["epoch"] == 1 and bestf["epoch"] == 0 and bestf["best_mAP50"] == 0.25 and last["best_mAP50"] == 0.25
        ok &= not any(k.startswith("module.") for k in last["model_state_dict"])
        plain = _Net()
        plain.load_state_dict(last["model_state_dict"])                                  # loads into an unwrapped model
        now = torch.cat([p.detach().flatten() for p in T.unwrap(model).parameters()])
        ok &= torch.equal(torch.cat([p.detach().flatten() for p in plain.parameters()]), now)
        ck = T.load_checkpoint(os.path.join(tmp, "best.pt"), model, opt)
        ok &= ck["val_metrics"]["mAP50"] == 0.25
    dist.barrier()
    b2 = T.save_checkpoint(os.path.join(tmp, "by_loss"), 0, model, opt, m1, {"loss": 1.5}, 2.0, 0.0)
    ok &= b2 == (1.5, 0.0)
    if rank == 0:
        ok &= os.path.exists(os.path.join(tmp, "by_loss", "best.pt"))
    else:
        ok &= sorted(os.listdir(tmp)) == ["best.pt", "by_loss", "last.pt"]              # shared dir: nothing rank-1 specific
    res = [None] * world
    dist.all_gather_object(res, bool(ok))
    if rank == 0:
        q.put(all(res))
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_epoch_and_checkpoint_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    with tempfile.TemporaryDirectory() as tmp:
        procs = [ctx.Process(target=_worker, args=(r, 2, port, tmp, q)) for r in range(2)]
        for p in procs:
            p.start()
        ok = q.get(timeout=240)
        for p in procs:
            p.join(timeout=60)
        assert ok
        assert all(p.exitcode == 0 for p in procs)


def test_cosine_schedule_and_single_process_paths():
    from yolo_scratch_b200 import ddp_train as T
    net = _Net()
    opt = torch.optim.SGD(net.parameters(), lr=1.0)
    # warm-up is linear from lr_min, then half-cosine down to lr_min (train_yolo11_cuda.py:440-451)
    assert T.cosine_lr_schedule(opt, 0, 10, lr_min=1e-5, lr_max=1e-3, warmup_epochs=3) == 1e-5
    assert abs(T.cosine_lr_schedule(opt, 3, 10, lr_min=1e-5, lr_max=1e-3, warmup_epochs=3) - 1e-3) < 1e-12
    mid = T.cosine_lr_schedule(opt, 10, 10, lr_min=1e-5, lr_max=1e-3, warmup_epochs=3)
    assert abs(mid - 1e-5) < 1e-12 and opt.param_groups[0]["lr"] == mid
    # single process: no sampler, model returned unwrapped, metrics are the plain per-step means
    loader, sampler = T.make_loader(_Set(6), batch_size=2, collate_fn=_collate, shuffle=False)
    assert sampler is None and T.wrap_model(net, torch.device("cpu")) is net and T.is_main()
    seen = []
    m = T.train_one_epoch(net, loader, opt, _criterion(seen), torch.device("cpu"), 1, 1)
    assert seen == list(range(6)) and math.isfinite(m["loss"])
    with tempfile.TemporaryDirectory() as tmp:
        assert T.save_checkpoint(tmp, 0, net, opt, m, {"loss": 0.5}, 1.0, 0.0) == (0.5, 0.0)
        assert sorted(os.listdir(tmp)) == ["best.pt", "last.pt"]
        assert torch.load(os.path.join(tmp, "last.pt"), weights_only=False)["best_loss"] == 1.0   # written before the update (:639)
        assert torch.load(os.path.join(tmp, "best.pt"), weights_only=False)["best_loss"] == 0.5
