"""CPU, world_size 2 over gloo: the N>1 host logic.  Each rank takes its image shard (sharding.shard_batch),
evaluates the path on it (the ORACLE stands in for the kernels here -- no GPU in this container), and the
results are gathered: per-shard assignments equal the corresponding rows of the single-process run, and the
image-weighted loss items all-reduce is consistent.  This is the property the multi-GPU bench relies on
(no data-path collective)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    from oracle import ref_port as orc
    from yolo_scratch_b200 import synth
    from yolo_scratch_b200.sharding import shard_batch, shard_bounds, allreduce_loss_items
    B, nc = 5, 8                                             # odd batch: uneven shards
    feats = synth.make_head_levels(B, 128, nc, 16, seed=500)
    batch = synth.make_gt_batch(B, nc, 6, seed=501, imgsz=128, p_empty=0.2, p_tiny=0.3)
    f_r, b_r = shard_batch(feats, batch, rank, world)
    lo, hi = shard_bounds(B, rank, world)
    total, items, parts = orc.detection_loss(f_r, b_r, nc, 16, synth.LEVEL_STRIDES, return_parts=True)
    mean_items = allreduce_loss_items(items, hi - lo)
    gathered = [None] * world
    dist.all_gather_object(gathered, {"lo": lo, "hi": hi, "tgi": parts["target_gt_idx"], "fg": parts["fg_mask"],
                                      "items": items, "mean": mean_items})
    if rank == 0:
        _, _, full = orc.detection_loss(feats, batch, nc, 16, synth.LEVEL_STRIDES, return_parts=True)
        ok = True
        cover = []
        for g in gathered:
            ok &= torch.equal(full["target_gt_idx"][g["lo"]:g["hi"]], g["tgi"])
            ok &= torch.equal(full["fg_mask"][g["lo"]:g["hi"]], g["fg"])
            cover += list(range(g["lo"], g["hi"]))
        ok &= cover == list(range(B))
        w = torch.tensor([g["hi"] - g["lo"] for g in gathered], dtype=torch.float64)
        ref_mean = sum(g["items"].double() * wi for g, wi in zip(gathered, w)) / w.sum()
        ok &= all(torch.allclose(g["mean"].double(), ref_mean, rtol=1e-6) for g in gathered)
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_matches_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_shard_bounds_cover():
    from yolo_scratch_b200.sharding import shard_bounds
    for B in (1, 5, 64, 67):
        for W in (1, 2, 4, 8):
            spans = [shard_bounds(B, r, W) for r in range(W)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
