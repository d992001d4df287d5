#!/usr/bin/env python
"""bench.py -- headline benchmark of the detection-head hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload at every N (weak scaling, one process per GPU, no data-path collective -- the path shards by
image): BASELINE.json configs[1] = YOLO11n assign + DFL/CIoU/BCE loss forward+backward,
batch 64 per GPU, 640 px (A=8400), 80 classes, synthetic COCO-shaped GT (<=20 per image).

One JSON line on rank 0:
  value      images/s of the whole job; a step = one ys_detection_loss C-ABI call (3 kernels, forward AND
             backward, gradients written) on inputs resident in HBM, timed with CUDA events, max over ranks.
  e2e        the same metric through the drop-in Python API (v8DetectionLoss.__call__ + backward) with the
             head tensors and targets coming from pinned HOST memory each step and the loss items read back.
  roofline   the dominant kernel (k_loss_stream) timed in place: the step launched stage by stage with a CUDA
             event between the kernels;
             algorithmic bytes (2 * A * (4*reg_max+nc) * 4 per image) / CUDA-event time vs MEASURED_PEAKS.json.
  cpu_baseline  the oracle port of the reference's PyTorch path (oracle/ref_port.py) on the host cores.
  nms        side measurement of configs[3] (batched NMS, B=256, conf 0.001 / IoU 0.7), N=1 only.
`--impl reference` prints the CPU reference arm: the oracle port timed on all host cores (the reference is
pure Python and /root/reference does not exist on the GPU box).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from yolo_scratch_b200 import synth  # noqa: E402

NC, REG_MAX, IMGSZ, B_PER_GPU, MAX_GT = 80, 16, 640, 64, 20
A = sum((IMGSZ // s) ** 2 for s in synth.LEVEL_STRIDES)
ALG_BYTES_PER_IMG = 2 * A * (4 * REG_MAX + NC) * 4          # read every head value, write every gradient
METRIC = "YOLO11 640px assign+loss+NMS images/sec (1/2/4/8 B200) vs CPU ref; HBM GB/s"


class Detect(torch.nn.Module):
    def __init__(self, nc):
        super().__init__()
        self.nc, self.reg_max = nc, REG_MAX
        self.stride = torch.tensor([float(s) for s in synth.LEVEL_STRIDES])
        self.p = torch.nn.Parameter(torch.zeros(1))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons = [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        hot = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(hot) if hot else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_inputs(seed):
    feats = synth.make_head_levels(B_PER_GPU, IMGSZ, NC, REG_MAX, seed=seed)
    batch = synth.make_gt_batch(B_PER_GPU, NC, MAX_GT, seed=seed + 1)
    return feats, batch


# --------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference(steps, warmup, budget_s=None):
    """The reference's PyTorch path restated (oracle/ref_port.py), all host threads, same config."""
    from oracle import ref_port as orc
    torch.set_num_threads(os.cpu_count() or 1)
    feats, batch = make_inputs(1000)

    def step():
        leaves = [f.requires_grad_(True) for f in (x.detach() for x in feats)]
        total, items = orc.detection_loss(leaves, batch, NC, REG_MAX, synth.LEVEL_STRIDES)
        total.backward()
        return items

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return {"value": B_PER_GPU * done / dt, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{done} step(s) of B={B_PER_GPU} assign+loss fwd+bwd (oracle/ref_port.py, torch CPU, "
                      f"os.cpu_count()={os.cpu_count()})", "ms_per_step": 1e3 * dt / done}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference(args.steps, min(args.warmup, 1), budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": r["value"], "unit": "images/s", "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(n):
    return {"workload": "configs[1]: YOLO11n assign + DFL/CIoU/BCE loss fwd+bwd, B=64 per GPU, 640px (A=8400), "
                        "nc=80, <=20 GT/img, synthetic COCO-shaped GT",
            "batch_per_gpu": B_PER_GPU, "global_batch": B_PER_GPU * n, "anchors": A, "nc": NC, "reg_max": REG_MAX,
            "parallelism": f"image-sharded x{n}, no data-path collective",
            "l2_policy": "inputs+gradients per step = 619 MB > 126 MB L2 (no explicit flush needed)"}


# --------------------------------------------------------------------------------------- B200 arm
def run_b200(args):
    from yolo_scratch_b200 import _lib
    from yolo_scratch_b200.losses import v8DetectionLoss
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback for the product path)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()

    feats_h, batch_h = make_inputs(1000 + 17 * rank)
    feats_h = [f.pin_memory() for f in feats_h]
    batch_h = {k: v.pin_memory() for k, v in batch_h.items()}
    feats = [f.to(dev) for f in feats_h]
    batch = {k: v.to(dev) for k, v in batch_h.items()}
    grads = [torch.empty_like(f) for f in feats]
    n = len(feats)
    n_gt = int(batch["cls"].shape[0])
    xp = (C.c_void_p * n)(*[f.data_ptr() for f in feats])
    gp = (C.c_void_p * n)(*[g.data_ptr() for g in grads])
    hs = (C.c_int * n)(*[f.shape[2] for f in feats])
    ws_ = (C.c_int * n)(*[f.shape[3] for f in feats])
    st = (C.c_float * n)(*[float(s) for s in synth.LEVEL_STRIDES])
    bi = batch["batch_idx"].contiguous()
    cl = batch["cls"].contiguous().view(-1)
    bb = batch["bboxes"].contiguous()
    out = torch.zeros(4, device=dev)
    wbytes = lib.ys_detection_loss_workspace_bytes(B_PER_GPU, A, n_gt, NC)
    wsb = torch.zeros(wbytes + 4096, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def call(stages=0):
        rc = lib.ys_detection_loss(n, xp, gp, hs, ws_, st, B_PER_GPU, NC, REG_MAX, bi.data_ptr(), cl.data_ptr(),
                                   bb.data_ptr(), n_gt, float(IMGSZ), float(IMGSZ), 7.5, 0.5, 1.5, 0.5, 4.0,
                                   out.data_ptr(), None, None, None, stages, wsb.data_ptr(), wsb.numel(),
                                   stream.cuda_stream)
        _lib.check(rc, "ys_detection_loss")

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value: K steps of the C-ABI call on resident inputs
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        call()
    # clock settle: the step is ~0.2 ms, so W steps alone neither bring the SM clock up from idle nor give
    # nvidia-smi (200 ms period) a sample under load; ~0.6 s of extra untimed steps do both
    t_settle = time.perf_counter()
    while time.perf_counter() - t_settle < 0.6:
        for _ in range(50):
            call()
        torch.cuda.synchronize(dev)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        call()
    e1.record(stream)
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    value = world * B_PER_GPU * args.steps / (ms_total * 1e-3)
    loss_vec = out.tolist()

    # ---- roofline: the same step, launched stage by stage with a CUDA event between the kernels, so that every
    #      kernel is timed where it runs in the pipeline (a kernel re-launched back to back with itself starts on an
    #      L2 full of its own dirty gradient lines and reads ~15 % slower than it does after k_assign).
    reps = max(args.steps, 20)
    names = (("k_pack_gt", 1), ("k_assign", 2), ("k_loss_stream", 4))
    for _ in range(3):
        for _, bit in names:
            call(bit)
    torch.cuda.synchronize(dev)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(reps)]
    for r in range(reps):
        evs[r][0].record(stream)
        for k, (_, bit) in enumerate(names):
            call(bit)
            evs[r][k + 1].record(stream)
    torch.cuda.synchronize(dev)
    stage_ms = {name: sum(evs[r][k].elapsed_time(evs[r][k + 1]) for r in range(reps)) / reps
                for k, (name, _) in enumerate(names)}
    loss_ms = stage_ms["k_loss_stream"]
    peak, peak_src = peaks()
    achieved = ALG_BYTES_PER_IMG * B_PER_GPU / (loss_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("k_loss_stream_dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "k_loss_stream", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms": loss_ms, "algorithmic_bytes_per_launch": ALG_BYTES_PER_IMG * B_PER_GPU,
                "whole_step_achieved_gbs_per_gpu": ALG_BYTES_PER_IMG * B_PER_GPU * args.steps / (ms_total * 1e-3) / 1e9,
                "stage_ms": stage_ms}

    # ---- e2e: drop-in Python API, inputs from pinned host memory every step, loss items read back
    crit = v8DetectionLoss(Detect(NC).to(dev))
    e2e_steps = max(3, min(args.steps, 20))

    def e2e_step():
        leaves = [f.to(dev, non_blocking=True).requires_grad_(True) for f in feats_h]
        bt = {k: v.to(dev, non_blocking=True) for k, v in batch_h.items()}
        total, items = crit(leaves, bt)
        total.backward()
        return items.cpu(), total.item()

    for _ in range(2):
        e2e_step()
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    for _ in range(e2e_steps):
        items_host, _ = e2e_step()
    g1.record(stream)
    barrier()
    e2e_ms = max_over_ranks(g0.elapsed_time(g1))
    clocks = sampler.stop()
    h2d = sum(f.numel() * 4 for f in feats_h) + sum(v.numel() * v.element_size() for v in batch_h.values())
    e2e = {"value": world * B_PER_GPU * e2e_steps / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": 16, "steps": e2e_steps,
           "api": "v8DetectionLoss.__call__ + backward (drop-in Python API over the C ABI)"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world), "clocks": clocks, "e2e": e2e, "gpu_launches": 3 * args.steps,
            "roofline": roofline, "loss_items": loss_vec}

    if world == 1 and not args.no_cpu:
        cb = cpu_reference(steps=12, warmup=1, budget_s=12.0)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if world == 1 and not args.no_nms:
        try:
            line["nms"] = bench_nms(dev, lib)
        except Exception as ex:  # the side measurement must not take the headline down
            line["nms"] = {"error": repr(ex)}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def bench_nms(dev, lib, B=256, reps=5):
    """configs[3]: batched NMS, 8400 candidates x 80 classes, conf 0.001 / IoU 0.7, batch 256."""
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    base = synth.make_nms_predictions(16, A, NC, IMGSZ, seed=2000)
    pred = base.repeat(B // 16, 1, 1).to(dev)
    pred[:, :, :2] += torch.arange(B, device=dev).view(B, 1, 1) * 0.03125      # distinct images, same statistics
    for _ in range(2):
        out = nms_batched(pred, 0.001, 0.7, IMGSZ)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = nms_batched(pred, 0.001, 0.7, IMGSZ)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    kept = out[4].float().mean().item()
    alg = A * (4 + NC) * 4 * B
    pairs = float((out[4].double() * (A - out[4].double() / 2)).sum().item())      # kept rows x later columns, approx.
    return {"workload": "configs[3]: B=256, 8400 candidates x 80 classes, conf 0.001 / IoU 0.7, class-agnostic",
            "images_per_s": B / (ms * 1e-3), "ms": ms, "kept_per_image": kept,
            "algorithmic_gbs": alg / (ms * 1e-3) / 1e9, "approx_pair_tests_per_s": pairs / (ms * 1e-3),
            "note": "not HBM bound (SURVEY §8d caveat): greedy tile chain, barrier/latency bound after the spatial culling of "
                    "pair tests (DESIGN.md §4 NMS); pair figure = kept rows x later columns the reference would test"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-nms", action="store_true", help="skip the NMS side measurement")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
