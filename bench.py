#!/usr/bin/env python
"""bench.py -- headline benchmark of the detection-head hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c2|c5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload at every N (weak scaling, one process per GPU, no data-path collective -- the path shards by image, SURVEY
§8e).  Default `--config c2` = BASELINE.json configs[1]: YOLO11n assign + DFL/CIoU/BCE loss forward+backward, batch 64
per GPU, 640 px (A=8400), 80 classes, synthetic COCO-shaped GT (<=20 per image).  `--config c5` = configs[4]: 1280 px
(A=33600), exactly 300 GT per image, batch 64 per GPU (pairs/s reported beside GB/s: that shape is pair-rate bound).

One JSON line on rank 0:
  value      images/s of the whole job.  A step = one ys_detection_loss C-ABI call (3 kernels, forward AND backward,
             gradients written) on inputs resident in HBM.  K steps are captured ONCE into a CUDA graph (host jitter
             cannot enter); the graph is replayed `--windows` times (default 25), every replay bracketed by a barrier +
             synchronize and timed with CUDA events on the launch stream; per window the MAX over ranks, over windows
             the MEDIAN (min and per-rank medians are printed too).
  e2e        the same metric through the drop-in Python API (v8DetectionLoss.__call__ + backward) with the head
             tensors and targets coming from pinned HOST memory each step and the loss items read back.
  roofline   the dominant kernel (k_loss_stream) timed in place: the step launched stage by stage with a CUDA event
             between the kernels; algorithmic bytes (2 * A * (4*reg_max+nc) * 4 per image) / CUDA-event time vs
             MEASURED_PEAKS.json.
  cpu_baseline / parity_check / c1 / nms / decode   (N=1 only) the reference's own CPU code (oracle/_ref, the unmodified
             reference package copied there by oracle/make_ref.py; kind "reference") or, without it, the oracle port (kind
             "port") on the host cores: the same B=64 step (its loss items must agree with the GPU's to 1e-5:
             parity_check), configs[0] (B=2 forward only), configs[3] (batched NMS, B=256) with its own roofline and CPU
             baseline objects, and the inference decode.
`--impl reference` prints the CPU reference arm: that same reference code timed on all host cores, rank 0 only.
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

NC, REG_MAX, B_PER_GPU = 80, 16, 64
METRIC = "YOLO11 640px assign+loss+NMS images/sec (1/2/4/8 B200) vs CPU ref; HBM GB/s"
CONFIGS = {
    "c2": {"imgsz": 640, "max_gt": 20, "fixed": None,
           "name": "configs[1]: YOLO11n assign + DFL/CIoU/BCE loss fwd+bwd, B=64 per GPU, 640px (A=8400), nc=80, "
                   "<=20 GT/img, synthetic COCO-shaped GT"},
    "c5": {"imgsz": 1280, "max_gt": 300, "fixed": 300,
           "name": "configs[4]: YOLO11m-shaped 1280px dense scene (A=33600), exactly 300 GT/img, assign + loss fwd+bwd, "
                   "B=64 per GPU"},
}


def anchors_of(imgsz):
    return sum((imgsz // s) ** 2 for s in synth.LEVEL_STRIDES)


def alg_bytes_per_img(imgsz):
    return 2 * anchors_of(imgsz) * (4 * REG_MAX + NC) * 4      # read every head value, write every gradient


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
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", 1965.0


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


def bind_host_memory_near_gpu(local):
    """Put this rank's pinned staging buffers on the NUMA node its GPU hangs off (at N > 1 the H2D copies of all ranks
    otherwise meet on whichever socket the ranks were started on).  Returns (restore, info): the CPU affinity is narrowed
    to the GPU's node (first-touch then allocates there) and, where the container allows it, the memory policy is set
    to prefer that node; restore() undoes both after the buffers exist.  Every step is best effort."""
    info = {"node": None}
    noop = (lambda: None)
    try:
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        info.update({"pci": bdf, "node": node})
        if node < 0:
            return noop, info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        before = os.sched_getaffinity(0)
        mine = cpus & before
        info["cpus_on_node"] = len(mine)
        if not mine:
            return noop, info
        os.sched_setaffinity(0, mine)
        libc, policy_set = None, False
        try:
            libc = C.CDLL(None, use_errno=True)
            mask = C.c_ulong(1 << node)
            policy_set = libc.syscall(238, 1, C.byref(mask), 65) == 0          # set_mempolicy(MPOL_PREFERRED, {node})
        except Exception:
            policy_set = False
        info["mempolicy_preferred"] = bool(policy_set)

        def restore():
            try:
                os.sched_setaffinity(0, before)
                if policy_set:
                    libc.syscall(238, 0, None, 0)                               # MPOL_DEFAULT
            except Exception:
                pass
        return restore, info
    except Exception as ex:
        info["error"] = repr(ex)
        return noop, info


def make_inputs(cfg, seed, batch=B_PER_GPU):
    feats = synth.make_head_levels(batch, cfg["imgsz"], NC, REG_MAX, seed=seed)
    if cfg["fixed"]:
        gt = synth.make_gt_batch(batch, NC, cfg["max_gt"], seed=seed + 1, imgsz=cfg["imgsz"], fixed_count=cfg["fixed"], dense=True)
    else:
        gt = synth.make_gt_batch(batch, NC, cfg["max_gt"], seed=seed + 1, imgsz=cfg["imgsz"])
    return feats, gt


def workload_config(cfg, n):
    A = anchors_of(cfg["imgsz"])
    return {"workload": cfg["name"], "batch_per_gpu": B_PER_GPU, "global_batch": B_PER_GPU * n, "anchors": A, "nc": NC,
            "reg_max": REG_MAX, "parallelism": f"image-sharded x{n}, no data-path collective",
            "l2_policy": f"inputs+gradients per step = {alg_bytes_per_img(cfg['imgsz']) * B_PER_GPU / 1e6:.0f} MB > 126 MB L2 "
                         "(no explicit flush needed)"}


# --------------------------------------------------------------------------------------- CPU reference code
class CpuReference:
    """The reference's own CPU implementation of the path: oracle/_ref (unmodified reference package; kind "reference")
    when present, else the oracle port (kind "port")."""

    def __init__(self):
        from oracle import make_ref
        torch.set_num_threads(os.cpu_count() or 1)
        self.ref_dir = make_ref.path()
        self.kind = "reference" if self.ref_dir else "port"
        self.cores = torch.get_num_threads()
        if self.ref_dir:
            sys.path.insert(0, self.ref_dir)
            import importlib
            self.loss_mod = importlib.import_module("losses.yolo_v8_loss")                       # reference, unmodified
            self.train_mod = importlib.import_module("train_yolo11_cuda")                       # reference, unmodified
            self.crit = self.loss_mod.v8DetectionLoss(Detect(NC))
        else:
            from oracle import ref_port
            self.port = ref_port

    def describe(self):
        return ("oracle/_ref (unmodified reference: losses/yolo_v8_loss.py v8DetectionLoss.__call__ + backward, "
                "train_yolo11_cuda.py decode_predictions_for_metrics)" if self.ref_dir else "oracle/ref_port.py (torch CPU port)")

    def loss_step(self, feats, batch, backward=True):
        leaves = [f.detach().requires_grad_(backward) for f in feats]
        if self.ref_dir:
            total, items = self.crit(leaves, {k: v.clone() for k, v in batch.items()})
        else:
            total, items = self.port.detection_loss(leaves, batch, NC, REG_MAX, synth.LEVEL_STRIDES)
        if backward:
            total.backward()
        return items.detach()

    def nms(self, pred, conf, iou, imgsz):
        if self.ref_dir:
            return self.train_mod.decode_predictions_for_metrics(pred, imgsz, conf, iou, torch.device("cpu"))
        return self.port.decode_for_metrics(pred, imgsz, conf, iou)

    def time_loss(self, cfg, batch_size, steps, warmup, backward=True, budget_s=None, seed=1000):
        feats, batch = make_inputs(cfg, seed, batch_size)
        items = None
        for _ in range(warmup):
            items = self.loss_step(feats, batch, backward)
        t0 = time.perf_counter()
        done = 0
        for _ in range(steps):
            items = self.loss_step(feats, batch, backward)
            done += 1
            if budget_s is not None and time.perf_counter() - t0 > budget_s:
                break
        dt = time.perf_counter() - t0
        what = "assign+loss fwd+bwd" if backward else "assign+loss forward only"
        return {"value": batch_size * done / dt, "unit": "images/s", "cores": self.cores, "kind": self.kind,
                "sample": f"{done} step(s) of B={batch_size} {what}, {cfg['imgsz']} px, after {warmup} warm-up step(s); "
                          f"{self.describe()}, torch CPU, os.cpu_count()={os.cpu_count()}",
                "ms_per_step": 1e3 * dt / done, "steps_done": done, "loss_items": [float(x) for x in items]}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    ref = CpuReference()
    # a bounded sample of the workload: the dense 1280 px scene costs seconds per image on the CPU, so its sample is B=2
    bsz = B_PER_GPU if args.config == "c2" else 2
    r = ref.time_loss(cfg, bsz, args.steps, args.warmup, backward=True, budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "images/s", "n_gpus": args.gpus,
            "steps": r["steps_done"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(cfg, args.gpus),
            "cpu_baseline": {"value": r["value"], "unit": "images/s", "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "loss_items": r["loss_items"]}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- B200 arm
class LossCall:
    """ctypes argument block of ys_detection_loss for resident tensors."""

    def __init__(self, lib, feats, grads, batch, imgsz, dev):
        self.lib = lib
        n = len(feats)
        self.n = n
        self.keep = (feats, grads, batch)
        self.B = feats[0].shape[0]
        self.A = sum(f.shape[2] * f.shape[3] for f in feats)
        self.n_gt = int(batch["cls"].shape[0])
        self.xp = (C.c_void_p * n)(*[f.data_ptr() for f in feats])
        self.gp = (C.c_void_p * n)(*[g.data_ptr() for g in grads]) if grads is not None else None
        self.hs = (C.c_int * n)(*[f.shape[2] for f in feats])
        self.ws = (C.c_int * n)(*[f.shape[3] for f in feats])
        self.st = (C.c_float * n)(*[float(s) for s in synth.LEVEL_STRIDES])
        self.bi = batch["batch_idx"].contiguous()
        self.cl = batch["cls"].contiguous().view(-1)
        self.bb = batch["bboxes"].contiguous()
        self.out = torch.zeros(4, device=dev)
        wbytes = lib.ys_detection_loss_workspace_bytes(self.B, self.A, self.n_gt, NC)
        self.wsb = torch.zeros(wbytes + 4096, dtype=torch.uint8, device=dev)
        self.imgsz = float(imgsz)
        self.dev = dev

    def __call__(self, stages=0, stream=None):
        from yolo_scratch_b200 import _lib
        s = stream if stream is not None else torch.cuda.current_stream(self.dev)
        rc = self.lib.ys_detection_loss(self.n, self.xp, self.gp, self.hs, self.ws, self.st, self.B, NC, REG_MAX,
                                        self.bi.data_ptr(), self.cl.data_ptr(), self.bb.data_ptr(), self.n_gt, self.imgsz,
                                        self.imgsz, 7.5, 0.5, 1.5, 0.5, 4.0, self.out.data_ptr(), None, None, None, stages,
                                        self.wsb.data_ptr(), self.wsb.numel(), s.cuda_stream)
        _lib.check(rc, "ys_detection_loss")


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
    cfg = CONFIGS[args.config]
    imgsz = cfg["imgsz"]
    A = anchors_of(imgsz)
    alg_step = alg_bytes_per_img(imgsz) * B_PER_GPU

    feats_h, batch_h = make_inputs(cfg, 1000 + 17 * rank)
    restore_affinity, host_numa = bind_host_memory_near_gpu(local)
    feats_h = [f.pin_memory() for f in feats_h]
    batch_h = {k: v.pin_memory() for k, v in batch_h.items()}
    restore_affinity()
    feats = [f.to(dev) for f in feats_h]
    batch = {k: v.to(dev) for k, v in batch_h.items()}
    grads = [torch.empty_like(f) for f in feats]
    call = LossCall(lib, feats, grads, batch, imgsz, dev)
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: K steps captured once into a CUDA graph, replayed `windows` times
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        call()
    torch.cuda.synchronize(dev)
    side = torch.cuda.Stream(dev)
    graph, timing_mode = None, "cuda graph of K steps, replayed per window"
    try:
        if args.no_graph:
            raise RuntimeError("--no-graph")
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            for _ in range(args.steps):
                call(stream=torch.cuda.current_stream(dev))
    except Exception as ex:      # capture is a convenience of the measurement, not of the product
        graph, timing_mode = None, f"eager launches (graph capture failed: {ex!r})"
        torch.cuda.synchronize(dev)

    def window():
        if graph is not None:
            graph.replay()
        else:
            for _ in range(args.steps):
                call()

    # clock settle: a step is ~0.2 ms, so W steps alone neither bring the SM clock up from idle nor give nvidia-smi
    # (200 ms period) a sample under load; ~0.6 s of extra untimed windows do both
    t_settle = time.perf_counter()
    while time.perf_counter() - t_settle < 0.6:
        window()
        torch.cuda.synchronize(dev)
    win_ms = []
    for _ in range(args.windows):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        window()
        e1.record(stream)
        barrier()
        win_ms.append(e0.elapsed_time(e1))
    mine = torch.tensor(win_ms, device=dev, dtype=torch.float64)
    if dist is not None:
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        allr = torch.stack(allr)
    else:
        allr = mine.unsqueeze(0)
    per_window_max = allr.max(0).values                           # max over ranks, window by window
    ms_total = float(per_window_max.median())
    value = world * B_PER_GPU * args.steps / (ms_total * 1e-3)
    loss_vec = call.out.tolist()
    timing = {"mode": timing_mode, "windows": args.windows, "steps_per_window": args.steps,
              "ms_per_step_median": ms_total / args.steps, "ms_per_step_min": float(per_window_max.min()) / args.steps,
              "ms_per_step_max": float(per_window_max.max()) / args.steps,
              "per_rank_ms_per_step_median": [float(x) / args.steps for x in allr.median(1).values],
              "per_rank_ms_per_step_min": [float(x) / args.steps for x in allr.min(1).values]}

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
    stage_ms = {name: statistics.median(evs[r][k].elapsed_time(evs[r][k + 1]) for r in range(reps))
                for k, (name, _) in enumerate(names)}
    loss_ms = stage_ms["k_loss_stream"]
    peak, peak_src, sm_max_mhz = peaks()
    achieved = alg_step / (loss_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp) and args.config == "c2":
        try:
            traffic = json.load(open(tp)).get("k_loss_stream_dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "k_loss_stream", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms": loss_ms, "algorithmic_bytes_per_launch": alg_step,
                "whole_step_achieved_gbs_per_gpu": alg_step * args.steps / (ms_total * 1e-3) / 1e9,
                "whole_step_frac": alg_step * args.steps / (ms_total * 1e-3) / 1e9 / peak,
                "stage_ms": stage_ms}
    if cfg["fixed"]:
        pairs = float(B_PER_GPU) * A * cfg["fixed"]      # (anchor, GT) pairs the reference's assigner evaluates per step
        roofline["pairs_per_step"] = pairs
        roofline["pairs_per_s_per_gpu"] = pairs * args.steps / (ms_total * 1e-3)
        roofline["note"] = "pair-rate bound shape (SURVEY §8d caveat): 10.1 M (anchor, GT) pairs per image"

    # ---- e2e: drop-in Python API, inputs from pinned host memory every step, loss items read back
    crit = v8DetectionLoss(Detect(NC).to(dev))
    e2e_steps = max(3, min(args.steps, 20 if args.config == "c2" else 5))

    # Two device-side input slots: the copy of step k+1 runs on a copy stream under the kernels of step k (every step's
    # inputs still cross PCIe inside the timed region, and every step's loss items are read back on the host).
    copy_stream = torch.cuda.Stream(dev)
    slots = [([torch.empty_like(f, device=dev) for f in feats_h], {k: torch.empty_like(v, device=dev) for k, v in batch_h.items()})
             for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def e2e_upload(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])          # the step that last read this slot has finished
            for d, h in zip(slots[slot][0], feats_h):
                d.copy_(h, non_blocking=True)
            for k, h in batch_h.items():
                slots[slot][1][k].copy_(h, non_blocking=True)
            copied[slot].record(copy_stream)

    def e2e_step(slot):
        stream.wait_event(copied[slot])
        leaves = [d.detach().requires_grad_(True) for d in slots[slot][0]]
        total, items = crit(leaves, slots[slot][1])
        total.backward()
        consumed[slot].record(stream)
        e2e_upload(slot ^ 1)                                # next step's inputs, while this step computes
        return items.cpu(), total.item()

    for sl in range(2):
        consumed[sl].record(stream)
    e2e_upload(0)
    for i in range(2):
        e2e_step(i & 1)
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    for i in range(e2e_steps):
        items_host, _ = e2e_step(i & 1)
    stream.wait_stream(copy_stream)      # the window holds K uploads: steps 1..K-1 and the one issued by the last step
    g1.record(stream)
    barrier()
    e2e_ms = g0.elapsed_time(g1)
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    clocks = sampler.stop()
    h2d = sum(f.numel() * 4 for f in feats_h) + sum(v.numel() * v.element_size() for v in batch_h.values())
    e2e = {"value": world * B_PER_GPU * e2e_steps / (e2e_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": 16, "steps": e2e_steps,
           "api": "v8DetectionLoss.__call__ + backward (drop-in Python API over the C ABI)",
           "pipeline": "step k+1's H2D copy (copy stream, second input slot) under step k's kernels; loss items read "
                       "back every step", "h2d_gbs_per_gpu": h2d * e2e_steps / (e2e_ms * 1e-3) / 1e9,
           "host_numa": host_numa}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(cfg, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": 3 * args.steps * args.windows, "timing": timing, "roofline": roofline, "loss_items": loss_vec}

    if world == 1 and not args.no_cpu:
        ref = CpuReference()
        if args.config == "c2":
            cb = ref.time_loss(cfg, B_PER_GPU, steps=12, warmup=1, budget_s=15.0)
        else:
            cb = ref.time_loss(cfg, 2, steps=3, warmup=0, budget_s=20.0)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if args.config == "c2":
            # the CPU reference ran on the very tensors the GPU path was timed on (seed 1000): its loss items must agree
            gi, ci = loss_vec[1:4], cb["loss_items"]
            rel = max(abs(a - b) / max(abs(b), 1e-12) for a, b in zip(gi, ci))
            line["parity_check"] = {"ok": bool(rel <= 1e-5), "max_rel_diff_loss_items": rel, "gpu": gi, "cpu": ci,
                                    "cpu_kind": cb["kind"], "tolerance": 1e-5}
            try:
                line["c1"] = bench_c1(dev, lib, ref)
            except Exception as ex:
                line["c1"] = {"error": repr(ex)}
    if world == 1 and not args.no_nms and args.config == "c2":
        try:
            ref = CpuReference() if not args.no_cpu else None
            line["nms"] = bench_nms(dev, lib, ref, peak, sm_max_mhz)
        except Exception as ex:  # the side measurements must not take the headline down
            line["nms"] = {"error": repr(ex)}
        try:
            line["decode"] = bench_decode(dev, feats, peak)
        except Exception as ex:
            line["decode"] = {"error": repr(ex)}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def bench_c1(dev, lib, ref):
    """configs[0]: B=2, 640 px, forward only -- the reference's own CPU-runnable case, beside the same call on the GPU."""
    cfg = CONFIGS["c2"]
    cb = ref.time_loss(cfg, 2, steps=20, warmup=2, backward=False, budget_s=5.0, seed=1000)
    feats, batch = make_inputs(cfg, 1000, 2)
    feats = [f.to(dev) for f in feats]
    batch = {k: v.to(dev) for k, v in batch.items()}
    call = LossCall(lib, feats, None, batch, cfg["imgsz"], dev)
    for _ in range(5):
        call()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 200
    e0.record()
    for _ in range(reps):
        call()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    gi = call.out.tolist()[1:4]
    rel = max(abs(a - b) / max(abs(b), 1e-12) for a, b in zip(gi, cb["loss_items"]))
    return {"workload": "configs[0]: B=2, 640 px, nc=80, <=20 GT/img, assign + loss forward only",
            "gpu_images_per_s": 2 / (ms * 1e-3), "gpu_ms": ms, "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "loss_items_rel_diff": rel, "note": "launch-latency bound on the GPU (3 kernels for 2 images)"}


def bench_nms(dev, lib, ref, peak, sm_max_mhz, B=256, reps=7):
    """configs[3]: batched NMS, 8400 candidates x 80 classes, conf 0.001 / IoU 0.7, batch 256."""
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    imgsz = 640
    A = anchors_of(imgsz)
    base = synth.make_nms_predictions(16, A, NC, imgsz, seed=2000)
    pred = base.repeat(B // 16, 1, 1).to(dev)
    pred[:, :, :2] += torch.arange(B, device=dev).view(B, 1, 1) * 0.03125      # distinct images, same statistics
    for _ in range(2):
        out = nms_batched(pred, 0.001, 0.7, imgsz)
    torch.cuda.synchronize(dev)
    times = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = nms_batched(pred, 0.001, 0.7, imgsz)
        e1.record()
        torch.cuda.synchronize(dev)
        times.append(e0.elapsed_time(e1))
    ms = statistics.median(times)
    kept = out[4].float().mean().item()
    alg = A * (4 + NC) * 4 * B
    # pair tests the reference's greedy loop performs: every kept box against the candidates still alive behind it
    # (upper estimate: kept rows x later columns)
    pairs = float((out[4].double() * (A - out[4].double() / 2)).sum().item())
    # ceiling of the pair test itself: 4 FADD + 2 LOP3 + 1 shift per pair and thread (csrc/ys_nms.cu, k_nms_suppress_zreg),
    # 128 fp32 lanes per SM
    ceiling = 148 * 128 * sm_max_mhz * 1e6 / 7.0
    res = {"workload": "configs[3]: B=256, 8400 candidates x 80 classes, conf 0.001 / IoU 0.7, class-agnostic (reference default)",
           "images_per_s": B / (ms * 1e-3), "ms": ms, "ms_min": min(times), "kept_per_image": kept,
           "roofline": {"bound": "pair tests (fp32 issue), not HBM (SURVEY §8d caveat)",
                        "algorithmic_bytes": alg, "achieved_gbs": alg / (ms * 1e-3) / 1e9, "hbm_peak_gbs": peak,
                        "frac_hbm": alg / (ms * 1e-3) / 1e9 / peak,
                        "reference_pair_tests": pairs, "pair_tests_per_s": pairs / (ms * 1e-3),
                        "pair_test_ceiling_per_s": ceiling, "frac_pairs": pairs / (ms * 1e-3) / ceiling,
                        "ceiling_note": "148 SMs x 128 lanes x max SM clock / 7 issue slots per pair test; the kernel culls "
                                        "most reference pairs spatially (Hilbert-ordered columns, per-warp bounding boxes), so "
                                        "frac_pairs counts pairs the reference evaluates, not pairs this kernel executes"}}
    if ref is not None:
        n_img = 4
        sample = base[:n_img].clone()
        t0 = time.perf_counter()
        r = ref.nms(sample, 0.001, 0.7, imgsz)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": n_img / dt, "unit": "images/s", "cores": ref.cores, "kind": ref.kind,
                               "sample": f"{n_img} images of the same distribution through {ref.describe()}; "
                                         f"{dt / n_img:.2f} s per image, {sum(len(d['scores']) for d in r) / n_img:.0f} kept per image"}
    return res


def bench_decode(dev, feats, peak, reps=20):
    """Inference decode (DFL + dist2bbox + sigmoid, Detect.inference): B=64 head of configs[1]."""
    from yolo_scratch_b200.head_decode import detect_inference
    B = feats[0].shape[0]
    A = sum(f.shape[2] * f.shape[3] for f in feats)
    alg = B * A * ((4 * REG_MAX + NC) + (4 + NC)) * 4           # 7.66 MB per image
    out = {"workload": f"Detect.inference decode, B={B}, 640 px head (A={A}), nc={NC}", "algorithmic_bytes": alg}
    for name, am in (("channels_first", False), ("anchor_major", True)):
        for _ in range(3):
            detect_inference(feats, NC, anchor_major=am)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            detect_inference(feats, NC, anchor_major=am)
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / reps
        out[name] = {"ms": ms, "images_per_s": B / (ms * 1e-3), "achieved_gbs": alg / (ms * 1e-3) / 1e9,
                     "frac_hbm": alg / (ms * 1e-3) / 1e9 / peak}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--windows", type=int, default=25, help="timed replays of the K-step window")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU legs (cpu_baseline, parity_check, c1, nms CPU)")
    ap.add_argument("--no-nms", action="store_true", help="skip the NMS and decode side measurements")
    ap.add_argument("--no-graph", action="store_true", help="eager launches in the timed windows (profiling under ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
