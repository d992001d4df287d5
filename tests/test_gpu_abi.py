"""GPU: properties of the C ABI itself -- CUDA-graph capture and replay of the fused loss, a second device in the same
thread, and loss calls on batch slices whose pointers are not 16-byte aligned."""
import pytest
import torch

from yolo_scratch_b200 import synth

pytestmark = pytest.mark.gpu


class Detect(torch.nn.Module):
    def __init__(self, nc):
        super().__init__()
        self.nc, self.reg_max = nc, 16
        self.stride = torch.tensor([float(s) for s in synth.LEVEL_STRIDES])
        self.p = torch.nn.Parameter(torch.zeros(1))


def _raw_call(lib, feats, grads, batch, nc, imgsz, out, ws, stream):
    import ctypes as C
    from yolo_scratch_b200 import _lib
    n = len(feats)
    xp = (C.c_void_p * n)(*[f.data_ptr() for f in feats])
    gp = (C.c_void_p * n)(*[g.data_ptr() for g in grads])
    hs = (C.c_int * n)(*[f.shape[2] for f in feats])
    ws_ = (C.c_int * n)(*[f.shape[3] for f in feats])
    st = (C.c_float * n)(*[float(s) for s in synth.LEVEL_STRIDES])
    rc = lib.ys_detection_loss(n, xp, gp, hs, ws_, st, feats[0].shape[0], nc, 16, batch["batch_idx"].data_ptr(),
                               batch["cls"].data_ptr(), batch["bboxes"].data_ptr(), int(batch["cls"].shape[0]),
                               float(imgsz), float(imgsz), 7.5, 0.5, 1.5, 0.5, 4.0, out.data_ptr(), None, None, None, 0,
                               ws.data_ptr(), ws.numel(), stream.cuda_stream)
    _lib.check(rc, "ys_detection_loss")


@pytest.mark.parametrize("imgsz,B,max_gt,kw", [(128, 3, 6, {"p_tiny": 0.3, "p_nested": 0.3}), (640, 8, 20, {}),
                                                (672, 2, 12, {}),
                                                # more than 1024 GT: the memset node + multi-block k_pack_gt in the graph
                                                (320, 8, 300, {"fixed_count": 300, "dense": True})])
def test_detection_loss_cuda_graph_capture_and_replay(cuda_lib, imgsz, B, max_gt, kw):
    """ys_detection_loss allocates nothing and never synchronises: the whole call (pack, assign, loss; programmatic
    dependent launches included) is captured into a CUDA graph and replayed; every replay gives bit-identical loss
    items and gradients to the eager call, also after the inputs change in place."""
    from yolo_scratch_b200 import _lib
    lib = cuda_lib
    nc = 8 if imgsz == 128 else 80
    feats = [f.cuda() for f in synth.make_head_levels(B, imgsz, nc, 16, seed=500 + imgsz)]
    batch = {k: v.cuda() for k, v in synth.make_gt_batch(B, nc, max_gt, seed=501 + imgsz, imgsz=imgsz, p_empty=0.0, **kw).items()}
    batch["cls"] = batch["cls"].view(-1).contiguous()
    grads = [torch.zeros_like(f) for f in feats]
    A = sum(f.shape[2] * f.shape[3] for f in feats)
    out = torch.zeros(4, device="cuda")
    ws = torch.zeros(lib.ys_detection_loss_workspace_bytes(B, A, int(batch["cls"].shape[0]), nc) + 256, dtype=torch.uint8,
                     device="cuda")
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        _raw_call(lib, feats, grads, batch, nc, imgsz, out, ws, side)           # eager (also raises the smem limits)
    side.synchronize()
    ref_out, ref_g = out.clone(), [g.clone() for g in grads]
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        _raw_call(lib, feats, grads, batch, nc, imgsz, out, ws, torch.cuda.current_stream())
    for _ in range(3):
        out.zero_()
        for g in grads:
            g.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, ref_out)
        for g, r in zip(grads, ref_g):
            assert torch.equal(g, r)
    # new values behind the same pointers: the replay must follow them
    for f in feats:
        f.mul_(0.5)
    graph.replay()
    torch.cuda.synchronize()
    replay_out, replay_g = out.clone(), [g.clone() for g in grads]
    _raw_call(lib, feats, grads, batch, nc, imgsz, out, ws, torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert not torch.equal(replay_out, ref_out)
    assert torch.equal(out, replay_out)
    for g, r in zip(grads, replay_g):
        assert torch.equal(g, r)


def test_loss_on_unaligned_batch_slices(cuda_lib):
    """feats[i][1:] of a head with odd h*w is contiguous but starts 4 bytes off a 16-byte boundary: those levels take
    the plain-load loaders; the result equals the call on fresh (aligned) copies bit for bit."""
    from yolo_scratch_b200.losses import v8DetectionLoss
    shapes = [(18, 18), (9, 9), (5, 5)]
    nc = 7
    feats = synth.make_head_levels(3, 0, nc, 16, seed=520, shapes=shapes)
    batch = synth.make_gt_batch(2, nc, 6, seed=521, imgsz=144, p_empty=0.0)
    crit = v8DetectionLoss(Detect(nc).cuda())
    big = [f.cuda() for f in feats]
    views = [f[1:].detach().requires_grad_(True) for f in big]
    assert any(v.data_ptr() % 16 for v in views) and all(v.is_contiguous() for v in views)
    t1, i1 = crit(views, {k: v.cuda() for k, v in batch.items()})
    t1.backward()
    fresh = [f[1:].clone().cuda().requires_grad_(True) for f in feats]
    t2, i2 = crit(fresh, {k: v.cuda() for k, v in batch.items()})
    t2.backward()
    assert torch.equal(i1, i2) and torch.equal(t1, t2)
    for a, b in zip(views, fresh):
        assert torch.allclose(a.grad, b.grad, rtol=0, atol=0)


def test_second_device_same_thread(cuda_lib):
    """Per-device caches (shared-memory opt-in, SM count): a second GPU used from the same thread must work."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from yolo_scratch_b200.losses import v8DetectionLoss
    from yolo_scratch_b200.train_yolo11_cuda import nms_batched
    feats = synth.make_head_levels(2, 128, 8, 16, seed=530)
    batch = synth.make_gt_batch(2, 8, 5, seed=531, imgsz=128, p_empty=0.0)
    res = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        crit = v8DetectionLoss(Detect(8).to(dev))
        leaves = [f.to(dev).requires_grad_(True) for f in feats]
        total, items = crit(leaves, {k: v.to(dev) for k, v in batch.items()})      # current device stays cuda:0
        total.backward()
        pred = synth.make_nms_predictions(2, 700, 8, 640, seed=532).to(dev)
        ka, _, _, _, kc = nms_batched(pred, 0.001, 0.7, 640)
        res.append((items.cpu(), [l.grad.cpu() for l in leaves], ka.cpu(), kc.cpu()))
    assert torch.equal(res[0][0], res[1][0])
    for a, b in zip(res[0][1], res[1][1]):
        assert torch.equal(a, b)
    assert torch.equal(res[0][2], res[1][2]) and torch.equal(res[0][3], res[1][3])
