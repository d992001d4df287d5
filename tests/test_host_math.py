"""CPU: the __host__ __device__ scalar math the kernels are built from (csrc/ys_math.cuh), exercised
through the library's host-side test hooks and checked against the oracle port."""
import ctypes as C

import numpy as np
import torch

from oracle import ref_port as orc
from yolo_scratch_b200 import _lib


def _hooks():
    lib = _lib.lib()
    fp = C.POINTER(C.c_float)
    lib.ys_host_iou_assign.restype = C.c_float
    lib.ys_host_iou_assign.argtypes = [fp, fp]
    lib.ys_host_iou_nms.restype = C.c_float
    lib.ys_host_iou_nms.argtypes = [fp, fp]
    lib.ys_host_centre_in_box.restype = C.c_int
    lib.ys_host_centre_in_box.argtypes = [C.c_float, C.c_float, fp]
    lib.ys_host_ciou_grad.restype = C.c_float
    lib.ys_host_ciou_grad.argtypes = [fp, fp, fp]
    lib.ys_host_bce.restype = None
    lib.ys_host_bce.argtypes = [C.c_float, C.c_float, fp, fp]
    lib.ys_host_align_norm.restype = C.c_float
    lib.ys_host_align_norm.argtypes = [C.c_float] * 5
    return lib


def _boxes(n, seed, scale=100.0):
    g = torch.Generator().manual_seed(seed)
    xy = torch.rand(n, 2, generator=g) * scale
    wh = torch.rand(n, 2, generator=g) * scale * 0.4 + 0.5
    b1 = torch.cat((xy, xy + wh), 1)
    xy2 = xy + torch.randn(n, 2, generator=g) * scale * 0.1
    wh2 = wh * (0.5 + torch.rand(n, 2, generator=g))
    return b1, torch.cat((xy2, xy2 + wh2), 1)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def test_iou_forms_bit_exact():
    lib = _hooks()
    b1, b2 = _boxes(2000, 1)
    b2[::7] = b1[::7] + 1000.0          # disjoint pairs
    b2[::11] = b1[::11]                 # identical pairs
    ref_a = orc.pair_iou_xyxy(b1, b2).squeeze(-1).clamp_(0).numpy()
    ref_n = torch.stack([orc.iou_matrix(b1[i:i + 1], b2[i:i + 1])[0, 0] for i in range(len(b1))]).numpy()
    p, q = b1.numpy().copy(), b2.numpy().copy()
    got_a = np.array([lib.ys_host_iou_assign(_fp(p[i]), _fp(q[i])) for i in range(len(p))], np.float32)
    got_n = np.array([lib.ys_host_iou_nms(_fp(p[i]), _fp(q[i])) for i in range(len(p))], np.float32)
    assert np.array_equal(got_a, ref_a)
    assert np.array_equal(got_n, ref_n)


def test_centre_in_box_matches_reference_predicate():
    lib = _hooks()
    g = torch.Generator().manual_seed(2)
    gt = torch.tensor([[8.0, 8.0, 24.0, 24.0], [3.9999, 3.9999, 4.0001, 4.0001], [0.0, 0.0, 4.0, 4.0]])
    pts = torch.cat((torch.rand(200, 2, generator=g) * 32, torch.tensor([[4.0, 4.0], [8.0, 12.0], [24.0, 12.0], [12.0, 12.0]])))
    for q in gt:
        lt = pts - q[:2]
        rb = q[2:] - pts
        ref = torch.cat((lt, rb), -1).amin(-1).gt(1e-9)
        qq = q.numpy().copy()
        got = [bool(lib.ys_host_centre_in_box(float(x), float(y), _fp(qq))) for x, y in pts.tolist()]
        assert got == ref.tolist()


def test_ciou_value_and_gradient():
    lib = _hooks()
    b1, b2 = _boxes(500, 3, scale=20.0)
    p = b1.clone().requires_grad_(True)
    ciou = orc.pair_iou_xyxy(p, b2, ciou=True).squeeze(-1)
    ciou.sum().backward()
    pn, qn = b1.numpy().copy(), b2.numpy().copy()
    g4 = np.zeros(4, np.float32)
    for i in range(len(pn)):
        v = lib.ys_host_ciou_grad(_fp(pn[i]), _fp(qn[i]), _fp(g4))
        assert abs(v - ciou[i].item()) <= 2e-6 * max(1.0, abs(ciou[i].item()))
        ref = p.grad[i].numpy()
        assert np.abs(g4 - ref).max() <= 1e-5 * max(1e-3, np.abs(ref).max()), (i, g4, ref)


def test_bce_and_sigmoid():
    lib = _hooks()
    xs = torch.linspace(-30, 30, 601)
    ts = torch.tensor([0.0, 0.3, 1.0])
    lo, sg = C.c_float(), C.c_float()
    for t in ts.tolist():
        ref = torch.nn.functional.binary_cross_entropy_with_logits(xs, torch.full_like(xs, t), reduction="none")
        for x, r in zip(xs.tolist(), ref.tolist()):
            lib.ys_host_bce(x, t, C.byref(lo), C.byref(sg))
            assert abs(lo.value - r) <= 2e-6 * max(1.0, abs(r))
            assert abs(sg.value - torch.sigmoid(torch.tensor(x)).item()) <= 2e-7


def test_align_norm():
    lib = _hooks()
    g = torch.Generator().manual_seed(4)
    s = torch.rand(200, generator=g)
    ov = torch.rand(200, generator=g)
    for a, b in ((0.5, 4.0), (1.0, 6.0), (0.7, 3.3)):
        am = s.pow(a) * ov.pow(b)
        ref = am * ov / (am + 1e-9)
        got = torch.tensor([lib.ys_host_align_norm(float(x), float(y), a, b, 1e-9) for x, y in zip(s, ov)])
        assert torch.allclose(got, ref, rtol=2e-6, atol=1e-12)
