"""Inference-time decode of the Detect head: DFL expectation + dist2bbox(xywh) * stride + sigmoid(cls)
in one kernel.  Mirrors `DFL.forward` / `Detect.inference` of the reference
(models/yolo11_modules.py:180-192, :248-266) -- the conv backbone, neck and head convs stay in PyTorch.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


def detect_inference(feats, nc, reg_max=16, strides=(8, 16, 32), anchor_major=False):
    """feats: list of NCHW level tensors (B, 4*reg_max+nc, h, w).
    Returns y = (B, 4+nc, A) like Detect.inference()[0]; with anchor_major=True the (B, A, 4+nc) layout that
    decode_predictions_for_metrics documents."""
    xs = []
    for i, f in enumerate(feats):
        _lib.require_cuda(f, f"feats[{i}]")
        xs.append(f.detach().float().contiguous())
    n = len(xs)
    B = xs[0].shape[0]
    A = sum(x.shape[2] * x.shape[3] for x in xs)
    dev = xs[0].device
    out = torch.empty((B, A, 4 + nc) if anchor_major else (B, 4 + nc, A), device=dev, dtype=torch.float32)
    xp = (C.c_void_p * n)(*[x.data_ptr() for x in xs])
    hs = (C.c_int * n)(*[x.shape[2] for x in xs])
    ws = (C.c_int * n)(*[x.shape[3] for x in xs])
    st = (C.c_float * n)(*[float(s) for s in list(strides)[:n]])
    with _lib.on_device(*xs):
        check(_lib.lib().ys_decode_infer(n, xp, hs, ws, st, B, nc, reg_max, ptr(out), 1 if anchor_major else 0,
                                         stream_ptr(dev)), "ys_decode_infer")
    return out


class DetectInference(torch.nn.Module):
    """`y, x = DetectInference(nc, strides)(x)` == reference `Detect.inference(x)`."""

    def __init__(self, nc=80, reg_max=16, strides=(8, 16, 32)):
        super().__init__()
        self.nc, self.reg_max, self.strides = nc, reg_max, tuple(float(s) for s in strides)

    def forward(self, x):
        return detect_inference(x, self.nc, self.reg_max, self.strides), x
